// Epilogue of the tcgen05 contraction kernels (gemm_tc.cu, gemm_tc2.cu): one 128-row accumulator tile, TMEM ->
// registers -> [row sums / bias / activation / residual / scale / rounding] -> swizzled staging -> whole-line stores.
#pragma once
#include "tc_common.cuh"

namespace dino {

struct TcEpilogue {
  int M, N, K;
  float* C; long long ldc;
  int sdim;
  const float* scale; long long ld_sc;
  const float* add_pre; long long ld_ap;
  const float* add_post; long long ld_apo;
  const float* bias;
  int act; float* s_out; long long ld_s;
  const float* yref; long long ld_y; float* rowsum; long long ld_rs;
  int round_tf32;
  // gemm_tc.cu only: A gathered per sample (a_rows rows each, a_tps row tiles per sample, sample numbers a_idx or 0..),
  // per-row dot products with yref, add_pre ahead of the row sums
  int a_rows, a_tps; const int* a_idx;
  float* rowdot; int pre_first;
  int b_lo;                  // fp32x3: the lo tile of B arrives by TMA (precomputed), the converters split A only
  int a_gran;                // > 0: consecutive-row tiles loaded as boxes of multiples of a_gran rows (gemm_tc.cu: TcGatherMaps)
};

// 4 floats of a row at column n (n % 4 == 0): vector load when the four lie inside N, guarded scalars at the edge
__device__ __forceinline__ float4 tc_ld4(const float* row, int n, int N) {
  if (n + 3 < N) return *reinterpret_cast<const float4*>(row + n);
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  if (n < N) v.x = row[n];
  if (n + 1 < N) v.y = row[n + 1];
  if (n + 2 < N) v.z = row[n + 2];
  return v;
}

// One accumulator tile of 128 rows x BN columns at tensor-memory address `tacc` (lane field 0), rows m0.. of C, columns
// n0..; executed by the epilogue warp (q = its TMEM lane quarter, half = which of the NEPI / 4 warps of that quarter).
// EPI: 0 plain store; 1 chain terms (rowsum / add_pre / scale / add_post / rounding); 2 network terms (bias / softplus +
// sigmoid / add_pre / add_post); 3 chain terms plus the per-row dot products and add_pre ahead of the row sums (gemm_tc.cu
// only: the first layer without its Jacobian).  SPLIT: the cross-term accumulator sits ACC_COLS columns above the main one.
template <int BN, bool SPLIT, int EPI_, int NEPI, uint32_t ACC_COLS>
__device__ __forceinline__ void tc_epilogue_tile(const TcEpilogue& ep, int m0, int n0, uint32_t tacc_base, uint32_t my_stage,
                                                 int q, int half, int lane, int m_lim) {
  using namespace tc;
  constexpr int EPI = EPI_ == 3 ? 1 : EPI_;          // class 3 = class 1 + the terms marked EXT
  constexpr bool EXT = EPI_ == 3;
  const uint32_t sw = (uint32_t)(lane & 7);
  {
      const int row = m0 + q * 32 + lane;
      const bool row_ok = row < m_lim;     // m_lim: first row past this tile's share of C (M, or the end of the sample)
      const int b = row_ok ? row / ep.sdim : 0;
      const float* sc = (EPI == 1 && ep.scale) ? ep.scale + (long long)b * ep.ld_sc : nullptr;
      const float* ap = (EPI != 0 && ep.add_pre) ? ep.add_pre + (long long)row * ep.ld_ap : nullptr;
      const float* apo = (EPI != 0 && ep.add_post) ? ep.add_post + (long long)row * ep.ld_apo : nullptr;
      const float* yr = (EPI == 1 && ep.yref) ? ep.yref + (long long)row * ep.ld_y : nullptr;
      float* so = (EPI == 2 && ep.s_out) ? ep.s_out + (long long)row * ep.ld_s : nullptr;
      const uint32_t tacc = tacc_base + ((uint32_t)(q * 32) << 16);
      const bool pre_first = EXT && ep.pre_first && ap;
      float rdot = 0.f;
#pragma unroll 1
      for (int c0 = half * 32; c0 < BN; c0 += 32 * (NEPI / 4)) {
        if (n0 + c0 >= ep.N) break;                  // warp-uniform
        const int nb = n0 + c0;
        const bool full = nb + 32 <= ep.N;           // warp-uniform: the block lies inside N
        float v[32];
        tmem_ld32(tacc + (uint32_t)c0, v);
        if (SPLIT) {
          float x[32];
          tmem_ld32(tacc + ACC_COLS + (uint32_t)c0, x);
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] += x[j];
        }
        // 32 floats of a row-indexed operand at this block's columns (zeros outside the matrix)
        auto load32 = [&](float (&d)[32], const float* rowp, bool ok) {
          if (full) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              const float4 t = ok ? *reinterpret_cast<const float4*>(rowp + nb + j) : make_float4(0.f, 0.f, 0.f, 0.f);
              d[j] = t.x; d[j + 1] = t.y; d[j + 2] = t.z; d[j + 3] = t.w;
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              const float4 t = ok ? tc_ld4(rowp, nb + j, ep.N) : make_float4(0.f, 0.f, 0.f, 0.f);
              d[j] = t.x; d[j + 1] = t.y; d[j + 2] = t.z; d[j + 3] = t.w;
            }
          }
        };
        if (pre_first) {
          float t[32];
          load32(t, ap, row_ok);
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] += t[j];
        }
        if (EPI == 1 && yr) {
          // d sigma term: rowsum[sample, n] += sum over the sample's chain rows of acc * yref  (segmented over the
          // samples under this warp's 32 rows; lane j ends up with column j)
          float p[32];
          load32(p, yr, row_ok);
#pragma unroll
          for (int j = 0; j < 32; ++j) p[j] *= v[j];
          if (EXT && ep.rowdot) {
#pragma unroll
            for (int j = 0; j < 32; ++j) rdot += p[j];
          }
          const int r_first = m0 + q * 32;
          if ((!EXT || ep.rowsum) && r_first < m_lim) {
            const int b_lo = r_first / ep.sdim, b_hi = min(r_first + 31, m_lim - 1) / ep.sdim;
            for (int bb = b_lo; bb <= b_hi; ++bb) {
              float mm[32];
#pragma unroll
              for (int j = 0; j < 32; ++j) mm[j] = (row_ok && b == bb) ? p[j] : 0.f;
              const float tot = colsum32(mm, lane);
              if (nb + lane < ep.N) atomicAdd(ep.rowsum + (long long)bb * ep.ld_rs + nb + lane, tot);
            }
          }
          if (!ep.C) continue;
        }
        if (EPI == 2) {
          if (ep.bias) {
            float t[32];
            load32(t, ep.bias, true);
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] += t[j];
          }
          if (ep.act) {
            // sigmoid and softplus from one exponential: e = exp(-|x|)
            float sg[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const float e = expf(-fabsf(v[j])), r1 = 1.f / (1.f + e);
              sg[j] = v[j] >= 0.f ? r1 : e * r1;
              v[j] = fmaxf(v[j], 0.f) + log1pf(e);
            }
            if (so && row_ok) {
              if (full) {
#pragma unroll
                for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(so + nb + j) = make_float4(sg[j], sg[j + 1], sg[j + 2], sg[j + 3]);
              } else {
#pragma unroll
                for (int j = 0; j < 32; ++j) if (nb + j < ep.N) so[nb + j] = sg[j];
              }
            }
          }
        }
        if (EPI != 0) {
          if (ap && !pre_first) {
            float t[32];
            load32(t, ap, row_ok);
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] += t[j];
          }
          if (sc) {
            float t[32];
            load32(t, sc, row_ok);
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] *= t[j];
          }
          if (apo) {
            float t[32];
            load32(t, apo, row_ok);
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] += t[j];
          }
          if (ep.round_tf32) {
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] = round_tf32_f(v[j]);
          }
        }
        // stage the block (thread = row, 16-byte chunks XOR-swizzled by the row), then store it four rows per
        // instruction: 8 lanes x 16 bytes = one whole 128-byte line of C per row
        __syncwarp();                                // the previous block's reads of the staging buffer are done
        const uint32_t rowb = my_stage + (uint32_t)(lane * 128);
#pragma unroll
        for (int j = 0; j < 8; ++j) sts128(rowb + ((((uint32_t)j) ^ sw) << 4), v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
        __syncwarp();
        const int ch = lane & 7, n = nb + 4 * ch;
        float* cbase = ep.C + (long long)(m0 + q * 32 + (lane >> 3)) * ep.ldc + n;
        const int rows_left = m_lim - (m0 + q * 32 + (lane >> 3));      // this lane's rows: rows_left > 4 * it
        if (full) {
#pragma unroll
          for (int it = 0; it < 8; ++it) {
            const int rr = it * 4 + (lane >> 3);
            const float4 o = lds128(my_stage + (uint32_t)(rr * 128) + ((((uint32_t)ch) ^ (uint32_t)(rr & 7)) << 4));
            if (rows_left > 4 * it) *reinterpret_cast<float4*>(cbase + (long long)(4 * it) * ep.ldc) = o;
          }
        } else {
#pragma unroll
          for (int it = 0; it < 8; ++it) {
            const int rr = it * 4 + (lane >> 3);
            const float4 o = lds128(my_stage + (uint32_t)(rr * 128) + ((((uint32_t)ch) ^ (uint32_t)(rr & 7)) << 4));
            if (rows_left > 4 * it && n < ep.N) {
              float* dst = cbase + (long long)(4 * it) * ep.ldc;
              if (n + 3 < ep.N) *reinterpret_cast<float4*>(dst) = o;
              else { dst[0] = o.x; if (n + 1 < ep.N) dst[1] = o.y; if (n + 2 < ep.N) dst[2] = o.z; }
            }
          }
        }
      }
      if (EXT && ep.rowdot && row_ok) atomicAdd(ep.rowdot + row, rdot);
  }
}

}  // namespace dino
