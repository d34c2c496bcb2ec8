// Weight gradients on the tensor cores:
//     dW_l[n_in, n_out] += dR_l^T Q_l = X[rows, n_in]^T * Y[rows, n_out]          (rows = B*s, tens of thousands)
// for every layer in ONE launch, split over the rows.  X / Y are the chain states exactly as the chain kernels keep
// them: ROW-major.  The reduction index is then the strided one, i.e. both factors are MN-major UMMA operands; for
// 4-byte elements that takes the 128-byte swizzle with 32-byte atoms on the TMA side (CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B)
// and shared-memory descriptor layout type 1 (SWIZZLE_128B_BASE32B: LBO = pitch between 32-column groups, SBO = 512 B
// = 4 reduction rows) plus the two "major" bits of the instruction descriptor.  Transposed factors [n, rows] (K-major,
// plain SWIZZLE_128B) are still accepted by the 128-wide kernels.
//
// Pipeline: TMA ring -> [fp32x3: four converter warps write the lo tile x - trunc_tf32(x) of every landed stage beside
// it] -> tcgen05.mma.kind::tf32 (one MMA per 8 reduction rows, or three on the hi/lo splits) -> TMEM -> red.global.add.
// Output tiles of 128 x TN (TN = 256 for wide layers: a 128 x 128 tile needs 128 bytes of operands per clock from
// shared memory, which is all there is; 128 x 256 needs 96); blockIdx.x enumerates op, tile, split.
// fp32x3: the cross terms have a TMEM accumulator of their own and one accumulator takes at most WG_X3_ROWS reduction
// rows (the tensor core's accumulation truncates: profiles/r02a_precision.md); partial sums meet in fp32 atomics.
#include <cuda.h>

#include <cstdlib>

#include "tc_common.cuh"

namespace dino {

constexpr int WG_NCONV = 8;                    // fp32x3: converter warps
constexpr int WG_X3_ROWS = 2048;                // fp32x3: longest reduction one TMEM accumulator takes (256 truncating additions, two chain contractions' worth; 1024 cost 20 % more time in per-CTA set-up and drain)

struct WgOp {
  float* C;                 // gradient block [n_in, n_out]
  long long ldc;            // its row pitch (n_out for a block of the parameter arena)
  int n_in, n_out, rows;    // rows = reduction length
  int z_start;              // first blockIdx.x of this op
  int chunks_per_cta;       // reduction chunks (of BK rows) per CTA
  int n_tiles, splits;      // CTAs of this op: m_tiles * n_tiles * splits, split fastest
  int mn;                   // 1: operands are the ROW-major chain states [rows, n] (MN-major UMMA operands), 0: transposed copies
  // X gathered per sample from a dataset [samples, x_rows, n_in] (3-D tensor map): the reduction runs over chunks that
  // do not cross samples, x_cps of them per sample (the last one zero-filled past the sample's rows)
  const int* x_idx; int x_rows, x_cps;
  float alpha;              // C += alpha * X^T Y
};
struct WgGroup {
  int n;
  WgOp op[GEMM_GROUP_MAX];
  CUtensorMap mx[GEMM_GROUP_MAX];   // X [rows, n_in]  (or XT [n_in, rows])
  CUtensorMap my[GEMM_GROUP_MAX];   // Y [rows, n_out] (or YT [n_out, rows])
};

template <int TN, int BK, int STAGES, bool SPLIT>
__global__ void __launch_bounds__(SPLIT ? 192 + WG_NCONV * 32 : 192, 1) wgrad_tc_kernel(const __grid_constant__ WgGroup g) {
  using namespace tc;
  constexpr int GROUP_BYTES = BK * 128;                    // one 32-column group: BK reduction rows of 128 bytes
  constexpr int A_BYTES = 4 * GROUP_BYTES, B_BYTES = (TN / 32) * GROUP_BYTES, RAW_BYTES = A_BYTES + B_BYTES;
  constexpr int STAGE_BYTES = SPLIT ? 2 * RAW_BYTES : RAW_BYTES;
  constexpr uint32_t TMEM_COLS = SPLIT ? 2 * TN : TN;      // fp32x3: main sum in columns [0, TN), cross terms in [TN, 2 TN)
  extern __shared__ uint8_t wg_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(wg_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* conv_bar = empty_bar + STAGES;                 // lo tiles written (fp32x3)
  uint64_t* done_bar = conv_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done_bar + 1);

  int oi = 0;
  while (oi + 1 < g.n && (int)blockIdx.x >= g.op[oi + 1].z_start) ++oi;
  const WgOp& op = g.op[oi];
  const int total_chunks = op.x_rows > 0 ? (op.rows / op.x_rows) * op.x_cps : (op.rows + BK - 1) / BK;
  const int local = blockIdx.x - op.z_start;
  const int split = local % op.splits, tile = local / op.splits;
  const int m0 = (tile / op.n_tiles) * 128, n0 = (tile % op.n_tiles) * TN;
  const int c_begin = split * op.chunks_per_cta;
  const int nchunks = min(total_chunks - c_begin, op.chunks_per_cta);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ncols = min(TN, op.n_out - n0);

  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < STAGES; ++i) { mbar_init(full_bar + i, 1); mbar_init(empty_bar + i, 1); mbar_init(conv_bar + i, WG_NCONV); }
      mbar_init(done_bar, 1);
      fence_mbar_init();
    }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (nchunks > 0) {
    if (warp == 0) {
      if (lane == 0) {
        const CUtensorMap* mx = &g.mx[oi];
        const CUtensorMap* my = &g.my[oi];
        // row-major states: one box of 32 columns x BK rows per column group that holds data
        const int ga = (min(128, op.n_in - m0) + 31) >> 5, gb = (ncols + 31) >> 5;
        for (int c = 0; c < nchunks; ++c) {
          const int st = c % STAGES;
          mbar_wait(empty_bar + st, ((c / STAGES) & 1) ^ 1);
          uint8_t* sx = smem + st * STAGE_BYTES;
          const int k0 = (c_begin + c) * BK;
          if (op.x_rows > 0) {
            const int b = (c_begin + c) / op.x_cps, r0 = ((c_begin + c) % op.x_cps) * BK;
            const int samp = op.x_idx ? op.x_idx[b] : b;
            mbar_expect_tx(full_bar + st, (uint32_t)((ga + gb) * GROUP_BYTES));
            for (int gI = 0; gI < ga; ++gI) tma_load_3d(sx + gI * GROUP_BYTES, mx, m0 + 32 * gI, r0, samp, full_bar + st);
            // (Y rows past the sample's last meet zero-filled X rows)
            for (int gI = 0; gI < gb; ++gI) tma_load_2d(sx + A_BYTES + gI * GROUP_BYTES, my, n0 + 32 * gI, b * op.x_rows + r0, full_bar + st);
          } else if (op.mn) {
            mbar_expect_tx(full_bar + st, (uint32_t)((ga + gb) * GROUP_BYTES));
            for (int gI = 0; gI < ga; ++gI) tma_load_2d(sx + gI * GROUP_BYTES, mx, m0 + 32 * gI, k0, full_bar + st);
            for (int gI = 0; gI < gb; ++gI) tma_load_2d(sx + A_BYTES + gI * GROUP_BYTES, my, n0 + 32 * gI, k0, full_bar + st);
          } else {
            // transposed copies (TN == 128, BK == 32): [128 operand rows][32 reduction elements], SWIZZLE_128B
            mbar_expect_tx(full_bar + st, (uint32_t)RAW_BYTES);
            tma_load_2d(sx, mx, k0, m0, full_bar + st);
            tma_load_2d(sx + A_BYTES, my, k0, n0, full_bar + st);
          }
        }
      }
    } else if (warp == 1) {
      if (lane == 0) {
        const uint32_t n16 = (uint32_t)((ncols + 15) & ~15);
        const uint32_t idesc = idesc_tf32(n16, op.mn != 0, op.mn != 0);
        const uint64_t kstep = op.mn ? 64 : 2;      // descriptor start-address advance per 8 reduction rows (16-byte units)
        for (int c = 0; c < nchunks; ++c) {
          const int st = c % STAGES;
          mbar_wait((SPLIT ? conv_bar : full_bar) + st, (c / STAGES) & 1);
          tc_fence_after();
          const uint32_t sx = smem_u32(smem + st * STAGE_BYTES), sy = sx + A_BYTES;
          const uint64_t ad = op.mn ? desc_mn_sw128_32(sx, GROUP_BYTES) : desc_k_sw128(sx);
          const uint64_t bd = op.mn ? desc_mn_sw128_32(sy, GROUP_BYTES) : desc_k_sw128(sy);
          constexpr uint64_t LO = RAW_BYTES >> 4;   // the lo tiles sit RAW_BYTES above their sources
#pragma unroll
          for (int k = 0; k < BK / 8; ++k) {
            umma_tf32(tmem_base, ad + kstep * k, bd + kstep * k, idesc, (c > 0 || k > 0) ? 1u : 0u);
            if (SPLIT) {
              umma_tf32(tmem_base + TN, ad + LO + kstep * k, bd + kstep * k, idesc, (c > 0 || k > 0) ? 1u : 0u);
              umma_tf32(tmem_base + TN, ad + kstep * k, bd + LO + kstep * k, idesc, 1u);
            }
          }
          umma_commit(empty_bar + st);
        }
        umma_commit(done_bar);
      }
    } else if (SPLIT && warp >= 6) {
      // converters (fp32x3): lo tile of every landed stage (whole stage: the layout does not matter elementwise)
      for (int c = 0; c < nchunks; ++c) {
        const int st = c % STAGES;
        mbar_wait(full_bar + st, (c / STAGES) & 1);
        const uint32_t sx = smem_u32(smem + st * STAGE_BYTES);
        split_lo_tile(sx, sx + RAW_BYTES, RAW_BYTES, (int)threadIdx.x - 192, WG_NCONV * 32);
        fence_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive(conv_bar + st);
      }
    } else {
      // epilogue: TMEM lane = n_in index m, columns = n_out index; the partial sum joins C through fp32 atomics
      mbar_wait(done_bar, 0);
      tc_fence_after();
      const int q = warp & 3;
      const int m = m0 + q * 32 + lane;
      const bool vec = (op.ldc & 3) == 0 && (reinterpret_cast<uintptr_t>(op.C) & 15) == 0;
      for (int c0 = 0; c0 < ncols; c0 += 32) {
        float v[32];
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c0;
        tmem_ld32(taddr, v);
        if (SPLIT) {
          float x[32];
          tmem_ld32(taddr + TN, x);
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] += x[j];
        }
        if (op.alpha != 1.f) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] *= op.alpha;
        }
        if (m < op.n_in) {
          float* crow = op.C + (long long)m * op.ldc + n0 + c0;
          if (vec && c0 + 32 <= ncols) {
#pragma unroll
            for (int j = 0; j < 32; j += 4)
              asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(crow + j), "f"(v[j]), "f"(v[j + 1]), "f"(v[j + 2]), "f"(v[j + 3]) : "memory");
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (c0 + j < ncols) atomicAdd(crow + j, v[j]);
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
}

// ---- the same reduction on CTA pairs (tcgen05 cta_group::2): a pair owns a 256 x 256 tile of dW; each CTA stages its
// 128 columns of X and its 128 columns of Y per reduction chunk, the leader issues the M = 256 MMAs, each CTA's tensor
// memory holds its 128 rows of the tile.  Barrier scheme as in gemm_tc2.cu.  For layers with n_in, n_out > 128.
template <int BK, int STAGES, bool SPLIT>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(SPLIT ? 192 + WG_NCONV * 32 : 192, 1)
wgrad_tc2_kernel(const __grid_constant__ WgGroup g) {
  using namespace tc;
  constexpr int GROUP_BYTES = BK * 128;
  constexpr int A_BYTES = 4 * GROUP_BYTES, RAW_BYTES = 8 * GROUP_BYTES;
  constexpr int STAGE_BYTES = SPLIT ? 2 * RAW_BYTES : RAW_BYTES;
  constexpr uint32_t TMEM_COLS = SPLIT ? 512 : 256;
  extern __shared__ uint8_t wg_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(wg_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* conv_bar = empty_bar + STAGES;
  uint64_t* done_bar = conv_bar + STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done_bar + 1);

  const int pairx = blockIdx.x >> 1;
  const uint32_t rank = cluster_ctarank();
  int oi = 0;
  while (oi + 1 < g.n && pairx >= g.op[oi + 1].z_start) ++oi;
  const WgOp& op = g.op[oi];
  const int total_chunks = (op.rows + BK - 1) / BK;
  const int local = pairx - op.z_start;
  const int split = local % op.splits, tile = local / op.splits;
  const int m0 = (tile / op.n_tiles) * 256, n0 = (tile % op.n_tiles) * 256;
  const int c_begin = split * op.chunks_per_cta;
  const int nchunks = min(total_chunks - c_begin, op.chunks_per_cta);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ncols = min(256, op.n_out - n0);
  // 32-column groups that hold data, per CTA of the pair: X columns [m0 + 128 r, +128), Y columns [n0 + 128 r, +128)
  auto groups = [](int have) { return have <= 0 ? 0 : (min(128, have) + 31) >> 5; };
  const int ga0 = groups(op.n_in - m0), ga1 = groups(op.n_in - m0 - 128), gb0 = groups(op.n_out - n0), gb1 = groups(op.n_out - n0 - 128);
  const int ga = rank ? ga1 : ga0, gb = rank ? gb1 : gb0;

  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < STAGES; ++i) { mbar_init(full_bar + i, 1); mbar_init(empty_bar + i, 1); mbar_init(conv_bar + i, 2 * WG_NCONV); }
      mbar_init(done_bar, 1);
      fence_mbar_init();
    }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (nchunks > 0) {
    if (warp == 0) {
      if (lane == 0) {
        const CUtensorMap* mx = &g.mx[oi];
        const int mc = m0 + 128 * (int)rank;
        for (int c = 0; c < nchunks; ++c) {
          const int st = c % STAGES;
          mbar_wait(empty_bar + st, ((c / STAGES) & 1) ^ 1);
          uint8_t* sx = smem + st * STAGE_BYTES;
          const int k0 = (c_begin + c) * BK;
          // (this thread brings the X boxes; the Y boxes are issued by a lane of warp 2, idle until the epilogue: the
          //  issue rate of the 32-column boxes, not bandwidth, is what the feed runs out of)
          if (SPLIT) {
            mbar_expect_tx(full_bar + st, (uint32_t)((ga + gb) * GROUP_BYTES));
            for (int gI = 0; gI < ga; ++gI) tma_load_2d(sx + gI * GROUP_BYTES, mx, mc + 32 * gI, k0, full_bar + st);
          } else {
            if (rank == 0) mbar_expect_tx(full_bar + st, (uint32_t)((ga0 + ga1 + gb0 + gb1) * GROUP_BYTES));
            for (int gI = 0; gI < ga; ++gI) tma_load_2d_2sm(sx + gI * GROUP_BYTES, mx, mc + 32 * gI, k0, full_bar + st);
          }
        }
      }
    } else if (warp == 1) {
      if (rank == 0 && lane == 0) {
        const uint32_t idesc = idesc_tf32_m(256, 256) | (1u << 15) | (1u << 16);     // both operands MN-major
        for (int c = 0; c < nchunks; ++c) {
          const int st = c % STAGES;
          mbar_wait((SPLIT ? conv_bar : full_bar) + st, (c / STAGES) & 1);
          tc_fence_after();
          const uint32_t sx = smem_u32(smem + st * STAGE_BYTES), sy = sx + A_BYTES;
          const uint64_t ad = desc_mn_sw128_32(sx, GROUP_BYTES), bd = desc_mn_sw128_32(sy, GROUP_BYTES);
          constexpr uint64_t LO = RAW_BYTES >> 4;
#pragma unroll
          for (int k = 0; k < BK / 8; ++k) {
            umma_tf32_2sm(tmem_base, ad + 64 * k, bd + 64 * k, idesc, (c > 0 || k > 0) ? 1u : 0u);
            if (SPLIT) {
              umma_tf32_2sm(tmem_base + 256u, ad + LO + 64 * k, bd + 64 * k, idesc, (c > 0 || k > 0) ? 1u : 0u);
              umma_tf32_2sm(tmem_base + 256u, ad + 64 * k, bd + LO + 64 * k, idesc, 1u);
            }
          }
          umma_commit_2sm(empty_bar + st);
        }
        umma_commit_2sm(done_bar);
      }
    } else if (SPLIT && warp >= 6) {
      for (int c = 0; c < nchunks; ++c) {
        const int st = c % STAGES;
        // (a CTA whose halves are empty has nothing in flight: its full barrier never fires, it only reports in)
        if (ga + gb > 0) {
          mbar_wait(full_bar + st, (c / STAGES) & 1);
          const uint32_t sx = smem_u32(smem + st * STAGE_BYTES);
          split_lo_tile(sx, sx + RAW_BYTES, RAW_BYTES, (int)threadIdx.x - 192, WG_NCONV * 32);
          fence_async_smem();
        } else {
          mbar_wait(empty_bar + st, ((c / STAGES) & 1) ^ 1);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(conv_bar + st, 0);
      }
    } else {
      if (warp == 2 && lane == 0) {
        // second producer: the Y boxes of every stage
        const CUtensorMap* my = &g.my[oi];
        const int nc = n0 + 128 * (int)rank;
        for (int c = 0; c < nchunks; ++c) {
          const int st = c % STAGES;
          mbar_wait(empty_bar + st, ((c / STAGES) & 1) ^ 1);
          uint8_t* sx = smem + st * STAGE_BYTES;
          const int k0 = (c_begin + c) * BK;
          for (int gI = 0; gI < gb; ++gI) {
            if (SPLIT) tma_load_2d(sx + A_BYTES + gI * GROUP_BYTES, my, nc + 32 * gI, k0, full_bar + st);
            else tma_load_2d_2sm(sx + A_BYTES + gI * GROUP_BYTES, my, nc + 32 * gI, k0, full_bar + st);
          }
        }
      }
      __syncwarp();
      mbar_wait(done_bar, 0);
      tc_fence_after();
      const int q = warp & 3;
      const int m = m0 + 128 * (int)rank + q * 32 + lane;
      const bool vec = (op.n_out & 3) == 0 && (reinterpret_cast<uintptr_t>(op.C) & 15) == 0;
      for (int c0 = 0; c0 < ncols; c0 += 32) {
        float v[32];
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c0;
        tmem_ld32(taddr, v);
        if (SPLIT) {
          float x[32];
          tmem_ld32(taddr + 256u, x);
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] += x[j];
        }
        if (m < op.n_in) {
          float* crow = op.C + (long long)m * op.n_out + n0 + c0;
          if (vec && c0 + 32 <= ncols) {
#pragma unroll
            for (int j = 0; j < 32; j += 4)
              asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(crow + j), "f"(v[j]), "f"(v[j + 1]), "f"(v[j + 2]), "f"(v[j + 3]) : "memory");
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (c0 + j < ncols) atomicAdd(crow + j, v[j]);
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync();
  if (warp == 1)
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
}

// ---- host ---------------------------------------------------------------------------------------------
// The ops passed here describe C[M,N] += A[M,K] * Bt[N,K]^T with BOTH operands K-major (a_sk == b_sk == 1):
// A = XT (pitch a_sm), B = YT (pitch b_sn), or both MN-major (a_sm == b_sn == 1: the row-major chain states).
bool wgrad_tc_supported(const GemmOp& op) {
  if (!op.atomic_c || op.b_kmod > 0 || op.a_mod > 0) return false;
  if (op.M < 1 || op.N < 1 || op.K < 1) return false;
  const bool kmaj = op.a_sk == 1 && op.b_sk == 1;                      // transposed copies [n, rows]
  const bool mnmaj = op.a_sm == 1 && op.b_sn == 1;                     // row-major states [rows, n], pitches a_sk / b_sk
  if (!kmaj && !mnmaj) return false;
  if (kmaj && ((op.a_sm & 3) || (op.b_sn & 3))) return false;
  if (!kmaj && ((op.a_sk & 3) || (op.b_sk & 3))) return false;
  if ((reinterpret_cast<uintptr_t>(op.A) & 15) || (reinterpret_cast<uintptr_t>(op.B) & 15)) return false;
  if (op.a_rows > 0 && (!mnmaj || op.K % op.a_rows || (op.a_ld2 & 3) || op.a_nsamp < 1)) return false;
  return op.ldc >= op.N;
}

template <int TN, int BK, int STAGES, bool SPLIT>
static cudaError_t launch_wg(const GemmGroup& gg, int num_sms, cudaStream_t st) {
  WgGroup g;                             // kernel parameter (about 7 KB), built on the stack: nothing is shared between host threads
  g.n = gg.n;
  // work of an op ~ tiles * rows; share about one wave of CTAs between the ops in that proportion
  double total_work = 0.0;
  for (int i = 0; i < gg.n; ++i)
    total_work += (double)gg.ops[i].K * ((gg.ops[i].M + 127) / 128) * ((gg.ops[i].N + TN - 1) / TN);
  static const double wg_waves = getenv("DINO_WG_WAVES") ? atof(getenv("DINO_WG_WAVES")) : 1.0;   // read once per process
  const double target_ctas = wg_waves * num_sms;
  int z = 0;
  for (int i = 0; i < gg.n; ++i) {
    const GemmOp& o = gg.ops[i];
    WgOp& w = g.op[i];
    w.C = o.C; w.ldc = o.ldc; w.n_in = o.M; w.n_out = o.N; w.rows = o.K;
    w.x_idx = o.a_idx; w.x_rows = o.a_rows; w.x_cps = o.a_rows > 0 ? (o.a_rows + BK - 1) / BK : 0; w.alpha = o.alpha;
    const int chunks = o.a_rows > 0 ? (o.K / o.a_rows) * w.x_cps : (o.K + BK - 1) / BK;
    const int m_tiles = (o.M + 127) / 128;
    w.n_tiles = (o.N + TN - 1) / TN;
    const int tiles = m_tiles * w.n_tiles;
    long long splits = (long long)(target_ctas * ((double)o.K * tiles / total_work) / tiles);     // rounded down: no second wave
    if (splits < 1) splits = 1;
    // fp32x3: at most WG_X3_ROWS reduction rows per CTA (one TMEM accumulator)
    const int cap = WG_X3_ROWS / BK;
    if (SPLIT && splits * cap < chunks) splits = (chunks + cap - 1) / cap;
    if (splits > chunks) splits = chunks;
    w.chunks_per_cta = (int)((chunks + splits - 1) / splits);
    w.splits = (chunks + w.chunks_per_cta - 1) / w.chunks_per_cta;
    w.z_start = z;
    z += tiles * w.splits;
    w.mn = (o.a_sk == 1 && o.b_sk == 1) ? 0 : 1;
    if (o.a_rows > 0) {
      if (!w.mn) return cudaErrorNotSupported;
      if (!tc_make_map3(&g.mx[i], o.A, (unsigned long long)o.M, (unsigned long long)o.a_rows, (unsigned long long)o.a_nsamp, o.a_sk, o.a_ld2, 32, BK, (int)CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B) ||
          !tc_make_map(&g.my[i], o.B, (unsigned long long)o.N, (unsigned long long)o.K, o.b_sk, 32, BK, (int)CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B))
        return cudaErrorInvalidValue;
    } else if (w.mn) {
      // ROW-major [rows, n] with pitch a_sk / b_sk: box = 32 columns x BK reduction rows, 32-byte-atom swizzle
      if (!tc_make_map(&g.mx[i], o.A, (unsigned long long)o.M, (unsigned long long)o.K, o.a_sk, 32, BK, (int)CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B) ||
          !tc_make_map(&g.my[i], o.B, (unsigned long long)o.N, (unsigned long long)o.K, o.b_sk, 32, BK, (int)CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B))
        return cudaErrorInvalidValue;
    } else {
      // transposed [n, rows] with pitch a_sm / b_sn: box = 128 operand rows x 32 reduction elements
      if (TN != 128 || BK != 32) return cudaErrorNotSupported;
      if (!tc_make_map(&g.mx[i], o.A, (unsigned long long)o.K, (unsigned long long)o.M, o.a_sm, 32, 128, (int)CU_TENSOR_MAP_SWIZZLE_128B) ||
          !tc_make_map(&g.my[i], o.B, (unsigned long long)o.K, (unsigned long long)o.N, o.b_sn, 32, 128, (int)CU_TENSOR_MAP_SWIZZLE_128B))
        return cudaErrorInvalidValue;
    }
  }
  constexpr int RAW = (4 + TN / 32) * BK * 128;
  constexpr int SMEM = STAGES * (SPLIT ? 2 * RAW : RAW) + 128 + 1024;
  static_assert(SMEM <= 232448, "shared memory budget of one CTA");
  static std::atomic<unsigned long long> mask{0};
  cudaError_t e = tc_set_smem_once(reinterpret_cast<const void*>(wgrad_tc_kernel<TN, BK, STAGES, SPLIT>), SMEM, mask);
  if (e != cudaSuccess) return e;
  wgrad_tc_kernel<TN, BK, STAGES, SPLIT><<<z, SPLIT ? 192 + WG_NCONV * 32 : 192, SMEM, st>>>(g);
  return cudaGetLastError();
}

static cudaError_t launch_wgrad_tc_rest(const GemmGroup& gg, int split, int num_sms, cudaStream_t st);

template <int BK, int STAGES, bool SPLIT>
static cudaError_t launch_wg2(const GemmGroup& gg, int num_sms, cudaStream_t st) {
  WgGroup g;
  g.n = gg.n;
  double total_work = 0.0;
  for (int i = 0; i < gg.n; ++i)
    total_work += (double)gg.ops[i].K * ((gg.ops[i].M + 255) / 256) * ((gg.ops[i].N + 255) / 256);
  const double target_pairs = num_sms / 2;
  int z = 0;                               // in pairs
  for (int i = 0; i < gg.n; ++i) {
    const GemmOp& o = gg.ops[i];
    WgOp& w = g.op[i];
    w.C = o.C; w.ldc = o.N; w.n_in = o.M; w.n_out = o.N; w.rows = o.K;
    w.x_idx = nullptr; w.x_rows = 0; w.x_cps = 0; w.alpha = 1.f;      // (gathered / scaled ops never come here)
    const int chunks = (o.K + BK - 1) / BK;
    w.n_tiles = (o.N + 255) / 256;
    const int tiles = ((o.M + 255) / 256) * w.n_tiles;
    long long splits = (long long)(target_pairs * ((double)o.K * tiles / total_work) / tiles);
    if (splits < 1) splits = 1;
    const int cap = WG_X3_ROWS / BK;
    if (SPLIT && splits * cap < chunks) splits = (chunks + cap - 1) / cap;
    if (splits > chunks) splits = chunks;
    w.chunks_per_cta = (int)((chunks + splits - 1) / splits);
    w.splits = (chunks + w.chunks_per_cta - 1) / w.chunks_per_cta;
    w.z_start = z;
    z += tiles * w.splits;
    w.mn = 1;
    if (!tc_make_map(&g.mx[i], o.A, (unsigned long long)o.M, (unsigned long long)o.K, o.a_sk, 32, BK, (int)CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B) ||
        !tc_make_map(&g.my[i], o.B, (unsigned long long)o.N, (unsigned long long)o.K, o.b_sk, 32, BK, (int)CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B))
      return cudaErrorInvalidValue;
  }
  constexpr int RAW = 8 * BK * 128;
  constexpr int SMEM = STAGES * (SPLIT ? 2 * RAW : RAW) + 128 + 1024;
  static_assert(SMEM <= 232448, "shared memory budget of one CTA");
  static std::atomic<unsigned long long> mask{0};
  cudaError_t e = tc_set_smem_once(reinterpret_cast<const void*>(wgrad_tc2_kernel<BK, STAGES, SPLIT>), SMEM, mask);
  if (e != cudaSuccess) return e;
  wgrad_tc2_kernel<BK, STAGES, SPLIT><<<2 * z, SPLIT ? 192 + WG_NCONV * 32 : 192, SMEM, st>>>(g);
  return cudaGetLastError();
}

cudaError_t launch_wgrad_tc(const GemmGroup& gg, int split, int num_sms, cudaStream_t st) {
  if (gg.n <= 0) return cudaSuccess;
  // layers wider than 128 both ways go to the CTA-pair kernel (row-major factors only); the rest stays here
  static const int pairs_on = getenv("DINO_WG_CG2") ? atoi(getenv("DINO_WG_CG2")) : 1;     // read once per process
  if (pairs_on) {
    GemmGroup big, rest;
    big.n = rest.n = 0;
    for (int i = 0; i < gg.n; ++i) {
      const GemmOp& o = gg.ops[i];
      const bool mn = !(o.a_sk == 1 && o.b_sk == 1);
      if (mn && o.a_rows == 0 && o.alpha == 1.f && o.ldc == o.N && o.M > 128 && o.N > 128 && o.K >= 2048) big.ops[big.n++] = o; else rest.ops[rest.n++] = o;
    }
    if (big.n > 0) {
      cudaError_t e = split ? launch_wg2<32, 3, true>(big, num_sms, st) : launch_wg2<64, 3, false>(big, num_sms, st);
      // (tf32 with 64 reduction rows per stage: the row-major operands arrive as 32-column boxes, eight per CTA and
      //  stage, and the rate at which one thread can issue those boxes - not bandwidth - paced the 32-row version:
      //  142 -> 86 us at 25600 rows, 1024 x 1024, cuBLAS TF32 90 us)
      if (e != cudaSuccess || rest.n == 0) return e;
      return launch_wgrad_tc_rest(rest, split, num_sms, st);
    }
  }
  return launch_wgrad_tc_rest(gg, split, num_sms, st);
}

static cudaError_t launch_wgrad_tc_rest(const GemmGroup& gg, int split, int num_sms, cudaStream_t st) {
  // outputs wider than 128 take 128 x 256 tiles (row-major factors only: transposed, K-major factors exist for the 128-wide
  // kernels); everything else 128 x 128.  The group is split by width: a 100-wide layer in a 256-wide tile would spend
  // 60 % of its tensor cycles on padding (the 4225 x 100 first layer of the full-Jacobian network shares its step with a
  // 100 x 400 output layer)
  GemmGroup wide, narrow;
  wide.n = narrow.n = 0;
  for (int i = 0; i < gg.n; ++i) {
    const GemmOp& o = gg.ops[i];
    const bool kmajor = o.a_sk == 1 && o.b_sk == 1;
    if (o.N > 128 && !kmajor) wide.ops[wide.n++] = o; else narrow.ops[narrow.n++] = o;
  }
  if (wide.n > 0) {
    cudaError_t e = split ? launch_wg<256, 16, 4, true>(wide, num_sms, st) : launch_wg<256, 32, 4, false>(wide, num_sms, st);
    if (e != cudaSuccess) return e;
  }
  if (narrow.n > 0) return split ? launch_wg<128, 32, 3, true>(narrow, num_sms, st) : launch_wg<128, 32, 3, false>(narrow, num_sms, st);
  return cudaSuccess;
}

}  // namespace dino
