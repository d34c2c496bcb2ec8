// fp32 SIMT GEMM with the fused H1-step epilogue (exact-parity engine).
//
// Every dense contraction of the step (MLP forward/backward, the per-sample Jacobian chain, its
// data-gradient and its split-K weight gradient) is one GemmOp (common.cuh).  This engine
// computes it with FFMA only, so the result is fp32-exact up to summation order; the tcgen05
// engine (gemm_tc.cu) replaces it for the large chain contractions in the tf32 precision mode.
//
// Tiling: BMxBNx16 per CTA, 256 threads, TMxTN register micro-tile, double-buffered shared
// memory with register prefetch (one __syncthreads per K step).  Operands may be contiguous in
// either of their two indices (the four NN/NT/TN/TT layouts), chosen from the strides.
#include "common.cuh"

namespace dino {

constexpr int BK = 16;
constexpr int NT = 256;

template <int BM, int BN, int TM, int TN, bool A_KC, bool B_NC>
__device__ __forceinline__ void gemm_simt_tile(const GemmOp& op, const int bx, const int by, const int bz) {
  static_assert((BM / TM) * (BN / TN) == NT, "thread layout");
  constexpr int LDA_S = BM + 4, LDB_S = BN + 4;
  constexpr int A_PER_T = BM * BK / NT, B_PER_T = BK * BN / NT;
  constexpr int TILE_FLOATS = 2 * BK * LDA_S + 2 * BK * LDB_S, STAGE_FLOATS = BM * (BN + 1);
  __shared__ __align__(16) float smem[TILE_FLOATS > STAGE_FLOATS ? TILE_FLOATS : STAGE_FLOATS];
  float* As = smem;
  float* Bs = smem + 2 * BK * LDA_S;

  const int tid = threadIdx.x;
  const int tx = tid % (BN / TN), ty = tid / (BN / TN);
  const int m0 = by * BM, n0 = bx * BN;
  const int kbeg = bz * op.k_per_split;
  const int kend = min(op.K, kbeg + op.k_per_split);
  const int nk = (kend - kbeg + BK - 1) / BK;

  float acc[TM][TN];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

  float ra[A_PER_T], rb[B_PER_T];

  auto load_regs = [&](int k0) {
#pragma unroll
    for (int j = 0; j < A_PER_T; ++j) {
      const int e = tid + j * NT;
      const int m = A_KC ? e / BK : e % BM;
      const int k = A_KC ? e % BK : e / BM;
      const int gm = m0 + m, gk = k0 + k;
      float v = 0.f;
      if (gm < op.M && gk < kend) {
        const int r = op.a_mod > 0 ? gm % op.a_mod : gm;
        v = __ldg(op.A + (long long)r * op.a_sm + (long long)gk * op.a_sk);
      }
      ra[j] = v;
    }
#pragma unroll
    for (int j = 0; j < B_PER_T; ++j) {
      const int e = tid + j * NT;
      const int n = B_NC ? e % BN : e / BK;
      const int k = B_NC ? e / BN : e % BK;
      const int gn = n0 + n, gk = k0 + k;
      float v = 0.f;
      if (gn < op.N && gk < kend) {
        const int kr = op.b_kmod > 0 ? gk % op.b_kmod : gk;
        v = __ldg(op.B + (long long)kr * op.b_sk + (long long)gn * op.b_sn);
      }
      rb[j] = v;
    }
  };
  auto store_smem = [&](int buf) {
    float* as = As + buf * BK * LDA_S;
    float* bs = Bs + buf * BK * LDB_S;
#pragma unroll
    for (int j = 0; j < A_PER_T; ++j) {
      const int e = tid + j * NT;
      const int m = A_KC ? e / BK : e % BM;
      const int k = A_KC ? e % BK : e / BM;
      as[k * LDA_S + m] = ra[j];
    }
#pragma unroll
    for (int j = 0; j < B_PER_T; ++j) {
      const int e = tid + j * NT;
      const int n = B_NC ? e % BN : e / BK;
      const int k = B_NC ? e / BN : e % BK;
      bs[k * LDB_S + n] = rb[j];
    }
  };

  if (nk > 0) {
    load_regs(kbeg);
    store_smem(0);
  }
  __syncthreads();
  for (int kt = 0; kt < nk; ++kt) {
    const int buf = kt & 1;
    if (kt + 1 < nk) load_regs(kbeg + (kt + 1) * BK);
    const float* as = As + buf * BK * LDA_S + ty * TM;
    const float* bs = Bs + buf * BK * LDB_S + tx * TN;
#pragma unroll 4
    for (int kk = 0; kk < BK; ++kk) {
      float a[TM], b[TN];
      if constexpr (TM % 4 == 0) {
#pragma unroll
        for (int i = 0; i < TM; i += 4) {
          const float4 t = *reinterpret_cast<const float4*>(as + kk * LDA_S + i);
          a[i] = t.x; a[i + 1] = t.y; a[i + 2] = t.z; a[i + 3] = t.w;
        }
      } else {
#pragma unroll
        for (int i = 0; i < TM; i += 2) {
          const float2 t = *reinterpret_cast<const float2*>(as + kk * LDA_S + i);
          a[i] = t.x; a[i + 1] = t.y;
        }
      }
#pragma unroll
      for (int j = 0; j < TN; j += 4) {
        const float4 t = *reinterpret_cast<const float4*>(bs + kk * LDB_S + j);
        b[j] = t.x; b[j + 1] = t.y; b[j + 2] = t.z; b[j + 3] = t.w;
      }
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    if (kt + 1 < nk) store_smem(buf ^ 1);
    __syncthreads();
  }

  // ---- stage the accumulator tile in shared memory -----------------------------------------------
  // The main-loop buffers are free now.  Everything after this point is written as ROLLED loops over
  // the staged tile: the epilogue runs once per launch, so straight-line unrolled code would cost an
  // instruction fetch per instruction executed (measured: profiles/r01a_summary.md), and the staged
  // order also makes every global access row-contiguous.
  constexpr int LDC_S = BN + 1;
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN; ++j) smem[(ty * TM + i) * LDC_S + tx * TN + j] = acc[i][j];
  __syncthreads();
  const int rows_here = min(BM, op.M - m0);
  const int cols_here = min(BN, op.N - n0);

  // ---- rowsum[b,n] += sum over the sample's rows of acc * yref  (pre-epilogue accumulators) ----
  if (op.rowsum != nullptr) {
    constexpr int GROUPS = NT / BN, ROWS_PER_GROUP = BM / GROUPS;
    const int n = tid % BN, rg = tid / BN;
    if (n < cols_here) {
      const int gn = n0 + n;
      const int r_end = min(rows_here, (rg + 1) * ROWS_PER_GROUP);
      float part = 0.f;
      int cur_b = -1;
#pragma unroll 1
      for (int r0 = rg * ROWS_PER_GROUP; r0 < r_end; r0 += 4) {
        float y[4];                       // batch the global loads: 4 independent requests in flight
#pragma unroll
        for (int u = 0; u < 4; ++u)
          y[u] = (r0 + u < r_end) ? op.yref[(long long)(m0 + r0 + u) * op.ld_y + gn] : 0.f;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int r = r0 + u;
          if (r < r_end) {
            const int b = (m0 + r) / op.sdim;
            if (b != cur_b) {
              if (cur_b >= 0) atomicAdd(op.rowsum + (long long)cur_b * op.ld_rs + gn, part);
              cur_b = b;
              part = 0.f;
            }
            part = fmaf(smem[r * LDC_S + n], y[u], part);
          }
        }
      }
      if (cur_b >= 0) atomicAdd(op.rowsum + (long long)cur_b * op.ld_rs + gn, part);
    }
  }

  // ---- epilogue -------------------------------------------------------------------------------
  if (op.C == nullptr) return;
  const int total = rows_here * BN;
#pragma unroll 1
  for (int e0 = tid; e0 < total; e0 += 4 * NT) {
    // gather every operand of 4 tile elements first (independent loads in flight), then finish them
    float v[4], bi[4], ap[4], sc[4], apo[4];
    bool ok[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int e = e0 + u * NT;
      const int r = e / BN, n = e % BN;
      ok[u] = e < total && n < cols_here;
      const int gm = m0 + r, gn = n0 + n;
      v[u] = ok[u] ? smem[r * LDC_S + n] : 0.f;
      bi[u] = (ok[u] && op.bias != nullptr) ? op.bias[gn] : 0.f;
      ap[u] = (ok[u] && op.add_pre != nullptr) ? op.add_pre[(long long)gm * op.ld_ap + gn] : 0.f;
      sc[u] = (ok[u] && op.scale != nullptr) ? op.scale[(long long)(gm / op.sdim) * op.ld_sc + gn] : 1.f;
      apo[u] = (ok[u] && op.add_post != nullptr) ? op.add_post[(long long)gm * op.ld_apo + gn] : 0.f;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      if (!ok[u]) continue;
      const int e = e0 + u * NT;
      const int gm = m0 + e / BN, gn = n0 + e % BN;
      float x = v[u] + bi[u];
      if (op.act) {
        if (op.s_out != nullptr) op.s_out[(long long)gm * op.ld_s + gn] = sigmoid_f(x);
        x = softplus_f(x);
      }
      x = (x + ap[u]) * sc[u] + apo[u];
      if (op.round_tf32) x = round_tf32_f(x);
      float* c = op.C + (long long)gm * op.ldc + gn;
      if (op.atomic_c) atomicAdd(c, x);
      else *c = x;
    }
  }
}

template <int BM, int BN, int TM, int TN, bool A_KC, bool B_NC>
__global__ void __launch_bounds__(NT) gemm_simt_kernel(const GemmOp op) {
  gemm_simt_tile<BM, BN, TM, TN, A_KC, B_NC>(op, blockIdx.x, blockIdx.y, blockIdx.z);
}

// Several independent split-K contractions (the weight gradients of every layer) in ONE launch:
// blockIdx.z enumerates (op, K-split) pairs, CTAs outside an op's tile range exit.
template <int BM, int BN, int TM, int TN, bool A_KC, bool B_NC>
__global__ void __launch_bounds__(NT) gemm_simt_grouped_kernel(const __grid_constant__ GemmGroup g) {
  int i = 0;
  while (i + 1 < g.n && (int)blockIdx.z >= g.z_start[i + 1]) ++i;
  const GemmOp& op = g.ops[i];
  if ((int)blockIdx.x * BN >= op.N || (int)blockIdx.y * BM >= op.M) return;
  gemm_simt_tile<BM, BN, TM, TN, A_KC, B_NC>(op, blockIdx.x, blockIdx.y, blockIdx.z - g.z_start[i]);
}

template <int BM, int BN, int TM, int TN>
static cudaError_t launch_cfg(const GemmOp& op, int splits, cudaStream_t st) {
  dim3 grid((op.N + BN - 1) / BN, (op.M + BM - 1) / BM, splits);
  const bool akc = (op.a_sk == 1), bnc = (op.b_sn == 1);
  if (akc && bnc) gemm_simt_kernel<BM, BN, TM, TN, true, true><<<grid, NT, 0, st>>>(op);
  else if (akc && !bnc) gemm_simt_kernel<BM, BN, TM, TN, true, false><<<grid, NT, 0, st>>>(op);
  else if (!akc && bnc) gemm_simt_kernel<BM, BN, TM, TN, false, true><<<grid, NT, 0, st>>>(op);
  else gemm_simt_kernel<BM, BN, TM, TN, false, false><<<grid, NT, 0, st>>>(op);
  return cudaGetLastError();
}

cudaError_t launch_gemm_simt(const GemmOp& op_in, int num_sms, cudaStream_t st) {
  GemmOp op = op_in;
  if (op.M <= 0 || op.N <= 0) return cudaSuccess;
  if (op.sdim <= 0) op.sdim = 1;
  // largest row tile that still gives every SM a CTA (small-M contractions: the MLP of a wide network)
  const int ntile_n = (op.N + 63) / 64;
  int BM = 128;
  if (op.M <= 32 || ((op.M + 127) / 128) * ntile_n < num_sms) BM = 64;
  if (op.M <= 32 || (BM == 64 && ((op.M + 63) / 64) * ntile_n < num_sms)) BM = 32;
  if (op.M > 32 && op.M <= 64 && BM == 128) BM = 64;
  int splits = 1;
  if (op.atomic_c && op.k_per_split == 0) {
    const int tiles = ((op.M + BM - 1) / BM) * ((op.N + 63) / 64);
    const int kchunks = (op.K + BK - 1) / BK;
    splits = (2 * num_sms + tiles - 1) / tiles;
    if (splits > (kchunks + 1) / 2) splits = (kchunks + 1) / 2;
    if (splits < 1) splits = 1;
  }
  if (op.k_per_split == 0) {
    int kps = (op.K + splits - 1) / splits;
    kps = ((kps + BK - 1) / BK) * BK;
    if (kps < BK) kps = BK;
    op.k_per_split = kps;
  }
  splits = (op.K + op.k_per_split - 1) / op.k_per_split;
  if (splits < 1) splits = 1;
  if (splits > 1 && !op.atomic_c) return cudaErrorInvalidValue;
  if (BM == 32) return launch_cfg<32, 64, 2, 4>(op, splits, st);
  if (BM == 64) return launch_cfg<64, 64, 4, 4>(op, splits, st);
  return launch_cfg<128, 64, 8, 4>(op, splits, st);
}

// All ops must be atomic-accumulate weight gradients with the same operand layout (A M-contiguous,
// B N-contiguous).  Splits are sized so that the whole group fills the GPU about twice.
cudaError_t launch_gemm_simt_grouped(GemmGroup& g, int num_sms, cudaStream_t st) {
  if (g.n <= 0) return cudaSuccess;
  int maxM = 0, maxN = 0;
  long long total_work = 0;
  for (int i = 0; i < g.n; ++i) {
    const GemmOp& op = g.ops[i];
    if (!op.atomic_c || op.a_sm != 1 || op.b_sn != 1) return cudaErrorInvalidValue;
    maxM = op.M > maxM ? op.M : maxM;
    maxN = op.N > maxN ? op.N : maxN;
    total_work += (long long)op.K;
  }
  const int BM = 64;     // weight matrices are 50..100 wide at the headline configs: 64x64 tiles waste least
  int z = 0, tiles_xy_max = 1;
  for (int i = 0; i < g.n; ++i) {
    GemmOp& op = g.ops[i];
    const int tiles = ((op.M + BM - 1) / BM) * ((op.N + 63) / 64);
    tiles_xy_max = tiles > tiles_xy_max ? tiles : tiles_xy_max;
    const int kchunks = (op.K + BK - 1) / BK;
    // share 2*num_sms CTAs between the ops in proportion to their reduction length
    int splits = (int)((2LL * num_sms * op.K + total_work * tiles - 1) / (total_work * tiles));
    if (splits > (kchunks + 1) / 2) splits = (kchunks + 1) / 2;
    if (splits < 1) splits = 1;
    int kps = (op.K + splits - 1) / splits;
    kps = ((kps + BK - 1) / BK) * BK;
    op.k_per_split = kps < BK ? BK : kps;
    op.sdim = op.sdim > 0 ? op.sdim : 1;
    g.z_start[i] = z;
    z += (op.K + op.k_per_split - 1) / op.k_per_split;
  }
  g.z_start[g.n] = z;
  dim3 grid((maxN + 63) / 64, (maxM + BM - 1) / BM, z);
  gemm_simt_grouped_kernel<64, 64, 4, 4, false, true><<<grid, NT, 0, st>>>(g);
  return cudaGetLastError();
}

}  // namespace dino
