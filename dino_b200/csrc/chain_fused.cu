// Fused Jacobian-chain kernel (tcgen05 + TMA): network forward, forward chain, H1 loss and backward chain for a
// tile of samples in ONE CTA, for networks whose widths fit one UMMA tile (<= 128; every BASELINE reduced-basis
// config with r <= 128).
//
// A CTA owns G = floor(128/s) whole samples = up to 128 chain rows (s seed rows per sample; a sample's rows start at
// a multiple of 32 when the tile has room, so a warp never straddles two samples).  The current chain state X
// [128 rows, K <= 128] lives in shared memory in the UMMA K-major SWIZZLE_128B layout; it is the A operand of every
// contraction.  Roles: 8 epilogue warps (thread = chain row = TMEM lane, two threads per row splitting the column
// blocks), four row-sum warps, one control warp (an elected lane streams the TF32 weight tiles by TMA one step ahead and issues the
// tcgen05.mma of each step into one of two TMEM accumulators), one store warp (an elected lane TMA-stores every
// step result from X to its row-major array in global memory: the weight-gradient launch reads those arrays as
// MN-major operands, and the backward steps re-read Q_l from them).
//
//   network  l = 0..L-1 :  z_{l+1} = act(z_l W_l + b_l), sg_l = sigmoid(.)    (fp32 FMAs, 1 thread per output; the fp32
//                          weights arrive by bulk copy in the still idle weight-tile buffers) ; output loss
//   forward  l = L-1..0 :  R_l = X W_l^T (+ Q_skip) ;  X <- Q_{l-1} = R_l * sg_{l-1}      (X_0 = seed)
//   loss                :  X <- G = coef (R_0 - J) / ||J_b||^2 ; statistics
//   backward l = 0..L-1 :  dQ_l = X W_l ; rs_l[b,:] += sum_rows dQ_l * Q_l ; X <- dR_{l+1} = dQ_l * sg_l (+ dR_skip)
//
// The epilogue warps only do what the next contraction waits for (accumulator -> X).  The row sums of the backward
// steps belong to four more warps (one per TMEM lane quarter) that re-read the accumulator of step t while step t+1
// fills the other one.
#include <cuda.h>

#include "tc_common.cuh"

namespace dino {

// Per-phase %globaltimer stamps of CTA 0 (debug aid; enabled with -DDINO_CF_TRACE, read by scripts/trace_chain.py)
__device__ unsigned long long g_cf_trace[64 * 8];
#ifdef DINO_CF_TRACE
#define CF_STAMP(t, slot) do { if (blockIdx.x == 0) { unsigned long long _c; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(_c)); g_cf_trace[(t) * 8 + (slot)] = _c; } } while (0)
#else
#define CF_STAMP(t, slot) do {} while (0)
#endif

constexpr int CF_ROWS = 128, CF_MAXW = 128, CF_CHUNK_BYTES = CF_ROWS * 128;       // 16 KB per 32-float K chunk
constexpr int CF_X_BYTES = (CF_MAXW / 32) * CF_CHUNK_BYTES;                        // 64 KB
constexpr int CF_PARTS = 2;                                                        // threads per chain row (column split)
constexpr int CF_EPI = 128 * CF_PARTS;                                             // 8 epilogue warps: 2 per TMEM lane quarter
constexpr int CF_EPI_WARPS = CF_EPI / 32;
constexpr int CF_RS_WARPS = 4;                                                     // row-sum warps: one per TMEM lane quarter
constexpr int CF_THREADS = CF_EPI + 64 + 32 * CF_RS_WARPS + 32;                    // + control warp (MMA issue) + store warp (TMA stores of the results) + row-sum warps + weight warp (TMA loads of the weight tiles)
constexpr int CF_W_WARP = CF_EPI_WARPS + 2 + CF_RS_WARPS;                          // index of the weight warp
constexpr int CF_SLOTS = 8;                                                        // weight ring: 16 KB slots, one per 32-column K chunk of a weight tile (hi or lo)
constexpr int CF_XLO_COL = 2 * 128;                                                // fp32x3: TMEM columns of X_lo, the low part of the operand tile (A operand from tensor memory)
constexpr int CF_CW = 16;                                                          // columns per TMEM load / work item
constexpr int CF_ACC_COLS = 128;                                                   // TMEM columns per accumulator (two of them: steps alternate)
constexpr int CF_TILE_LD = 36;                                                     // floats per column of a warp's row-sum transposition tile

__device__ __forceinline__ uint32_t cf_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void cf_mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(cf_smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void cf_mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(cf_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cf_mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(cf_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void cf_mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "CF_WAIT:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra CF_DONE;\n\t"
      "bra CF_WAIT;\n\t"
      "CF_DONE:\n\t"
      "}" ::"r"(cf_smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void cf_tma_load(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(cf_smem_u32(dst)), "l"(map), "r"(cf_smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void cf_tma_store(const CUtensorMap* map, uint32_t src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
               ::"l"(map), "r"(src), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void cf_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(cf_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void cf_mma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void cf_sts128(uint32_t addr, float a, float b, float c, float d) {
  asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ void cf_sts32(uint32_t addr, float a) {
  asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(a) : "memory");
}
__device__ __forceinline__ float4 cf_lds128(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
  return v;
}
// round-to-nearest (ties away) to the 10-bit TF32 mantissa: what cvt.rna.tf32.f32 does for finite values, without its
// Inf/NaN guard (a non-finite chain state has already lost the step)
__device__ __forceinline__ float cf_round_tf32(float x) {
  return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u);
}
__device__ __forceinline__ bool cf_elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t"
      "}" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ uint64_t cf_desc(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3FFFF) >> 4) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
__device__ __forceinline__ void cf_tmem_ld16(uint32_t taddr, float (&v)[CF_CW]) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[j]);
}
__device__ __forceinline__ void cf_tmem_ld32(uint32_t taddr, float (&v)[32]) {
  uint32_t r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
}
__device__ __forceinline__ void cf_tmem_st16(uint32_t taddr, const float (&v)[CF_CW]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
      ::"r"(taddr), "r"(__float_as_uint(v[0])), "r"(__float_as_uint(v[1])), "r"(__float_as_uint(v[2])), "r"(__float_as_uint(v[3])),
        "r"(__float_as_uint(v[4])), "r"(__float_as_uint(v[5])), "r"(__float_as_uint(v[6])), "r"(__float_as_uint(v[7])),
        "r"(__float_as_uint(v[8])), "r"(__float_as_uint(v[9])), "r"(__float_as_uint(v[10])), "r"(__float_as_uint(v[11])),
        "r"(__float_as_uint(v[12])), "r"(__float_as_uint(v[13])), "r"(__float_as_uint(v[14])), "r"(__float_as_uint(v[15]))
      : "memory");
}
__device__ __forceinline__ void cf_tmem_st4(uint32_t taddr, float a, float b, float c, float d) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};"
               ::"r"(taddr), "r"(__float_as_uint(a)), "r"(__float_as_uint(b)), "r"(__float_as_uint(c)), "r"(__float_as_uint(d)) : "memory");
}
__device__ __forceinline__ void cf_tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// byte offset of (row r, float column k..k+3, k % 4 == 0) inside X (K-major, 128-byte rows, SWIZZLE_128B)
__device__ __forceinline__ uint32_t cf_xoff(int r, int k) {
  const int c = k >> 5, j = (k & 31) >> 2;
  return (uint32_t)(c * CF_CHUNK_BYTES + (r >> 3) * 1024 + (r & 7) * 128 + ((j ^ (r & 7)) << 4));
}
// After the call lane j (and lane j+16) holds the sum over the 32 lanes of v[j & 15]  (31 shuffles).
__device__ __forceinline__ float cf_colsum16(float (&v)[CF_CW], int lane) {
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] += __shfl_xor_sync(0xffffffffu, v[i], 16);
#pragma unroll
  for (int off = 8; off >= 1; off >>= 1) {
    const bool up = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < off; ++i) {
      const float send = up ? v[i] : v[i + off];
      const float keep = up ? v[i + off] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
    }
  }
  return v[0];
}
// 16 consecutive floats of a row (zero beyond N / for invalid rows); vector loads when the row is 16-byte aligned
__device__ __forceinline__ void cf_load16(float (&d)[CF_CW], const float* row, int c0, int N, bool ok, bool vec) {
#pragma unroll
  for (int j = 0; j < 16; j += 4) {
    float4 q4 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (ok && c0 + j < N) {
      if (vec && c0 + j + 3 < N) q4 = *reinterpret_cast<const float4*>(row + c0 + j);
      else {
        q4.x = row[c0 + j];
        if (c0 + j + 1 < N) q4.y = row[c0 + j + 1];
        if (c0 + j + 2 < N) q4.z = row[c0 + j + 2];
        if (c0 + j + 3 < N) q4.w = row[c0 + j + 3];
      }
    }
    d[j] = q4.x; d[j + 1] = q4.y; d[j + 2] = q4.z; d[j + 3] = q4.w;
  }
}

// 16 consecutive floats of a row (zero at and beyond `len`; `len`, the row pitch and c0 are multiples of 4)
__device__ __forceinline__ void cf_load_vec16(float (&d)[CF_CW], const float* row, int c0, int len, bool ok = true) {
#pragma unroll
  for (int j = 0; j < CF_CW; j += 4) {
    float4 q4 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (ok && c0 + j < len) q4 = *reinterpret_cast<const float4*>(row + c0 + j);
    d[j] = q4.x; d[j + 1] = q4.y; d[j + 2] = q4.z; d[j + 3] = q4.w;
  }
}
__device__ __forceinline__ void cf_store_vec16(float* row, int c0, int len, const float (&x)[CF_CW]) {
#pragma unroll
  for (int j = 0; j < CF_CW; j += 4)
    if (c0 + j < len) *reinterpret_cast<float4*>(row + c0 + j) = make_float4(x[j], x[j + 1], x[j + 2], x[j + 3]);
}

// the result of a step goes to global memory (row-major) when somebody reads it later: the weight-gradient launch
// after a training step, a residual connection, or the caller (the Jacobian itself when no loss is asked for)
__device__ __forceinline__ bool cf_out_needed(const ChainStep& sd, const ChainNetDev& net, int do_loss, int do_backward, int dbg) {
  return sd.out_off >= 0 && (do_backward || net.has_skip || (sd.mode == 1 && !do_loss)) && !(dbg & 1);
}

__global__ void __launch_bounds__(CF_THREADS, 1)
chain_fused_kernel(const __grid_constant__ ChainMaps maps, const __grid_constant__ ChainNetDev net,
                   const __grid_constant__ ChainSteps steps, float* __restrict__ ws, const float* __restrict__ jt,
                   const float* __restrict__ jn2, const int* __restrict__ idx, const float* __restrict__ seed, int B,
                   float coef, int do_loss, int do_backward, float* __restrict__ stats,
                   const __grid_constant__ MlpNetDev mlp, const float* __restrict__ params, const float* __restrict__ m_in,
                   const float* __restrict__ q_tgt, const float* __restrict__ qn2, float coef_q, int mlp_mode, int split) {
  extern __shared__ uint8_t cf_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(cf_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* X = smem;
  uint8_t* Wb[2] = {smem + CF_X_BYTES, smem + 2 * CF_X_BYTES};              // the weight ring (CF_SLOTS x 16 KB); its two halves first serve the MLP phase
  uint64_t* w_full = reinterpret_cast<uint64_t*>(smem + 3 * CF_X_BYTES);     // [2] MLP phase: fp32 weights of layer l landed in Wb[l & 1]
  uint64_t* mma_done = w_full + 2;
  uint64_t* x_ready = w_full + 3;
  uint64_t* x_free = w_full + 4;                                             // the stores of the previous result have read X
  uint64_t* fwd_stored = w_full + 5;                                         // every forward chain state has landed in global memory
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(w_full + 6);
  uint64_t* rs_done = w_full + 8;                                            // [2] the row-sum warps have read accumulator t&1 (step t)
  uint64_t* mma_done1 = w_full + 10;                                         // mma_done of the odd steps
  uint64_t* mlp_done = w_full + 7;                                           // the MLP forward no longer uses the weight-tile buffers
  uint64_t* slot_full = w_full + 16;                                         // [CF_SLOTS] weight ring: chunk landed
  uint64_t* slot_empty = w_full + 16 + CF_SLOTS;                             // [CF_SLOTS] weight ring: the MMAs that read the chunk are complete
  float* ex_s = reinterpret_cast<float*>(w_full + 16 + 2 * CF_SLOTS);        // [CF_PARTS][128][2] loss partials of the row parts
  float* tiles = ex_s + CF_PARTS * 128 * 2;                                   // [CF_EPI_WARPS][8][CF_TILE_LD] row-sum transposition tiles
  float* act_s = tiles + CF_EPI_WARPS * 8 * CF_TILE_LD;                       // [2][16][128] MLP activations of the tile's samples (ping-pong)

  // One completion barrier per accumulator: a waiter that lags a step behind still finds "its" phase in flight or just
  // completed (a parity wait cannot tell phase p from phase p+2 of the same barrier).
  auto mma_bar = [&](int t) { return (t & 1) ? mma_done1 : mma_done; };
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int s = net.s, G = net.G, L = net.L;
  const int b0 = blockIdx.x * G;
  const int nsamp = min(G, B - b0);
  const int srow = net.srow;                     // tile rows reserved per sample (>= s)
  const long long row0 = (long long)b0 * s;      // first global chain row
#ifdef DINO_CF_TRACE
  const int dbg = do_backward >> 8;      // experiment knobs: 1 skip transposed stores, 2 skip X writes, 8 skip rounding
  do_backward &= 0xff;
#else
  const int dbg = 0;
#endif
  const int n_steps = do_backward ? steps.n_all : steps.n_fwd;
  // residual networks re-read chain states inside the kernel right after they are produced: those keep per-thread stores
  const bool tma_out = net.has_skip == 0;

  // ---- set-up --------------------------------------------------------------------------------------
  if (tid == 0) CF_STAMP(0, 7);
  for (int e = tid; e < CF_X_BYTES / 16; e += CF_THREADS) reinterpret_cast<float4*>(X)[e] = make_float4(0.f, 0.f, 0.f, 0.f);
  if (warp == CF_EPI_WARPS) {
    if (lane == 0) {
      cf_mbar_init(w_full, 1); cf_mbar_init(w_full + 1, 1); cf_mbar_init(mma_done, 1); cf_mbar_init(mma_done1, 1); cf_mbar_init(x_ready, CF_EPI_WARPS);
      cf_mbar_init(x_free, 1); cf_mbar_init(fwd_stored, 1); cf_mbar_init(mlp_done, CF_EPI_WARPS); cf_mbar_init(rs_done, CF_RS_WARPS); cf_mbar_init(rs_done + 1, CF_RS_WARPS);
      for (int i = 0; i < CF_SLOTS; ++i) { cf_mbar_init(slot_full + i, 1); cf_mbar_init(slot_empty + i, 1); }
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(cf_smem_u32(tmem_slot)), "r"(split ? 512u : 2u * CF_ACC_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // the zero fill of X must be visible to the tensor core
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;
  if (tid == 0) CF_STAMP(1, 7);

  const int warp_u = __shfl_sync(0xffffffffu, warp, 0);   // tells the compiler that everything derived from it is warp-uniform
  if (warp_u == CF_EPI_WARPS) {
    // ===== control warp: one elected lane issues the weight TMA (one step ahead) and the MMAs =====
    // (elect.sync, not `lane == 0`: the compiler then knows a single lane is active and feeds the uniform
    //  datapath directly instead of looping over possibly-divergent lanes)
    if (cf_elect_one()) {
      const uint64_t xd = cf_desc(cf_smem_u32(X));
      const uint32_t w_u = cf_smem_u32(Wb[0]);
      const uint32_t xlo = tmem_base + (uint32_t)CF_XLO_COL;
      int slot = 0, round = 0;                              // position in the weight ring (the weight warp walks it in the same order)
#pragma unroll 1
      for (int t = 0; t < n_steps; ++t) {
        const int buf = t & 1;
        const int K = steps.st[t].K, N = steps.st[t].N;
        const uint32_t n16 = (uint32_t)((N + 15) & ~15);
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((n16 >> 3) << 17) | ((128u >> 4) << 24);
        const int ksteps = (K + 7) >> 3;
        cf_mbar_wait(x_ready, t & 1);                       // operand X written by the epilogue threads
        CF_STAMP(t, 0);
        if (t >= 2) cf_mbar_wait(rs_done + buf, ((t >> 1) - 1) & 1);     // accumulator t&1 was last read by the row-sum warps in step t-2
        CF_STAMP(t, 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        // descriptor start addresses advance by 32 bytes per 8-wide K step, 16 KB per 32-column chunk (no carry out
        // of the 14-bit address field: the tiles sit below 256 KB)
        const uint32_t dcol = tmem_base + (uint32_t)(buf * CF_ACC_COLS);
        const int nch = (ksteps + 3) >> 2;                  // 32-column chunks of X / the weight tile (four K steps each)
        const int slot0 = slot, round0 = round;             // the step's chunks sit in consecutive ring slots: hi (, lo) per chunk
        auto slot_of = [&](int i) { int sl = slot0 + i; return sl >= CF_SLOTS ? sl - CF_SLOTS : sl; };
        auto round_of = [&](int i) { return slot0 + i >= CF_SLOTS ? round0 + 1 : round0; };
        const int per = split ? 2 : 1;
        if (split) {
          // fp32x3, pass 1: the two small cross terms of every K step, x_hi * w_lo and x_lo * w_hi (the tensor core reads
          // the hi parts out of the fp32 containers).  They go FIRST: the tensor core adds into its fp32 accumulator with
          // truncation, an error of ~2^-24 of the accumulator's magnitude per MMA, and while only cross terms (2^-11 of
          // the result) are in it that error is nil.  The main sum below then costs K/8 truncations, not 3K/8.
#pragma unroll 1
          for (int c = 0; c < nch; ++c) {
            const uint64_t oc = (uint64_t)(c * (CF_CHUNK_BYTES >> 4));
            cf_mbar_wait(slot_full + slot_of(2 * c), round_of(2 * c) & 1);
            cf_mbar_wait(slot_full + slot_of(2 * c + 1), round_of(2 * c + 1) & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint64_t bh = cf_desc(w_u + (uint32_t)(slot_of(2 * c) * CF_CHUNK_BYTES));
            const uint64_t bl = cf_desc(w_u + (uint32_t)(slot_of(2 * c + 1) * CF_CHUNK_BYTES));
#pragma unroll
            for (int j = 0; j < 4; ++j)
              if (c * 4 + j < ksteps) {
                cf_mma(dcol, xd + oc + 2 * j, bl + 2 * j, idesc, (c | j) ? 1u : 0u);
                tc::umma_tf32_ts(dcol, xlo + (uint32_t)(c * 32 + j * 8), bh + 2 * j, idesc, 1u);
              }
          }
        }
#pragma unroll 1
        for (int c = 0; c < nch; ++c) {                     // main products x_hi * w_hi (tf32: the only pass)
          const uint64_t oc = (uint64_t)(c * (CF_CHUNK_BYTES >> 4));
          if (!split) {
            cf_mbar_wait(slot_full + slot_of(c), round_of(c) & 1);       // weight chunk landed
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          }
          const uint64_t bh = cf_desc(w_u + (uint32_t)(slot_of(per * c) * CF_CHUNK_BYTES));
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (c * 4 + j < ksteps) cf_mma(dcol, xd + oc + 2 * j, bh + 2 * j, idesc, (split || (c | j)) ? 1u : 0u);
          cf_commit(slot_empty + slot_of(per * c));         // the slot is free once the MMAs issued so far have read it
          if (split) cf_commit(slot_empty + slot_of(per * c + 1));
        }
        slot = slot0 + per * nch;
        if (slot >= CF_SLOTS) { slot -= CF_SLOTS; ++round; }
        cf_commit(mma_bar(t));
        CF_STAMP(t, 2);
      }
    }
    __syncwarp();
  } else if (warp_u == CF_EPI_WARPS + 1) {
    // ===== store warp: row-major copy of every step result that somebody reads later =====
    // The result of step t sits in X as the operand of step t+1: one elected lane TMA-stores it from there (whole
    // 128-byte lines, no thread touches the data), one box per 32-column chunk (and per sample when rows are warp-aligned).
    if (do_loss) {
      // the target Jacobians of this tile are first needed ~15 us from now (last forward step): pull them into L2
      for (int gI = 0; gI < nsamp; ++gI) {
        const long long sb = idx ? idx[b0 + gI] : b0 + gI;
        const char* base = reinterpret_cast<const char*>(jt + sb * (long long)s * net.d_in);
        const long long bytes = (long long)s * net.d_in * 4;
        for (long long o = (long long)lane * 128; o < bytes; o += 32 * 128)
          asm volatile("prefetch.global.L2 [%0];" ::"l"(base + o));
        if (mlp_mode > 1) {                                  // ... and the output targets of the in-kernel output loss
          const int d_out = mlp.layer[mlp.L - 1].n_out;
          const char* qb = reinterpret_cast<const char*>(q_tgt + sb * d_out);
          if (lane * 128 < d_out * 4) asm volatile("prefetch.global.L2 [%0];" ::"l"(qb + lane * 128));
          if (lane == 31) asm volatile("prefetch.global.L2 [%0];" ::"l"(qn2 + sb));
        }
      }
    }
    if (tma_out && cf_elect_one()) {
      const uint32_t x_u = cf_smem_u32(X);
      auto store_x = [&](const CUtensorMap* map, int N) {
        const int nch = (N + 31) >> 5;
        for (int c = 0; c < nch; ++c) {
          if (srow == s) cf_tma_store(map, x_u + (uint32_t)(c * CF_CHUNK_BYTES), c * 32, (int)row0);
          else
            for (int gI = 0; gI < nsamp; ++gI)
              cf_tma_store(map, x_u + (uint32_t)(c * CF_CHUNK_BYTES + gI * srow * 128), c * 32, (int)(row0 + (long long)gI * s));
        }
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      };
      // X version 0 = the top chain state Q_{L-1} (operand of the top layer's weight gradient)
      cf_mbar_wait(x_ready, 0);
      if (net.top_materialised && do_backward) {
        store_x(&maps.top, net.layer[L - 1].n_out);
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      }
      cf_mbar_arrive(x_free);
#pragma unroll 1
      for (int t = 1; t <= n_steps; ++t) {
        const bool last = t == n_steps;
        const bool need = cf_out_needed(steps.st[t - 1], net, do_loss, do_backward, dbg);
        if (last && !need) break;
        cf_mbar_wait(x_ready, t & 1);                       // X holds the result of step t-1
        if (need) store_x(&maps.out[t - 1], steps.st[t - 1].N);
        if (do_backward && t == steps.n_fwd) {
          // the backward steps re-read the forward chain states with ordinary loads: all of them must have landed
          asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
          cf_mbar_arrive(fwd_stored);
        } else if (need) {
          asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        }
        if (!last) cf_mbar_arrive(x_free);                  // X may be rewritten
      }
      asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");   // shared memory must outlive the stores
    }
    __syncwarp();
  } else if (warp_u == CF_W_WARP) {
    // ===== weight warp: one elected lane streams the weight tiles of every step through the ring, as far ahead as
    // the ring has room (a tile = one 16 KB slot per 32 K columns; fp32x3 brings the lo tile of each chunk as well) =====
    if (cf_elect_one()) {
      // the in-kernel MLP forward borrows the ring's memory first: wait until it is done with it
      if (mlp_mode > 0) {
        cf_mbar_wait(mlp_done, 0);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      }
      int slot = 0, round = 0;
#pragma unroll 1
      for (int t = 0; t < n_steps; ++t) {
        const int nch = (steps.st[t].K + 31) >> 5;
        const CUtensorMap* mh = t < L ? &maps.fwd[L - 1 - t] : &maps.bwd[t - L];
        const CUtensorMap* ml = t < L ? &maps.fwd_lo[L - 1 - t] : &maps.bwd_lo[t - L];
        for (int c = 0; c < nch; ++c) {
          for (int h = 0; h < (split ? 2 : 1); ++h) {
            cf_mbar_wait(slot_empty + slot, (round & 1) ^ 1);
            cf_mbar_expect_tx(slot_full + slot, (uint32_t)CF_CHUNK_BYTES);
            cf_tma_load(Wb[0] + slot * CF_CHUNK_BYTES, h ? ml : mh, c * 32, 0, slot_full + slot);
            if (++slot == CF_SLOTS) { slot = 0; ++round; }
          }
        }
      }
    }
    __syncwarp();
  } else if (warp_u >= CF_EPI_WARPS + 2 && warp_u < CF_W_WARP) {
    // ===== row-sum warps: rs_l[b, n] += sum over the sample's rows of dQ_l * Q_l  (backward steps through a softplus) =====
    // One warp per TMEM lane quarter, all column blocks of its 32 rows.  They re-read the accumulator of step t while the
    // epilogue warps and the tensor core are already on step t+1 (two accumulators; the control warp waits for rs_done
    // before it lets step t+2 overwrite the one read here).
    const int qw = warp_u & 3;
    const int r = qw * 32 + lane;
    const int g_raw = r / srow;
    const int i_seed = r - g_raw * srow;
    const bool row_ok = g_raw < nsamp && i_seed < s;
    const int g = row_ok ? g_raw : 0;
    const long long grow = row0 + (long long)g * s + i_seed;
    const uint32_t taddr = tmem_base + ((uint32_t)(qw * 32) << 16);
    const int g_lo = (32 * qw) / srow, g_hi = min((32 * qw + 31) / srow, nsamp - 1);
    const bool aligned = (srow & 31) == 0;
    // 32 columns x 36 floats per warp: the four tiles run over `tiles` into `act_s`, which the MLP phase no longer needs
    const uint32_t tile_u = cf_smem_u32(tiles) + (uint32_t)((warp_u - CF_EPI_WARPS - 2) * (32 * CF_TILE_LD) * 4);
#pragma unroll 1
    for (int t = 0; t < n_steps; ++t) {
      const ChainStep& sd = steps.st[t];
      const int N = sd.N;
      const bool work = sd.mode == 2 && !(dbg & 16);
      if (tma_out && do_backward && t == steps.n_fwd) cf_mbar_wait(fwd_stored, 0);   // Q_l are read from global memory from here on
      const float* const q_p = ws + sd.q_off + grow * sd.q_ld;         // Q_l rows
      float* const rs_p = ws + sd.rs_off + (long long)b0 * sd.rs_ld;   // row sums of the layer [B, rs_ld] (zeroed by the caller; added to)
      // 32 columns per pass: Q_l[row, c0 .. c0+32) (two 64-byte halves per row)
      float q[32];
      auto load_q32 = [&](int c0) {
        float h0[CF_CW], h1[CF_CW];
        cf_load_vec16(h0, q_p, c0, sd.q_ld, row_ok);
        cf_load_vec16(h1, q_p, c0 + CF_CW, sd.q_ld, row_ok);
#pragma unroll
        for (int j = 0; j < CF_CW; ++j) { q[j] = h0[j]; q[CF_CW + j] = h1[j]; }
      };
#pragma unroll
      for (int j = 0; j < 32; ++j) q[j] = 0.f;
      if (work) load_q32(0);                                            // first pass, requested while the MMA runs
      cf_mbar_wait(mma_bar(t), (t >> 1) & 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      if (work) {
        const uint32_t acc = taddr + (uint32_t)((t & 1) * CF_ACC_COLS);
#pragma unroll 1
        for (int c0 = 0; c0 < N; c0 += 32) {
          float v[32];
          cf_tmem_ld32(acc + (uint32_t)c0, v);
          if (aligned) {
            // the warp's 32 rows belong to one sample: transpose the 32 columns through the warp's tile, lane (c, i)
            // adds rows 8i..8i+7 of column c of each 8-column group; two shuffles per group finish (four independent
            // chains), lane j publishes column c0 + j
#pragma unroll
            for (int j = 0; j < 32; ++j) cf_sts32(tile_u + (uint32_t)((j * CF_TILE_LD + lane) * 4), v[j] * q[j]);
            if (c0 + 32 < N) load_q32(c0 + 32);                         // next pass, in flight during the reduction
            __syncwarp();
            const uint32_t tp = tile_u + (uint32_t)(((lane & 7) * CF_TILE_LD + (lane >> 3) * 8) * 4);
            float sg4[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const float4 a = cf_lds128(tp + (uint32_t)(i * 8 * CF_TILE_LD * 4)), c = cf_lds128(tp + (uint32_t)(i * 8 * CF_TILE_LD * 4) + 16);
              sg4[i] = ((a.x + a.y) + (a.z + a.w)) + ((c.x + c.y) + (c.z + c.w));
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) sg4[i] += __shfl_xor_sync(0xffffffffu, sg4[i], 8);
#pragma unroll
            for (int i = 0; i < 4; ++i) sg4[i] += __shfl_xor_sync(0xffffffffu, sg4[i], 16);
            const int grp = lane >> 3;
            const float sum = grp == 0 ? sg4[0] : grp == 1 ? sg4[1] : grp == 2 ? sg4[2] : sg4[3];
            if (g_lo < nsamp && c0 + lane < N) atomicAdd(rs_p + (long long)g_lo * sd.rs_ld + c0 + lane, sum);
            __syncwarp();
          } else {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              if (c0 + CF_CW * h < N) {
                for (int gg = g_lo; gg <= g_hi; ++gg) {           // segmented over the samples under this warp's rows
                  float m[CF_CW];
#pragma unroll
                  for (int j = 0; j < CF_CW; ++j) m[j] = (row_ok && g == gg) ? v[CF_CW * h + j] * q[CF_CW * h + j] : 0.f;
                  const float tot = cf_colsum16(m, lane);
                  if (lane < 16 && c0 + CF_CW * h + lane < N) atomicAdd(rs_p + (long long)gg * sd.rs_ld + c0 + CF_CW * h + lane, tot);
                }
              }
            }
            if (c0 + 32 < N) load_q32(c0 + 32);
          }
        }
      }
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) cf_mbar_arrive(rs_done + (t & 1));
    }
  } else {
    // ===== epilogue threads: CF_PARTS threads per chain row (= TMEM lane), alternating 16-column blocks =====
    // Per step: accumulator -> per-sample scaling / residual / loss gradient -> next operand X, then X is handed to
    // the tensor core (and to the store warp).  The row sums of the backward steps are the row-sum warps' business.
    const int qw = warp_u & 3;                              // TMEM lane quarter of this warp
    const int part = warp_u >> 2;
    const int r = qw * 32 + lane;
    const int g_raw = r / srow;                             // local sample
    const int i_seed = r - g_raw * srow;                    // seed row within the sample
    const bool row_ok = g_raw < nsamp && i_seed < s;
    const int g = row_ok ? g_raw : 0;
    const long long b = b0 + g;                             // batch sample
    const long long grow = row0 + (long long)g * s + i_seed;   // global chain row
    const uint32_t taddr = tmem_base + ((uint32_t)(qw * 32) << 16);
    const uint32_t xrow = cf_smem_u32(X) + (uint32_t)((r >> 3) * 1024 + (r & 7) * 128);
    const int xsw = r & 7;
    constexpr int STRIDE = CF_CW * CF_PARTS;
    const int c_first = part * CF_CW;
    const uint32_t xlo_t = taddr + (uint32_t)CF_XLO_COL;   // this row's X_lo columns (fp32x3)
    if (split) {
      // X_lo starts as zeros (the contractions read whole 8-column K steps of it); each row part clears its half
      float z16[CF_CW];
#pragma unroll
      for (int j = 0; j < CF_CW; ++j) z16[j] = 0.f;
#pragma unroll
      for (int c0 = 0; c0 < CF_MAXW / CF_PARTS; c0 += CF_CW) cf_tmem_st16(xlo_t + (uint32_t)(part * (CF_MAXW / CF_PARTS) + c0), z16);
      cf_tmem_st_wait();
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      asm volatile("bar.sync 1, %0;" ::"n"(CF_EPI) : "memory");
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    }

    // ---- the network itself for the tile's samples: z_{l+1} = act(z_l W_l + b_l), plain fp32 FMAs ----
    // (a few thousand multiply-adds per sample: not worth a launch of its own.  One thread per (sample, output column),
    //  activations ping-pong in shared memory, the fp32 weights of layer l+1 / l+2 arrive by bulk copy in the two
    //  weight-tile buffers -- idle until the chain starts -- while layer l is computed.)
    if (mlp_mode > 0) {
      float* a_cur = act_s;
      float* a_nxt = act_s + 16 * 128;
      auto fetch_w = [&](int l) {
        const MlpLayerDev& ly = mlp.layer[l];
        const uint32_t bytes = (uint32_t)(((ly.n_in * ly.n_out + 3) & ~3) * 4);
        cf_mbar_expect_tx(w_full + (l & 1), bytes);
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(cf_smem_u32(Wb[l & 1])), "l"(params + ly.koff), "r"(bytes), "r"(cf_smem_u32(w_full + (l & 1))) : "memory");
      };
      if (tid == 0) { fetch_w(0); if (mlp.L > 1) fetch_w(1); }
      for (int e = tid; e < nsamp * 128; e += CF_EPI) {
        const int gI = e >> 7, k = e & 127;
        float v = 0.f;
        if (k < mlp.d_in) {
          const long long row = idx ? idx[b0 + gI] : (b0 + gI);
          v = __ldg(m_in + row * mlp.d_in + k);
          ws[mlp.z0_off + (long long)(b0 + gI) * mlp.ld_in + k] = v;
        }
        a_cur[e] = v;
      }
      asm volatile("bar.sync 1, %0;" ::"n"(CF_EPI) : "memory");
#pragma unroll 1
      for (int l = 0; l < mlp.L; ++l) {
        const MlpLayerDev& ly = mlp.layer[l];
        const float* W = reinterpret_cast<const float*>(Wb[l & 1]);
        const int n_in = ly.n_in, n_out = ly.n_out;
        // 256 threads = 4 samples x 64 columns per pass, or 2 samples x 128 columns for the wider layers
        const int cw = n_out > 64 ? 128 : 64, su = CF_EPI / cw;
        const int cl = tid & (cw - 1), u = tid / cw;
        cf_mbar_wait(w_full + (l & 1), (l >> 1) & 1);
#pragma unroll 1
        for (int c0 = 0; c0 < n_out; c0 += cw) {
          const int c = c0 + cl;
#pragma unroll 1
          for (int g0 = 0; g0 < nsamp; g0 += su) {
            const int gI = g0 + u;
            if (c < n_out && gI < nsamp) {
              const float* a = a_cur + gI * 128;
              const float* w = W + c;
              const float bias = ly.boff >= 0 ? __ldg(params + ly.boff + c) : 0.f;   // an L2 round trip: consumed after the sums
              float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, acc3 = 0.f;
              int k = 0;
#pragma unroll 2
              for (; k + 3 < n_in; k += 4) {                  // four chains: the FMA latency is the cost here
                const float4 a4 = *reinterpret_cast<const float4*>(a + k);
                acc0 = fmaf(a4.x, w[k * n_out], acc0);
                acc1 = fmaf(a4.y, w[(k + 1) * n_out], acc1);
                acc2 = fmaf(a4.z, w[(k + 2) * n_out], acc2);
                acc3 = fmaf(a4.w, w[(k + 3) * n_out], acc3);
              }
              for (; k < n_in; ++k) acc0 = fmaf(a[k], w[k * n_out], acc0);
              float v = ((acc0 + acc1) + (acc2 + acc3)) + bias;
              if (ly.act) {
                // sigmoid and softplus from one exponential: e = exp(-|v|)
                const float e = expf(-fabsf(v)), r1 = 1.f / (1.f + e);
                ws[ly.sg_off + (long long)(b0 + gI) * ly.ld_out + c] = v >= 0.f ? r1 : e * r1;
                v = fmaxf(v, 0.f) + log1pf(e);
              }
              a_nxt[gI * 128 + c] = v;
              ws[ly.z_off + (long long)(b0 + gI) * ly.ld_out + c] = v;
            }
          }
        }
        asm volatile("bar.sync 1, %0;" ::"n"(CF_EPI) : "memory");      // layer done: activations visible, weight buffer free
        if (tid == 0 && l + 2 < mlp.L) fetch_w(l + 2);
        float* sw = a_cur; a_cur = a_nxt; a_nxt = sw;
      }
      // output loss normalized_mse (trainingUtilities.py:25-29): one warp per sample
      if (mlp_mode > 1) {
        const int d_out = mlp.layer[mlp.L - 1].n_out;
        for (int gI = warp_u; gI < nsamp; gI += CF_EPI_WARPS) {
          const long long sb = idx ? idx[b0 + gI] : (b0 + gI);
          const float* qt = q_tgt + sb * d_out;
          float e = 0.f;
          for (int cI = lane; cI < d_out; cI += 32) {
            const float d = a_cur[gI * 128 + cI] - __ldg(qt + cI);
            e = fmaf(d, d, e);
          }
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) e += __shfl_xor_sync(0xffffffffu, e, o);
          const float nq = __ldg(qn2 + sb);
          if (mlp_mode > 2) {
            const float gq = coef_q / nq;
            for (int cI = lane; cI < d_out; cI += 32)
              ws[mlp.dz_top_off + (long long)(b0 + gI) * mlp.ld_dz + cI] = gq * (a_cur[gI * 128 + cI] - __ldg(qt + cI));
          }
          if (lane == 0) {
            atomicAdd(stats + ST_L2_SUM, e / nq);
            atomicAdd(stats + ST_EQ_SUM, e);
            atomicAdd(stats + ST_NQ_SUM, nq);
            atomicAdd(stats + ST_COUNT, 1.f);
          }
        }
      }
      // the chain epilogue of OTHER threads reads these sigmoids from global memory; the weight-tile buffers and their
      // barriers now belong to the chain (the control warp waits for this before its first TMA)
      __threadfence_block();
      asm volatile("bar.sync 1, %0;" ::"n"(CF_EPI) : "memory");
      if (lane == 0) cf_mbar_arrive(mlp_done);
    }

    // X_0 = Q_{L-1} = seed (* sg_{L-1})
    {
      const ChainLayerDev& top = net.layer[L - 1];
      const int n16 = (top.n_out + 15) & ~15;
      for (int k = part * 4; k < n16; k += 4 * CF_PARTS) {
        float t4[4] = {0.f, 0.f, 0.f, 0.f};
        if (row_ok) {
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int kk = k + u;
            float x = 0.f;
            if (kk < top.n_out) {
              x = seed ? __ldg(seed + (long long)i_seed * top.n_out + kk) : (kk == i_seed ? 1.f : 0.f);
              if (top.act) x *= ws[top.sg_off + b * top.ld_out + kk];
            }
            t4[u] = x;
          }
          if (!tma_out && net.top_materialised && (do_backward || net.has_skip) && k < top.ld_out)
            *reinterpret_cast<float4*>(ws + top.q_off + grow * top.ld_out + k) = make_float4(t4[0], t4[1], t4[2], t4[3]);
        }
        cf_sts128(cf_smem_u32(X) + cf_xoff(r, k), t4[0], t4[1], t4[2], t4[3]);
        if (split) cf_tmem_st4(xlo_t + (uint32_t)k, tc::tf32_lo(t4[0]), tc::tf32_lo(t4[1]), tc::tf32_lo(t4[2]), tc::tf32_lo(t4[3]));
      }
      if (split) { cf_tmem_st_wait(); asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      if (lane == 0) cf_mbar_arrive(x_ready);
    }

    float st_h1 = 0.f, st_rr = 0.f, st_rows = 0.f;
#pragma unroll 1
    for (int t = 0; t < n_steps; ++t) {
      // ---- this step's descriptor: kernel parameter, uniform index -> uniform registers ----
      const ChainStep& sd = steps.st[t];
      const int N = sd.N;
      int mode = sd.mode;
      bool next_reads_x = sd.write_x != 0;          // the result is the next contraction's operand
      if (mode == 1) {
        if (!do_loss) mode = 0;                     // Jacobian only: plain row-major output
        if (!do_backward) next_reads_x = false;
      }
      const bool out_on = cf_out_needed(sd, net, do_loss, do_backward, dbg);
      // with TMA stores the global copy is taken from X, so X is written whenever the copy is needed
      const bool write_x = (next_reads_x || (tma_out && out_on)) && !(dbg & 2);
      const bool thread_out = out_on && !tma_out;
      const bool add_on = sd.add_off >= 0;
      const bool add_first = sd.add_first != 0;
      const bool sc_on = sd.sc_off >= 0 && !(dbg & 64);
      const bool do_round = (next_reads_x || do_backward) && !(dbg & 8) && !split;  // tf32: what feeds a later contraction is rounded to nearest; fp32x3 keeps fp32 and splits it
      float* const out_p = ws + sd.out_off + grow * sd.out_ld;
      const float* const add_p = ws + sd.add_off + grow * sd.add_ld;   // residual rows
      const float* const sc_p = ws + sd.sc_off + b * sd.sc_ld;      // this row's sample: per-sample scale vector
      const int sc_len = (sd.sc_ld < ((N + 3) & ~3)) ? sd.sc_ld : ((N + 3) & ~3);
      const float* prow = nullptr;                  // target-Jacobian row (loss)
      float gscale = 0.f, inv_nj = 0.f;
      if (mode == 1 && row_ok) {
        const long long sb = idx ? idx[b] : b;
        prow = jt + (sb * s + i_seed) * (long long)net.d_in;
        inv_nj = 1.f / __ldg(jn2 + sb);
        gscale = coef * inv_nj;
      }
      const bool vec_ok = prow != nullptr && ((reinterpret_cast<uintptr_t>(prow) & 15) == 0);
      const uint32_t acc = taddr + (uint32_t)((t & 1) * CF_ACC_COLS);

      // value of a 16-column block (identical in both phases): residual and per-sample scaling ...
      auto block_pre = [&](float (&x)[CF_CW], const float (&v)[CF_CW], const float (&sc)[CF_CW], int c0) {
#pragma unroll
        for (int j = 0; j < CF_CW; ++j) x[j] = v[j];
        if (add_on) {
          float ad[CF_CW];
          cf_load_vec16(ad, add_p, c0, sd.add_ld, row_ok);
          if (add_first) {
#pragma unroll
            for (int j = 0; j < CF_CW; ++j) x[j] = (x[j] + ad[j]) * sc[j];
          } else {
#pragma unroll
            for (int j = 0; j < CF_CW; ++j) x[j] = fmaf(x[j], sc[j], ad[j]);
          }
        } else if (sc_on) {
#pragma unroll
          for (int j = 0; j < CF_CW; ++j) x[j] *= sc[j];
        }
      };
      // ... then the loss gradient (last forward step) and the TF32 rounding of whatever feeds a later contraction
      auto block_post = [&](float (&x)[CF_CW], const float (&tg)[CF_CW]) {
        if (mode == 1) {
#pragma unroll
          for (int j = 0; j < CF_CW; ++j) x[j] = gscale * (x[j] - tg[j]);
        }
        if (do_round) {
#pragma unroll
          for (int j = 0; j < CF_CW; ++j) x[j] = cf_round_tf32(x[j]);
        }
      };
      // per-block operands: scale vector block, target-Jacobian block
      auto load_sc_tg = [&](float (&sc)[CF_CW], float (&tg)[CF_CW], int c0) {
        if (sc_on) cf_load_vec16(sc, sc_p, c0, sc_len);
        if (mode == 1) cf_load16(tg, prow, c0, N, row_ok, vec_ok);
      };

      // ---- while the MMA runs: request the first block's operands ----
      float sc0[CF_CW], tg0[CF_CW];
#pragma unroll
      for (int j = 0; j < CF_CW; ++j) { sc0[j] = 1.f; tg0[j] = 0.f; }
      if (c_first < N) {
        load_sc_tg(sc0, tg0, c_first);
      }
      cf_mbar_wait(mma_bar(t), (t >> 1) & 1);
      if (tid == 0) CF_STAMP(t, 4);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

      // ---- phase A: accumulator -> next operand X (and its row-major copy in global memory), hand X over ----
      // Blocks go in pairs: both values are computed first, then X is written.  X still feeds the TMA store of the
      // PREVIOUS result when the contraction completes; computing before waiting for that store hides its latency.
      float e_row = 0.f, n_row = 0.f;
      if (write_x || thread_out || mode == 1) {
        auto compute_block = [&](float (&x)[CF_CW], int c0) {
          float v[CF_CW], sc[CF_CW], tg[CF_CW];
          cf_tmem_ld16(acc + (uint32_t)c0, v);
#pragma unroll
          for (int j = 0; j < CF_CW; ++j) { sc[j] = sc0[j]; tg[j] = tg0[j]; }
          if (c0 != c_first) load_sc_tg(sc, tg, c0);
          block_pre(x, v, sc, c0);
          if (mode == 1) {                           // loss statistics of the row: squared error and target norm
#pragma unroll
            for (int j = 0; j < CF_CW; ++j) {
              const float d = x[j] - tg[j];
              e_row = fmaf(d, d, e_row);
              n_row = fmaf(tg[j], tg[j], n_row);
            }
          }
          block_post(x, tg);
        };
        auto write_block = [&](const float (&x)[CF_CW], int c0) {
          if (write_x) {
            const uint32_t xb = xrow + (uint32_t)((c0 >> 5) * CF_CHUNK_BYTES);
            const int j0 = (c0 & 31) >> 2;
#pragma unroll
            for (int j = 0; j < CF_CW; j += 4)
              cf_sts128(xb + (uint32_t)(((j0 + (j >> 2)) ^ xsw) << 4), x[j], x[j + 1], x[j + 2], x[j + 3]);
            if (split) {
              float lo[CF_CW];
#pragma unroll
              for (int j = 0; j < CF_CW; ++j) lo[j] = tc::tf32_lo(x[j]);
              cf_tmem_st16(xlo_t + (uint32_t)c0, lo);
            }
          }
          if (thread_out && row_ok) cf_store_vec16(out_p, c0, sd.out_ld, x);     // padded columns are exact zeros
        };
        bool x_is_free = !tma_out;
#pragma unroll 1
        for (int c0 = c_first; c0 < N; c0 += 2 * STRIDE) {
          float xa[CF_CW], xb[CF_CW];
          const bool two = c0 + STRIDE < N;
          compute_block(xa, c0);
          if (two) compute_block(xb, c0 + STRIDE);
          if (!x_is_free) { cf_mbar_wait(x_free, t & 1); x_is_free = true; }   // the copy of the previous result has left X
          write_block(xa, c0);
          if (two) write_block(xb, c0 + STRIDE);
        }
        if (!x_is_free) cf_mbar_wait(x_free, t & 1);          // (a thread without blocks still follows every phase)
      } else if (tma_out) {
        cf_mbar_wait(x_free, t & 1);
      }
      if (split) cf_tmem_st_wait();
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      if (t + 1 < n_steps || (tma_out && out_on)) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) cf_mbar_arrive(x_ready);
      }
      if (tid == 0) CF_STAMP(t, 5);

      // ---- cross-warp leftovers of the step ----
      if (mode == 1) {
        {
          if (row_ok) st_h1 += e_row * inv_nj;
          ex_s[(part * 128 + r) * 2 + 0] = e_row;               // the per-row metric needs the FULL row sums
          ex_s[(part * 128 + r) * 2 + 1] = n_row;
        }
        asm volatile("bar.sync 1, %0;" ::"n"(CF_EPI) : "memory");
        if (mode == 1 && part == 0 && row_ok) {
          float e_all = 0.f, n_all = 0.f;
#pragma unroll
          for (int pp = 0; pp < CF_PARTS; ++pp) { e_all += ex_s[(pp * 128 + r) * 2 + 0]; n_all += ex_s[(pp * 128 + r) * 2 + 1]; }
          st_rr += e_all / n_all;
          st_rows += 1.f;
        }
      }
      if (tid == 0) CF_STAMP(t, 6);
    }
    if (tid == 0) CF_STAMP(2, 7);
    if (do_loss) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        st_h1 += __shfl_xor_sync(0xffffffffu, st_h1, o);
        st_rr += __shfl_xor_sync(0xffffffffu, st_rr, o);
        st_rows += __shfl_xor_sync(0xffffffffu, st_rows, o);
      }
      if (lane == 0) {                 // every column part of a row carries part of its squared error
        if (st_h1 != 0.f) atomicAdd(stats + ST_H1_SUM, st_h1);
        if (st_rows > 0.f) { atomicAdd(stats + ST_JROW_SUM, st_rr); atomicAdd(stats + ST_JROWS, st_rows); }
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (tid == 0) CF_STAMP(3, 7);
  if (warp_u == CF_EPI_WARPS)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(split ? 512u : 2u * CF_ACC_COLS) : "memory");
}

// ---- host ---------------------------------------------------------------------------------------------
typedef CUresult (*PFN_cfEncode)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                 const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

bool chain_fused_encode_maps(ChainMaps* maps, const ChainNetDev& net, const ChainSteps& steps, long long max_rows, const float* ws,
                             const long long* wp_off, const long long* ld_wp, const long long* wtp_off, const long long* ld_wtp,
                             const long long* wp_lo_off, const long long* wtp_lo_off) {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) != cudaSuccess ||
      qres != cudaDriverEntryPointSuccess)
    return false;
  PFN_cfEncode enc = reinterpret_cast<PFN_cfEncode>(fn);
  auto mk = [&](CUtensorMap* m, const float* base, int rows, int cols, long long ld) {
    cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t gstr[1] = {(cuuint64_t)ld * sizeof(float)};
    cuuint32_t box[2] = {32, (cuuint32_t)CF_ROWS};
    cuuint32_t es[2] = {1, 1};
    return enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), gdim, gstr, box, es,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
  };
  // row-major destinations of the step results [max_rows, pitch]: box = 32 columns x the rows of one sample (or of the
  // whole tile when the samples are packed back to back); rows / columns beyond the array are clipped by the hardware
  const cuuint32_t box_rows = (cuuint32_t)(net.srow == net.s ? (net.G * net.s < CF_ROWS ? net.G * net.s : CF_ROWS) : net.s);
  for (int t = 0; t < steps.n_all; ++t) {
    const ChainStep& st = steps.st[t];
    if (st.out_off < 0) continue;
    cuuint64_t gdim[2] = {(cuuint64_t)st.out_ld, (cuuint64_t)max_rows};
    cuuint64_t gstr[1] = {(cuuint64_t)st.out_ld * sizeof(float)};
    cuuint32_t box[2] = {32, box_rows};
    cuuint32_t es[2] = {1, 1};
    if (enc(&maps->out[t], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(ws + st.out_off), gdim, gstr, box, es,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return false;
  }
  if (net.top_materialised && net.layer[net.L - 1].q_off >= 0) {
    const ChainLayerDev& top = net.layer[net.L - 1];
    cuuint64_t gdim[2] = {(cuuint64_t)top.ld_out, (cuuint64_t)max_rows};
    cuuint64_t gstr[1] = {(cuuint64_t)top.ld_out * sizeof(float)};
    cuuint32_t box[2] = {32, box_rows};
    cuuint32_t es[2] = {1, 1};
    if (enc(&maps->top, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(ws + top.q_off), gdim, gstr, box, es,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return false;
  }
  for (int l = 0; l < net.L; ++l) {
    const ChainLayerDev& ly = net.layer[l];
    // forward: B operand stored [N = n_in][K = n_out] = padded W_l ; backward: [N = n_out][K = n_in] = padded W_l^T
    if (!mk(&maps->fwd[l], ws + wp_off[l], ly.n_in, ly.n_out, ld_wp[l])) return false;
    if (!mk(&maps->bwd[l], ws + wtp_off[l], ly.n_out, ly.n_in, ld_wtp[l])) return false;
    if (wp_lo_off && wp_lo_off[l] >= 0) {     // fp32x3: the low parts of the same tiles
      if (!mk(&maps->fwd_lo[l], ws + wp_lo_off[l], ly.n_in, ly.n_out, ld_wp[l])) return false;
      if (!mk(&maps->bwd_lo[l], ws + wtp_lo_off[l], ly.n_out, ly.n_in, ld_wtp[l])) return false;
    }
  }
  return true;
}

// Execution-order step table: forward l = L-1..0, then backward l = 0..(L-1 or L-2).
void chain_fused_build_steps(ChainSteps* steps, const ChainNetDev& net) {
  const int L = net.L;
  auto blank = [] { ChainStep st; st.K = st.N = 0; st.mode = 0; st.add_first = 0; st.write_x = 1;
                    st.sc_ld = st.out_ld = st.add_ld = st.q_ld = st.rs_ld = 0;
                    st.sc_off = st.out_off = st.add_off = st.q_off = st.rs_off = -1; return st; };
  int t = 0;
  for (int l = L - 1; l >= 0; --l, ++t) {
    const ChainLayerDev& ly = net.layer[l];
    ChainStep st = blank();
    st.K = ly.n_out; st.N = ly.n_in;
    if (l > 0) {
      const ChainLayerDev& below = net.layer[l - 1];
      if (below.act) { st.sc_off = below.sg_off; st.sc_ld = below.ld_out; }
      st.out_off = below.q_off; st.out_ld = below.ld_out;
    } else {
      st.mode = 1;
      st.out_off = net.r0_off; st.out_ld = net.ld_in;
    }
    if (ly.lands_skip >= 0) { st.add_off = net.layer[ly.lands_skip].q_off; st.add_ld = net.layer[ly.lands_skip].ld_out; st.add_first = 1; }   // Q of the skip layer
    steps->st[t] = st;
  }
  steps->n_fwd = t;
  // the backward contraction of a LINEAR top layer produces nothing anybody reads (R_L is the constant seed)
  const int n_bwd = net.layer[L - 1].act ? L : L - 1;
  for (int l = 0; l < n_bwd; ++l, ++t) {
    const ChainLayerDev& ly = net.layer[l];
    ChainStep st = blank();
    st.K = ly.n_in; st.N = ly.n_out;
    if (ly.act) {
      st.mode = 2;
      st.sc_off = ly.sg_off; st.sc_ld = ly.ld_out;
      st.q_off = ly.q_off; st.q_ld = ly.ld_out;                 // Q_l
      st.rs_off = ly.rs_off; st.rs_ld = ly.ld_out;
    }
    if (l < L - 1) {
      st.out_off = ly.dr_off; st.out_ld = ly.ld_out;
      if (ly.skip_from >= 0) {
        const int j = ly.skip_from;
        st.add_off = j == 0 ? net.r0_off : net.layer[j - 1].dr_off;        // dR_j
        st.add_ld = j == 0 ? net.ld_in : net.layer[j - 1].ld_out;
      }
    } else {
      st.write_x = 0;
    }
    steps->st[t] = st;
  }
  steps->n_all = t;
}

bool chain_fused_supported(const ChainNetDev& net) {
  if (net.L < 1 || net.L > CHAIN_FUSED_MAX_LAYERS) return false;
  if (net.s < 1 || net.s > CF_ROWS || net.G < 1 || net.G > 16) return false;
  if (net.d_in > CF_MAXW) return false;
  for (int l = 0; l < net.L; ++l)
    if (net.layer[l].n_in > CF_MAXW || net.layer[l].n_out > CF_MAXW) return false;
  return true;
}

cudaError_t chain_fused_read_trace(unsigned long long* out) {
  return cudaMemcpyFromSymbol(out, g_cf_trace, sizeof(unsigned long long) * 64 * 8);
}

cudaError_t launch_chain_fused(const ChainMaps& maps, const ChainNetDev& net, const ChainSteps& steps, float* ws, const float* jt, const float* jn2,
                               const int* idx, const float* seed, int B, float coef, int do_loss, int do_backward,
                               float* stats, const MlpNetDev& mlp, const float* params, const float* m_in, const float* q_tgt,
                               const float* qn2, float coef_q, int mlp_mode, int split, cudaStream_t st) {
  constexpr int SMEM = 2 * 16 * 128 * 4 + 3 * CF_X_BYTES + (16 + 2 * CF_SLOTS) * 8 + CF_PARTS * 128 * 2 * 4 + CF_EPI_WARPS * 8 * CF_TILE_LD * 4 + 1024;
  static_assert(SMEM <= 232448, "shared memory budget of one CTA");
  static std::atomic<unsigned long long> attr_mask{0};
  cudaError_t e = tc_set_smem_once(reinterpret_cast<const void*>(chain_fused_kernel), SMEM, attr_mask);
  if (e != cudaSuccess) return e;
  const int ctas = (B + net.G - 1) / net.G;
  chain_fused_kernel<<<ctas, CF_THREADS, SMEM, st>>>(maps, net, steps, ws, jt, jn2, idx, seed, B, coef, do_loss, do_backward, stats,
                                                      mlp, params, m_in, q_tgt, qn2, coef_q, mlp_mode, split);
  return cudaGetLastError();
}

}  // namespace dino
