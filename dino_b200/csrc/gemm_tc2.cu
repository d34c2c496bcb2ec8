// tcgen05 engine, cta_group::2 form: C[M,N] = epilogue(A[M,K] * Bt[N,K]^T) on PAIRS of CTAs (a cluster of two = the two
// SMs of a TPC).  A pair owns a 256 x 256 output tile: each CTA stages ITS 128 rows of A and ITS 128-row half of the B
// tile, the leader CTA issues tcgen05.mma.cta_group::2 (M = 256), the hardware shares the B halves between the two
// tensor cores, and each CTA's tensor memory holds its own 128 x 256 half of the accumulator.  Per SM that halves the
// shared-memory operand reads of a 128 x 256 tile (64 instead of 96 bytes per clock in tf32) and the TMA traffic for B -
// the two things the single-CTA kernel (gemm_tc.cu) runs out of: fp32x3 is bound by shared-memory bandwidth there.
//
// Barriers (same offsets in both CTAs):
//   tf32   both producers credit the LEADER's full[stage] (cp.async.bulk.tensor.cta_group::2; no multicast, the maps
//          cover this GPU's memory); the leader's MMA thread waits on it
//   fp32x3 each producer uses its own full[stage]; the converter warps of BOTH CTAs then arrive on the leader's
//          conv[stage] (the peer's with a cluster-scope remote arrive) once the lo tiles are written
//   empty[stage], tmem_full[acc]: tcgen05.commit.cta_group::2 with multicast: both CTAs see them
//   tmem_empty[acc] (leader's): the epilogue warps of both CTAs arrive (2 x NEPI)
// Everything else (epilogue classes, staging, hi/lo split, the separate cross-term accumulator) is gemm_tc.cu's.
#include <cuda.h>

#include <cstdlib>

#include "tc_epilogue.cuh"

namespace dino {

using namespace tc;

constexpr int T2_BK = 32;

// NCONV: converter warps of the fp32x3 form (a stage of 32 KB per 1536 MMA clocks: four keep up)
template <bool SPLIT, int EPI, int NEPI, int STAGES, int NCONV, int BN2>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__((2 + NEPI + (SPLIT ? NCONV : 0)) * 32, 1)
gemm_tc2_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b, const TcEpilogue ep,
                const __grid_constant__ CUtensorMap tma_b_lo) {
  // BN2: columns of the pair's tile (256, or 128: half the operand reuse, but main + cross accumulators of fp32x3 then
  // fit tensor memory twice and the epilogue of a tile overlaps the next tile's main loop)
  constexpr int A_BYTES = 128 * T2_BK * 4, B_BYTES = (BN2 / 2) * T2_BK * 4, RAW_BYTES = A_BYTES + B_BYTES;
  constexpr int STAGE_BYTES = SPLIT ? 2 * RAW_BYTES : RAW_BYTES;
  constexpr uint32_t ACC_COLS = BN2;
  // fp32x3: main sum + cross-term accumulator (as in gemm_tc.cu) fill tensor memory: one set, the epilogue does not
  // overlap the next tile.  (Tried: one accumulator and two passes over K - cross terms first, then the main products -
  // which frees the second set; the extra operand pass cost more than the overlap gave: 384 us against 240 us at
  // [25600,1024] x [1024,1024].  The feed - TMA, conversion, hand-over - not the tensor pipe, paces this kernel.)
  constexpr uint32_t SET_COLS = SPLIT ? 2 * ACC_COLS : ACC_COLS;
  constexpr int NACC = 2 * SET_COLS <= 512 ? 2 : 1;
  constexpr int EPI_BUF = 32 * 128;
  constexpr int CONV0 = 2 + NEPI;

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* epi_stage = smem + STAGES * STAGE_BYTES;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(epi_stage + NEPI * EPI_BUF);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* conv_bar = empty_bar + STAGES;
  uint64_t* tmem_full_bar = conv_bar + STAGES;
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();                   // 0 = leader
  const int pair = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
  const int nkb = (ep.K + T2_BK - 1) / T2_BK;
  const int tiles_n = (ep.N + BN2 - 1) / BN2, n_tiles = tiles_n * ((ep.M + 255) / 256);

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_b) : "memory");
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < STAGES; ++i) { mbar_init(full_bar + i, 1); mbar_init(empty_bar + i, 1); mbar_init(conv_bar + i, NCONV > 0 ? 2 * NCONV : 1); }
      mbar_init(tmem_full_bar, 1); mbar_init(tmem_full_bar + 1, 1);
      mbar_init(tmem_empty_bar, 2 * NEPI); mbar_init(tmem_empty_bar + 1, 2 * NEPI);
      fence_mbar_init();
    }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync();                                            // both CTAs' barriers exist before anybody signals across
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===== TMA producer (both CTAs): own 128 rows of A, own 128-row half of the B tile =====
    if (lane == 0) {
      int it = 0;
      for (int tile = pair; tile < n_tiles; tile += n_pairs) {
        const int m0 = (tile / tiles_n) * 256 + (int)rank * 128, n0 = (tile % tiles_n) * BN2 + (int)rank * (BN2 / 2);
        for (int kb = 0; kb < nkb; ++kb, ++it) {
          const int st = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1;
          mbar_wait(empty_bar + st, ph ^ 1);
          uint8_t* sa = smem + st * STAGE_BYTES;
          if (SPLIT) {
            mbar_expect_tx(full_bar + st, ep.b_lo ? RAW_BYTES + B_BYTES : RAW_BYTES);
            tma_load_2d(sa, &tma_a, kb * T2_BK, m0, full_bar + st);
            tma_load_2d(sa + A_BYTES, &tma_b, kb * T2_BK, n0, full_bar + st);
            if (ep.b_lo) tma_load_2d(sa + RAW_BYTES + A_BYTES, &tma_b_lo, kb * T2_BK, n0, full_bar + st);
          } else {
            if (rank == 0) mbar_expect_tx(full_bar + st, 2 * RAW_BYTES);     // the bytes of both CTAs land on the leader's barrier
            tma_load_2d_2sm(sa, &tma_a, kb * T2_BK, m0, full_bar + st);
            tma_load_2d_2sm(sa + A_BYTES, &tma_b, kb * T2_BK, n0, full_bar + st);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: the leader CTA's elected thread, for the pair =====
    if (rank == 0 && lane == 0) {
      const uint32_t idesc = idesc_tf32_m(256, BN2);
      int it = 0, i = 0;
      for (int tile = pair; tile < n_tiles; tile += n_pairs, ++i) {
        const int acc = i % NACC;
        mbar_wait(tmem_empty_bar + acc, ((i / NACC) & 1) ^ 1);    // both CTAs' epilogues have drained this accumulator set
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)acc * SET_COLS;
        for (int kb = 0; kb < nkb; ++kb, ++it) {
          const int st = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1;
          mbar_wait((SPLIT ? conv_bar : full_bar) + st, ph);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + st * STAGE_BYTES);
          const uint64_t adesc = desc_k_sw128(sa), bdesc = desc_k_sw128(sa + A_BYTES);
          constexpr uint64_t LO = RAW_BYTES >> 4;
#pragma unroll
          for (int k = 0; k < T2_BK / 8; ++k) {
            umma_tf32_2sm(tmem_d, adesc + 2 * k, bdesc + 2 * k, idesc, (kb > 0 || k > 0) ? 1u : 0u);
            if (SPLIT) {
              umma_tf32_2sm(tmem_d + ACC_COLS, adesc + LO + 2 * k, bdesc + 2 * k, idesc, (kb > 0 || k > 0) ? 1u : 0u);
              umma_tf32_2sm(tmem_d + ACC_COLS, adesc + 2 * k, bdesc + LO + 2 * k, idesc, 1u);
            }
          }
          umma_commit_2sm(empty_bar + st);             // frees the stage in BOTH CTAs
        }
        umma_commit_2sm(tmem_full_bar + acc);          // accumulator halves complete in both CTAs
      }
    }
  } else if (SPLIT && warp >= CONV0) {
    // ===== converters (fp32x3, both CTAs): lo tile of every landed stage, reported to the leader =====
    const int t = threadIdx.x - CONV0 * 32;
    int it = 0;
    for (int tile = pair; tile < n_tiles; tile += n_pairs) {
      for (int kb = 0; kb < nkb; ++kb, ++it) {
        const int st = it % STAGES;
        mbar_wait(full_bar + st, (it / STAGES) & 1);
        const uint32_t sa = smem_u32(smem + st * STAGE_BYTES);
        split_lo_tile(sa, sa + RAW_BYTES, ep.b_lo ? A_BYTES : RAW_BYTES, t, NCONV * 32);
        fence_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(conv_bar + st, 0);
      }
    }
  } else {
    // ===== epilogue (both CTAs): own 128 x 256 half of the accumulator =====
    const int q = warp & 3, ew = warp - 2, half = ew >> 2;
    const uint32_t my_stage = smem_u32(epi_stage + ew * EPI_BUF);
    int i = 0;
#pragma unroll 1
    for (int tile = pair; tile < n_tiles; tile += n_pairs, ++i) {
      const int m0 = (tile / tiles_n) * 256 + (int)rank * 128, n0 = (tile % tiles_n) * BN2;
      const int acc = i % NACC;
      mbar_wait(tmem_full_bar + acc, (i / NACC) & 1);
      tc_fence_after();
      tc_epilogue_tile<BN2, SPLIT, EPI, NEPI, ACC_COLS>(ep, m0, n0, tmem_base + (uint32_t)acc * SET_COLS, my_stage, q, half, lane, ep.M);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(tmem_empty_bar + acc, 0);
    }
  }
  // ===== teardown: nobody leaves (or frees tensor memory) while the pair still works =====
  tc_fence_before();
  __syncthreads();
  cluster_sync();
  if (warp == 1)
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u) : "memory");
}

template <bool SPLIT, int EPI, int NEPI, int STAGES, int NCONV, int BN2 = 256>
static cudaError_t launch_tc2(const GemmOp& op, cudaStream_t st) {
  CUtensorMap ma, mb, mbl;
  if (!tc_make_map(&ma, op.A, op.K, op.M, op.a_sm, T2_BK, 128, (int)CU_TENSOR_MAP_SWIZZLE_128B) ||
      !tc_make_map(&mb, op.B, op.K, op.N, op.b_sn, T2_BK, BN2 / 2, (int)CU_TENSOR_MAP_SWIZZLE_128B))
    return cudaErrorInvalidValue;
  const bool b_lo = SPLIT && op.B_lo != nullptr && (reinterpret_cast<uintptr_t>(op.B_lo) & 15) == 0;
  if (b_lo) {
    if (!tc_make_map(&mbl, op.B_lo, op.K, op.N, op.b_sn, T2_BK, BN2 / 2, (int)CU_TENSOR_MAP_SWIZZLE_128B)) return cudaErrorInvalidValue;
  } else mbl = mb;
  TcEpilogue ep;
  ep.M = op.M; ep.N = op.N; ep.K = op.K; ep.C = op.C; ep.ldc = op.ldc; ep.sdim = op.sdim > 0 ? op.sdim : 1;
  ep.scale = op.scale; ep.ld_sc = op.ld_sc; ep.add_pre = op.add_pre; ep.ld_ap = op.ld_ap;
  ep.add_post = op.add_post; ep.ld_apo = op.ld_apo; ep.bias = op.bias; ep.act = op.act;
  ep.s_out = op.act ? op.s_out : nullptr; ep.ld_s = op.ld_s;
  ep.yref = op.rowsum ? op.yref : nullptr; ep.ld_y = op.ld_y; ep.rowsum = op.rowsum; ep.ld_rs = op.ld_rs;
  ep.round_tf32 = SPLIT ? 0 : op.round_tf32;
  ep.a_rows = 0; ep.a_tps = 0; ep.a_idx = nullptr; ep.rowdot = nullptr; ep.pre_first = 0; ep.a_gran = 0; ep.b_lo = b_lo ? 1 : 0;
  constexpr int RAW = (128 + BN2 / 2) * T2_BK * 4;
  constexpr int SMEM = STAGES * (SPLIT ? 2 * RAW : RAW) + NEPI * 32 * 128 + 256 + 1024;
  static_assert(SMEM <= 232448, "shared memory budget of one CTA");
  static std::atomic<unsigned long long> attr_mask{0};
  cudaError_t e = tc_set_smem_once(reinterpret_cast<const void*>(gemm_tc2_kernel<SPLIT, EPI, NEPI, STAGES, NCONV, BN2>), SMEM, attr_mask);
  if (e != cudaSuccess) return e;
  const int num_sms = tc_current_num_sms();
  const long long n_tiles = (long long)((op.N + BN2 - 1) / BN2) * ((op.M + 255) / 256);
  const int pairs = (int)(n_tiles < num_sms / 2 ? n_tiles : num_sms / 2);
  gemm_tc2_kernel<SPLIT, EPI, NEPI, STAGES, NCONV, BN2><<<2 * pairs, (2 + NEPI + (SPLIT ? NCONV : 0)) * 32, SMEM, st>>>(ma, mb, ep, mbl);
  return cudaGetLastError();
}

// the pair kernel pays off for contractions with many 256 x 256 tiles and a long reduction
bool gemm_tc2_wanted(const GemmOp& op) {
  return op.N >= 256 && op.K >= 256 && (long long)((op.N + 255) / 256) * ((op.M + 255) / 256) >= 2LL * (tc_current_num_sms() / 2);
}

cudaError_t launch_gemm_tc2(const GemmOp& op, int split, cudaStream_t st) {
  const bool net = op.bias || op.act;
  const bool chain = op.rowsum || op.scale || op.add_pre || op.add_post || (op.round_tf32 && !split);
  if (split) {
    // the accumulator pair (main + cross terms) fills tensor memory, so the epilogue does not overlap the next tile's
    // main loop here: eight epilogue warps keep it short; four converter warps keep up at this tile size.
    // (Tried: the two cross terms as ONE bf16 contraction over [a | a_lo] x [b_lo | b] - a third fewer tensor cycles, the
    //  same accuracy - which came out slower, 274 against 241 us: the conversion got heavier and the kernel is not
    //  bound by the tensor pipe but by the clocks it can hold under the power cap.)
    // (eight converter warps instead of four: 576 threads leave 96 registers each, the chain epilogue spills, 3 % slower)
    // (Measured: launch_tc2<true, 1, 8, 4, 4, 128> - 256 x 128 tiles, whose two accumulator sets let the epilogue overlap the
    //  next tile, four 48 KB stages - runs the cfg5 chain stage in 3.6 ms against 2.9 ms: at half the operand reuse the
    //  main loop loses more than the exposed epilogue costs.  Not instantiated.)
    if (net) return launch_tc2<true, 2, 8, 3, 4>(op, st);
    if (chain) return launch_tc2<true, 1, 8, 3, 4>(op, st);
    return launch_tc2<true, 0, 8, 3, 4>(op, st);
  }
  if (net) return launch_tc2<false, 2, 8, 6, 0>(op, st);
  if (chain) return launch_tc2<false, 1, 8, 6, 0>(op, st);
  return launch_tc2<false, 0, 8, 6, 0>(op, st);
}

}  // namespace dino
