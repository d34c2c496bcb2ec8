// Weight gradients on the tensor cores:
//     dW_l[n_in, n_out] += dR_l^T Q_l = X[rows, n_in]^T * Y[rows, n_out]          (rows = B*s, tens of thousands)
// for every layer in ONE launch, split over the rows.  X / Y are the chain states exactly as the chain kernel keeps
// them: ROW-major.  The reduction index is then the strided one, i.e. both factors are MN-major UMMA operands; for
// 4-byte elements that takes the 128-byte swizzle with 32-byte atoms on the TMA side (CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B)
// and shared-memory descriptor layout type 1 (SWIZZLE_128B_BASE32B: LBO = pitch between 32-column groups, SBO = 512 B
// = 4 reduction rows) plus the two "major" bits of the instruction descriptor.  Transposed factors [n, rows] (K-major,
// plain SWIZZLE_128B) are still accepted.  Pipeline: TMA 3-stage ring -> tcgen05.mma.kind::tf32 -> TMEM ->
// red.global.add.  Output tiles of 128 x 128; wide layers are covered by several tiles (blockIdx.x enumerates op,
// tile, split).  The launch is sized to ONE wave of CTAs: every extra split adds 128 x n_out atomics.
// fp32x3 mode: four converter warps write the lo tile (x - trunc_tf32(x)) of every landed stage beside it and each
// 8-row reduction step issues three MMAs (tc_common.cuh), which makes the reduction fp32-grade.
#include <cuda.h>

#include <cstdlib>
#include <mutex>

#include "tc_common.cuh"

namespace dino {

constexpr int WG_BK = 32;                       // reduction rows per stage (one 128-byte swizzle row)
constexpr int WG_STAGES = 3;   // tf32: 96 KB, leaves room for a fused-MLP CTA on the same SM (the two overlap on separate streams); fp32x3: 192 KB
constexpr int WG_OPER_BYTES = 128 * WG_BK * 4;  // 16 KB: up to 128 operand rows x 32 reduction elements
constexpr int WG_RAW_BYTES = 2 * WG_OPER_BYTES;
constexpr int WG_X3_CHUNKS = 32;                // fp32x3: longest reduction (in chunks of WG_BK rows) one TMEM accumulator takes

struct WgOp {
  float* C;                 // gradient block [n_in, n_out], ldc = n_out
  int n_in, n_out, rows;    // rows = reduction length
  int z_start;              // first blockIdx.x of this op
  int chunks_per_cta;       // reduction chunks (of WG_BK rows) per CTA
  int n_tiles, splits;      // CTAs of this op: m_tiles * n_tiles * splits, split fastest
  int mn;                   // 1: operands are the ROW-major chain states [rows, n] (MN-major UMMA operands), 0: transposed copies
};
struct WgGroup {
  int n;
  WgOp op[GEMM_GROUP_MAX];
  CUtensorMap mx[GEMM_GROUP_MAX];   // XT [n_in, rows]
  CUtensorMap my[GEMM_GROUP_MAX];   // YT [n_out, rows]
};

__device__ __forceinline__ uint32_t wg_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void wg_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "WG_WAIT:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra WG_DONE;\n\t"
      "bra WG_WAIT;\n\t"
      "WG_DONE:\n\t"
      "}" ::"r"(wg_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ uint64_t wg_desc(uint32_t saddr) {    // K-major, SWIZZLE_128B (as gemm_tc.cu)
  return (uint64_t)((saddr & 0x3FFFF) >> 4) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}

// MN-major TF32 operand tile as TMA (SWIZZLE_128B_ATOM_32B) lays it down: per 32-column group WG_BK reduction rows of
// 128 bytes; canonical UMMA layout ((8,n),(4,k)):((1,LBO),(8,SBO)) in 16-byte units: LBO = group pitch, SBO = 4-row atom
__device__ __forceinline__ uint64_t wg_desc_mn(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3FFFF) >> 4) | ((uint64_t)((WG_BK * 128) >> 4) << 16) | ((uint64_t)(512 >> 4) << 32) | (1ull << 46) | (1ull << 61);
}

template <bool SPLIT>
__global__ void __launch_bounds__(SPLIT ? 320 : 192, 1) wgrad_tc_kernel(const __grid_constant__ WgGroup g) {
  constexpr int WG_STAGE_BYTES = SPLIT ? 2 * WG_RAW_BYTES : WG_RAW_BYTES;
  extern __shared__ uint8_t wg_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(wg_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + WG_STAGES * WG_STAGE_BYTES);
  uint64_t* empty_bar = full_bar + WG_STAGES;
  uint64_t* done_bar = empty_bar + WG_STAGES;
  uint64_t* conv_bar = done_bar + 1;                      // [WG_STAGES] lo tiles written (fp32x3)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(conv_bar + WG_STAGES);

  int oi = 0;
  while (oi + 1 < g.n && (int)blockIdx.x >= g.op[oi + 1].z_start) ++oi;
  const WgOp& op = g.op[oi];
  const int total_chunks = (op.rows + WG_BK - 1) / WG_BK;
  const int local = blockIdx.x - op.z_start;
  const int split = local % op.splits, tile = local / op.splits;
  const int m0 = (tile / op.n_tiles) * 128, n0 = (tile % op.n_tiles) * 128;
  const int c_begin = split * op.chunks_per_cta;
  const int nchunks = min(total_chunks - c_begin, op.chunks_per_cta);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < WG_STAGES; ++i) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(wg_u32(full_bar + i)), "r"(1u));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(wg_u32(empty_bar + i)), "r"(1u));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(wg_u32(conv_bar + i)), "r"(4u));
      }
      asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(wg_u32(done_bar)), "r"(1u));
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(wg_u32(tmem_slot)), "r"(SPLIT ? 256u : 128u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;

  if (nchunks > 0) {
    if (warp == 0) {
      if (lane == 0) {
        const CUtensorMap* mx = &g.mx[oi];
        const CUtensorMap* my = &g.my[oi];
        for (int c = 0; c < nchunks; ++c) {
          const int st = c % WG_STAGES;
          wg_wait(empty_bar + st, ((c / WG_STAGES) & 1) ^ 1);
          uint8_t* sx = smem + st * WG_STAGE_BYTES;
          const int k0 = (c_begin + c) * WG_BK;
          if (op.mn) {
            // row-major states: one box of 32 columns x WG_BK rows per column group that holds data
            const int ga = (min(128, op.n_in - m0) + 31) >> 5, gb = (min(128, op.n_out - n0) + 31) >> 5;
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(wg_u32(full_bar + st)),
                         "r"((uint32_t)((ga + gb) * WG_BK * 128)) : "memory");
            for (int gI = 0; gI < ga; ++gI)
              asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                           ::"r"(wg_u32(sx + gI * WG_BK * 128)), "l"(mx), "r"(wg_u32(full_bar + st)), "r"(m0 + 32 * gI), "r"(k0) : "memory");
            for (int gI = 0; gI < gb; ++gI)
              asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                           ::"r"(wg_u32(sx + WG_OPER_BYTES + gI * WG_BK * 128)), "l"(my), "r"(wg_u32(full_bar + st)), "r"(n0 + 32 * gI), "r"(k0) : "memory");
            continue;
          }
          asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(wg_u32(full_bar + st)),
                       "r"((uint32_t)WG_RAW_BYTES) : "memory");
          asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                       ::"r"(wg_u32(sx)), "l"(mx), "r"(wg_u32(full_bar + st)), "r"(k0), "r"(m0) : "memory");
          asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                       ::"r"(wg_u32(sx + WG_OPER_BYTES)), "l"(my), "r"(wg_u32(full_bar + st)), "r"(k0), "r"(n0) : "memory");
        }
      }
    } else if (warp == 1) {
      if (lane == 0) {
        const uint32_t n16 = (uint32_t)((min(128, op.n_out - n0) + 15) & ~15);
        // (bits 15 / 16: A / B are MN-major when the operands are the row-major states)
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((n16 >> 3) << 17) | ((128u >> 4) << 24) |
                               (op.mn ? ((1u << 15) | (1u << 16)) : 0u);
        const uint64_t kstep = op.mn ? 64 : 2;      // descriptor start-address advance per 8 reduction rows (16-byte units)
        for (int c = 0; c < nchunks; ++c) {
          const int st = c % WG_STAGES;
          wg_wait((SPLIT ? conv_bar : full_bar) + st, (c / WG_STAGES) & 1);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint32_t sx = wg_u32(smem + st * WG_STAGE_BYTES), sy = sx + WG_OPER_BYTES;
          const uint64_t ad = op.mn ? wg_desc_mn(sx) : wg_desc(sx), bd = op.mn ? wg_desc_mn(sy) : wg_desc(sy);
          constexpr uint64_t LO = WG_RAW_BYTES >> 4;        // the lo tiles sit WG_RAW_BYTES above their sources
#pragma unroll
          for (int k = 0; k < WG_BK / 8; ++k) {
            tc::umma_tf32(tmem_base, ad + kstep * k, bd + kstep * k, idesc, (c > 0 || k > 0) ? 1u : 0u);
            if (SPLIT) {
              // the cross terms go to an accumulator of their own (columns 128..255): the tensor core's accumulation
              // truncates, so the long main sum should see as few additions as possible (see gemm_tc.cu)
              tc::umma_tf32(tmem_base + 128u, ad + LO + kstep * k, bd + kstep * k, idesc, (c > 0 || k > 0) ? 1u : 0u);
              tc::umma_tf32(tmem_base + 128u, ad + kstep * k, bd + LO + kstep * k, idesc, 1u);
            }
          }
          asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(wg_u32(empty_bar + st)) : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(wg_u32(done_bar)) : "memory");
      }
    } else if (warp >= 6) {
      // converters (fp32x3): lo tile of every landed stage (whole stage: the layout does not matter elementwise)
      if (SPLIT) {
        for (int c = 0; c < nchunks; ++c) {
          const int st = c % WG_STAGES;
          wg_wait(full_bar + st, (c / WG_STAGES) & 1);
          const uint32_t sx = wg_u32(smem + st * WG_STAGE_BYTES);
          tc::split_lo_tile(sx, sx + WG_RAW_BYTES, WG_RAW_BYTES, (int)threadIdx.x - 192, 128);
          tc::fence_async_smem();
          __syncwarp();
          if (lane == 0) tc::mbar_arrive(conv_bar + st);
        }
      }
    } else {
      // epilogue: TMEM lane = n_in index m, columns = n_out index
      wg_wait(done_bar, 0);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const int q = warp & 3;
      const int m = m0 + q * 32 + lane;
      const int ncols = min(128, op.n_out - n0);
      for (int c0 = 0; c0 < ncols; c0 += 32) {
        uint32_t r[32];
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c0;
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
        if (SPLIT) {
          float x[32];
          tc::tmem_ld32(taddr + 128u, x);
#pragma unroll
          for (int j = 0; j < 32; ++j) r[j] = __float_as_uint(__uint_as_float(r[j]) + x[j]);
        }
        if (m < op.n_in) {
          float* crow = op.C + (long long)m * op.n_out + n0 + c0;
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (c0 + j < ncols) atomicAdd(crow + j, __uint_as_float(r[j]));
        }
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(SPLIT ? 256u : 128u) : "memory");
}

// ---- host ---------------------------------------------------------------------------------------------
// operand stored [n, rows] with pitch ld floats: box = 128 operand rows x 32 reduction elements
static bool wg_map(CUtensorMap* m, const float* base, int n, int rows, long long ld) {
  return tc_make_map(m, base, (unsigned long long)rows, (unsigned long long)n, ld, WG_BK, 128, (int)CU_TENSOR_MAP_SWIZZLE_128B);
}
// operand stored ROW-major [rows, n] with pitch ld floats: box = 32 columns x WG_BK reduction rows, 32-byte-atom swizzle
static bool wg_map_mn(CUtensorMap* m, const float* base, int n, int rows, long long ld) {
  return tc_make_map(m, base, (unsigned long long)n, (unsigned long long)rows, ld, 32, WG_BK, (int)CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B);
}

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
  return op.ldc == op.N;
}

cudaError_t launch_wgrad_tc(const GemmGroup& gg, int split, int num_sms, cudaStream_t st) {
  if (gg.n <= 0) return cudaSuccess;
  WgGroup g;                             // kernel parameter (about 7 KB), built on the stack: nothing is shared between host threads
  g.n = gg.n;
  // work of an op ~ tiles * rows; share about one wave of CTAs between the ops in that proportion
  double total_work = 0.0;
  for (int i = 0; i < gg.n; ++i)
    total_work += (double)gg.ops[i].K * ((gg.ops[i].M + 127) / 128) * ((gg.ops[i].N + 127) / 128);
  static const double wg_waves = getenv("DINO_WG_WAVES") ? atof(getenv("DINO_WG_WAVES")) : 1.0;   // read once per process
  const double target_ctas = wg_waves * num_sms;
  int z = 0;
  for (int i = 0; i < gg.n; ++i) {
    const GemmOp& o = gg.ops[i];
    WgOp& w = g.op[i];
    w.C = o.C; w.n_in = o.M; w.n_out = o.N; w.rows = o.K;
    const int chunks = (o.K + WG_BK - 1) / WG_BK;
    const int m_tiles = (o.M + 127) / 128;
    w.n_tiles = (o.N + 127) / 128;
    const int tiles = m_tiles * w.n_tiles;
    long long splits = (long long)(target_ctas * ((double)o.K * tiles / total_work) / tiles + 0.999);
    // fp32x3: at most WG_X3_CHUNKS chunks (4 x that many additions into the main accumulator) per CTA; the partial sums
    // of the CTAs meet in fp32 atomics, which round to nearest
    if (split && splits * WG_X3_CHUNKS < chunks) splits = (chunks + WG_X3_CHUNKS - 1) / WG_X3_CHUNKS;
    if (splits > chunks) splits = chunks;
    if (splits < 1) splits = 1;
    w.chunks_per_cta = (int)((chunks + splits - 1) / splits);
    w.splits = (chunks + w.chunks_per_cta - 1) / w.chunks_per_cta;
    w.z_start = z;
    z += tiles * w.splits;
    w.mn = (o.a_sk == 1 && o.b_sk == 1) ? 0 : 1;
    if (w.mn) {
      if (!wg_map_mn(&g.mx[i], o.A, o.M, o.K, o.a_sk) || !wg_map_mn(&g.my[i], o.B, o.N, o.K, o.b_sk)) return cudaErrorInvalidValue;
    } else if (!wg_map(&g.mx[i], o.A, o.M, o.K, o.a_sm) || !wg_map(&g.my[i], o.B, o.N, o.K, o.b_sn)) return cudaErrorInvalidValue;
  }
  if (split) {
    constexpr int SMEM = WG_STAGES * 2 * WG_RAW_BYTES + 128 + 1024;
    static std::atomic<unsigned long long> mask{0};
    cudaError_t e = tc_set_smem_once(reinterpret_cast<const void*>(wgrad_tc_kernel<true>), SMEM, mask);
    if (e != cudaSuccess) return e;
    wgrad_tc_kernel<true><<<z, 320, SMEM, st>>>(g);
  } else {
    constexpr int SMEM = WG_STAGES * WG_RAW_BYTES + 128 + 1024;
    static std::atomic<unsigned long long> mask{0};
    cudaError_t e = tc_set_smem_once(reinterpret_cast<const void*>(wgrad_tc_kernel<false>), SMEM, mask);
    if (e != cudaSuccess) return e;
    wgrad_tc_kernel<false><<<z, 192, SMEM, st>>>(g);
  }
  return cudaGetLastError();
}

}  // namespace dino
