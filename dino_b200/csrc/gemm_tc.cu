// tcgen05 engine: C[M,N] = A[M,K] * Bt[N,K]^T in TF32 with fp32 accumulation in tensor memory.
//
// This is the Jacobian-chain contraction of the H1 step ([B*s, n] x [n, n'] with a per-sample column
// scale) on the 5th-generation tensor cores:
//   * operands are fp32 in HBM (the hardware reads them as TF32), both K-major; 128 x 32 / BN x 32
//     tiles are brought in by TMA (cp.async.bulk.tensor, SWIZZLE_128B) through a 4-stage mbarrier ring;
//   * one elected thread issues tcgen05.mma.kind::tf32 (UMMA 128 x BN x 8), accumulators live in TMEM;
//   * four epilogue warps read TMEM with tcgen05.ld (32 lanes x 32 columns per instruction) and apply
//     v = (acc + add_pre) * scale[sample] + add_post before storing fp32.
// Warp roles (192 threads): warp 0 TMA producer, warp 1 TMEM allocator + MMA issuer, warps 2-5 epilogue.
// Persistent CTAs (one per SM) walk the 128 x BN output tiles; two TMEM accumulators let the epilogue of a tile overlap
// the MMAs of the next.  Out-of-range rows/columns/K are zero-filled by TMA and masked
// in the epilogue, so M, N, K are arbitrary; leading dimensions must be multiples of 4 floats.
#include <cuda.h>

#include <mutex>

#include "common.cuh"

namespace dino {

constexpr int TC_BM = 128, TC_BK = 32, TC_STAGES = 4, TC_THREADS = 192;

struct TcEpilogue {
  int M, N, K;
  float* C; long long ldc;
  int sdim;
  const float* scale; long long ld_sc;
  const float* add_pre; long long ld_ap;
  const float* add_post; long long ld_apo;
  int round_tf32;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra WAIT_DONE;\n\t"
      "bra WAIT_LOOP;\n\t"
      "WAIT_DONE:\n\t"
      "}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// K-major operand tile, 128-byte rows, SWIZZLE_128B: 8-row groups 1024 B apart (SBO), LBO unused (=1).
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3FFFF) >> 4) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}

template <int BN>
__global__ void __launch_bounds__(TC_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b, const TcEpilogue ep) {
  constexpr int A_BYTES = TC_BM * TC_BK * 4, B_BYTES = BN * TC_BK * 4, STAGE_BYTES = A_BYTES + B_BYTES;
  constexpr uint32_t ACC_COLS = BN < 32 ? 32 : BN;           // columns of one accumulator
  constexpr uint32_t TMEM_COLS = 2 * ACC_COLS;               // two accumulators: the epilogue of tile i overlaps the MMAs of tile i+1
  // instruction descriptor: D fp32, A/B tf32, both K-major, N = BN, M = 128
  constexpr uint32_t IDESC = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + TC_STAGES * STAGE_BYTES);
  uint64_t* empty_bar = full_bar + TC_STAGES;
  uint64_t* tmem_full_bar = empty_bar + TC_STAGES;          // [2] accumulator complete
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;             // [2] accumulator read out by the four epilogue warps
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nkb = (ep.K + TC_BK - 1) / TC_BK;
  // persistent: this CTA walks the output tiles blockIdx.x, blockIdx.x + gridDim.x, ... (column tile fastest)
  const int tiles_n = (ep.N + BN - 1) / BN, n_tiles = tiles_n * ((ep.M + TC_BM - 1) / TC_BM);

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_b) : "memory");
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < TC_STAGES; ++i) { mbar_init(full_bar + i, 1); mbar_init(empty_bar + i, 1); }
      mbar_init(tmem_full_bar, 1); mbar_init(tmem_full_bar + 1, 1);
      mbar_init(tmem_empty_bar, 4); mbar_init(tmem_empty_bar + 1, 4);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(TMEM_COLS) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      int it = 0;                                    // running stage counter: the ring does not restart per tile
      for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int m0 = (tile / tiles_n) * TC_BM, n0 = (tile % tiles_n) * BN;
        for (int kb = 0; kb < nkb; ++kb, ++it) {
          const int st = it % TC_STAGES;
          const uint32_t ph = (it / TC_STAGES) & 1;
          mbar_wait(empty_bar + st, ph ^ 1);
          mbar_expect_tx(full_bar + st, STAGE_BYTES);
          uint8_t* sa = smem + st * STAGE_BYTES;
          tma_load_2d(sa, &tma_a, kb * TC_BK, m0, full_bar + st);
          tma_load_2d(sa + A_BYTES, &tma_b, kb * TC_BK, n0, full_bar + st);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (lane == 0) {
      int it = 0, i = 0;
      for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++i) {
        const int acc = i & 1;
        mbar_wait(tmem_empty_bar + acc, ((i >> 1) & 1) ^ 1);      // the epilogue has drained this accumulator (two tiles ago)
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t tmem_d = tmem_base + (uint32_t)acc * ACC_COLS;
        for (int kb = 0; kb < nkb; ++kb, ++it) {
          const int st = it % TC_STAGES;
          const uint32_t ph = (it / TC_STAGES) & 1;
          mbar_wait(full_bar + st, ph);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint32_t sa = smem_u32(smem + st * STAGE_BYTES);
          const uint64_t adesc = umma_desc_sw128(sa), bdesc = umma_desc_sw128(sa + A_BYTES);
#pragma unroll
          for (int k = 0; k < TC_BK / 8; ++k)        // UMMA_K = 8 tf32 = 32 bytes = 2 descriptor units
            umma_tf32(tmem_d, adesc + 2 * k, bdesc + 2 * k, IDESC, (kb > 0 || k > 0) ? 1u : 0u);
          umma_commit(empty_bar + st);               // frees the smem stage once these MMAs have read it
        }
        umma_commit(tmem_full_bar + acc);            // accumulator complete
      }
    }
  } else {
    // ===== epilogue: TMEM -> registers -> global, one accumulator behind the MMA issuer =====
    const int q = warp & 3;                        // TMEM lane quarter this warp may access
    int i = 0;
#pragma unroll 1
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++i) {
    const int m0 = (tile / tiles_n) * TC_BM, n0 = (tile % tiles_n) * BN;
    const int acc = i & 1;
    mbar_wait(tmem_full_bar + acc, (i >> 1) & 1);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const int row = m0 + q * 32 + lane;
    const bool row_ok = row < ep.M;
    const long long b = row_ok ? row / ep.sdim : 0;
    float* crow = ep.C + (long long)row * ep.ldc;
    const float* sc = ep.scale ? ep.scale + b * ep.ld_sc : nullptr;
    const float* ap = ep.add_pre ? ep.add_pre + (long long)row * ep.ld_ap : nullptr;
    const float* apo = ep.add_post ? ep.add_post + (long long)row * ep.ld_apo : nullptr;
#pragma unroll 1
    for (int c0 = 0; c0 < BN; c0 += 32) {
      if (n0 + c0 >= ep.N) break;                  // warp-uniform
      uint32_t r[32];
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)acc * ACC_COLS + (uint32_t)c0;
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
      if (row_ok) {
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          const int n = n0 + c0 + j;
          if (n + 3 < ep.N) {
            float4 v = make_float4(__uint_as_float(r[j]), __uint_as_float(r[j + 1]), __uint_as_float(r[j + 2]),
                                   __uint_as_float(r[j + 3]));
            if (ap) { const float4 t = *reinterpret_cast<const float4*>(ap + n); v.x += t.x; v.y += t.y; v.z += t.z; v.w += t.w; }
            if (sc) { const float4 t = *reinterpret_cast<const float4*>(sc + n); v.x *= t.x; v.y *= t.y; v.z *= t.z; v.w *= t.w; }
            if (apo) { const float4 t = *reinterpret_cast<const float4*>(apo + n); v.x += t.x; v.y += t.y; v.z += t.z; v.w += t.w; }
            if (ep.round_tf32) { v.x = round_tf32_f(v.x); v.y = round_tf32_f(v.y); v.z = round_tf32_f(v.z); v.w = round_tf32_f(v.w); }
            *reinterpret_cast<float4*>(crow + n) = v;
          } else {
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              if (n + u < ep.N) {
                float v = __uint_as_float(r[j + u]);
                if (ap) v += ap[n + u];
                if (sc) v *= sc[n + u];
                if (apo) v += apo[n + u];
                crow[n + u] = ep.round_tf32 ? round_tf32_f(v) : v;
              }
            }
          }
        }
      }
    }
    // this warp is done reading the accumulator: hand it back to the MMA issuer
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncwarp();
    if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(tmem_empty_bar + acc)) : "memory");
    }
  }
  // ===== teardown =====
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
}

// ---- host side ---------------------------------------------------------------------------------------
typedef CUresult (*PFN_tmapEncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                        const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                        CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static PFN_tmapEncodeTiled g_encode = nullptr;
static std::once_flag g_encode_once;

static PFN_tmapEncodeTiled get_encode() {
  std::call_once(g_encode_once, [] {
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      g_encode = reinterpret_cast<PFN_tmapEncodeTiled>(fn);
  });
  return g_encode;
}

// 2-D fp32 row-major [rows, cols] (leading dimension ld floats) -> tiles of box_rows x 32 columns.
static bool make_map(CUtensorMap* map, const float* base, long long rows, long long cols, long long ld, int box_rows) {
  PFN_tmapEncodeTiled enc = get_encode();
  if (!enc) return false;
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)ld * sizeof(float)};
  cuuint32_t box[2] = {(cuuint32_t)TC_BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  return enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), gdim, gstride, box, estr,
             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

bool gemm_tc_supported(const GemmOp& op) {
  if (op.a_sk != 1 || op.b_sk != 1) return false;                  // both operands K-major
  if (op.a_mod > 0 || op.b_kmod > 0) return false;
  if (op.rowsum || op.act || op.bias || op.atomic_c || op.k_per_split) return false;
  if (op.C == nullptr || (op.ldc & 3)) return false;
  if ((op.a_sm & 3) || (op.b_sn & 3)) return false;                // TMA: 16-byte row pitch
  if ((reinterpret_cast<uintptr_t>(op.A) & 15) || (reinterpret_cast<uintptr_t>(op.B) & 15) ||
      (reinterpret_cast<uintptr_t>(op.C) & 15)) return false;
  if (op.scale && ((op.ld_sc & 3) || (reinterpret_cast<uintptr_t>(op.scale) & 15))) return false;
  if (op.add_pre && ((op.ld_ap & 3) || (reinterpret_cast<uintptr_t>(op.add_pre) & 15))) return false;
  if (op.add_post && ((op.ld_apo & 3) || (reinterpret_cast<uintptr_t>(op.add_post) & 15))) return false;
  return op.M > 0 && op.N > 0 && op.K > 0;
}

template <int BN>
static cudaError_t launch_tc(const GemmOp& op, cudaStream_t st) {
  CUtensorMap ma, mb;
  // A[M,K] rows of a_sm floats; B operand is stored [N][K] (element (k,n) at n*b_sn + k)
  if (!make_map(&ma, op.A, op.M, op.K, op.a_sm, TC_BM) || !make_map(&mb, op.B, op.N, op.K, op.b_sn, BN))
    return cudaErrorInvalidValue;
  TcEpilogue ep;
  ep.M = op.M; ep.N = op.N; ep.K = op.K; ep.C = op.C; ep.ldc = op.ldc; ep.sdim = op.sdim > 0 ? op.sdim : 1;
  ep.scale = op.scale; ep.ld_sc = op.ld_sc; ep.add_pre = op.add_pre; ep.ld_ap = op.ld_ap;
  ep.add_post = op.add_post; ep.ld_apo = op.ld_apo; ep.round_tf32 = op.round_tf32;
  constexpr int SMEM = TC_STAGES * (TC_BM * TC_BK * 4 + BN * TC_BK * 4) + 256 + 1024;
  static bool attr_done = false;
  if (!attr_done) {
    cudaError_t e = cudaFuncSetAttribute(gemm_tc_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    if (e != cudaSuccess) return e;
    attr_done = true;
  }
  // persistent CTAs, one per SM: each walks its share of the 128 x BN output tiles
  static int num_sms = 0;
  if (num_sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    if (cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || num_sms <= 0) num_sms = 148;
  }
  const long long n_tiles = (long long)((op.N + BN - 1) / BN) * ((op.M + TC_BM - 1) / TC_BM);
  const int grid = (int)(n_tiles < num_sms ? n_tiles : num_sms);
  gemm_tc_kernel<BN><<<grid, TC_THREADS, SMEM, st>>>(ma, mb, ep);
  return cudaGetLastError();
}

cudaError_t launch_gemm_tc(const GemmOp& op, cudaStream_t st) {
  if (!gemm_tc_supported(op)) return cudaErrorNotSupported;
  if (op.N <= 64) return launch_tc<64>(op, st);
  if (op.N <= 128) return launch_tc<128>(op, st);
  return launch_tc<256>(op, st);
}

}  // namespace dino
