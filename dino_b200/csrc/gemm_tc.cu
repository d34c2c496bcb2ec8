// tcgen05 engine: C[M,N] = epilogue(A[M,K] * Bt[N,K]^T), fp32 accumulation in tensor memory, in two arithmetic modes:
//   tf32   one kind::tf32 MMA per 8-wide K step (operands read as TF32: 10 mantissa bits)
//   fp32x3 three MMAs per K step on on-chip hi/lo splits of BOTH operands (tc_common.cuh): fp32-grade products
//
// This is the Jacobian-chain contraction of the H1 step ([B*s, n] x [n, n'] with a per-sample column scale), the
// data-gradient contraction of the backward chain (with the cross-row d sigma sum fused into the epilogue) and the
// network's own forward / backward GEMMs (bias + softplus + sigmoid epilogue):
//   * operands are fp32 in HBM, both K-major; 128 x BK / BN x BK tiles arrive by TMA (cp.async.bulk.tensor, 128- or
//     64-byte swizzle) through an mbarrier ring;
//   * fp32x3: four converter warps turn every landed stage into its lo tile (x - trunc_tf32(x)) beside it;
//   * one elected thread issues tcgen05.mma.kind::tf32 (UMMA 128 x BN x 8), accumulators live in TMEM (two of them:
//     the epilogue of a tile overlaps the MMAs of the next);
//   * four epilogue warps read TMEM with tcgen05.ld (32 lanes x 32 columns per instruction), apply
//         rowsum[b,n] += acc * yref ; v = act(acc + bias) ; v = (v + add_pre) * scale[sample] + add_post
//     and write 32 x 32 blocks through a swizzled staging buffer, so that every store instruction covers whole
//     128-byte lines of C (a thread owns a ROW of the accumulator; stored directly, a warp would touch 32 lines per
//     instruction).  TMA stores were tried here first: with two 4 KB boxes in flight per warp their issue-to-read
//     latency, not bandwidth, set the pace (0.9 TB/s of output at K = 200).
// Warp roles: warp 0 TMA producer, warp 1 TMEM allocator + MMA issuer, then 4 or 8 epilogue warps, then (fp32x3) 8 converter warps.
// Persistent CTAs (one per SM) walk the 128 x BN output tiles.  Out-of-range rows/columns/K are zero-filled by TMA,
// so M, N, K are arbitrary; leading dimensions must be multiples of 4 floats.
#include <cuda.h>

#include <cstdlib>
#include <mutex>

#include "tc_epilogue.cuh"

namespace dino {

using namespace tc;

constexpr int TC_BM = 128;
// A gathered per sample, sample length a multiple of g = 16, 32, 64: a 128-row tile of consecutive minibatch rows is
// loaded as up to 128 / g boxes, each inside one sample; m[j] has a box of (j + 1) * g rows.  (Otherwise row tiles stop at
// sample ends and the last one of a sample is part empty: 4 tiles for 400 rows.)
struct TcGatherMaps { CUtensorMap m[8]; };
constexpr int TC_NCONV = 8;     // fp32x3: converter warps (a stage of 32-48 KB per 768 MMA cycles: four warps could not keep up)

// EPI: 0 plain store; 1 chain terms (rowsum / add_pre / scale / add_post / rounding); 2 network terms (bias / softplus +
// sigmoid / add_pre / add_post); 3 chain terms + row dot products (tc_epilogue.cuh)
// NEPI: epilogue warps (4 or 8): 8 split the column blocks of a lane quarter over two warps, for contractions whose
// main loop is short against their epilogue (tf32; fp32x3 with a short K)
template <int BN, int BK, int STAGES, bool SPLIT, int EPI, int NEPI, int NCONV>
__global__ void __launch_bounds__((2 + NEPI + (SPLIT ? NCONV : 0)) * 32, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b, const TcEpilogue ep,
               const __grid_constant__ TcGatherMaps gmaps, const __grid_constant__ CUtensorMap tma_b_lo) {
  constexpr int A_BYTES = TC_BM * BK * 4, B_BYTES = BN * BK * 4, RAW_BYTES = A_BYTES + B_BYTES;
  constexpr int STAGE_BYTES = SPLIT ? 2 * RAW_BYTES : RAW_BYTES;
  constexpr uint32_t ACC_COLS = BN < 32 ? 32 : BN;           // columns of one accumulator
  // fp32x3 keeps TWO accumulators per tile: the hi*hi products in one, the two small cross terms in the other.  The
  // tensor core adds into its fp32 accumulator with truncation, a bias of ~2^-25 of the accumulator's magnitude per
  // MMA; the cross terms are 2^-11 of the main sum, so their MMAs cost nothing in a separate accumulator, and the main
  // one sees K/8 additions instead of 3K/8.  The epilogue adds the two in fp32.
  constexpr uint32_t SET_COLS = SPLIT ? 2 * ACC_COLS : ACC_COLS;
  constexpr int NACC = 2 * SET_COLS <= 512 ? 2 : 1;          // two sets when they fit: the epilogue of tile i overlaps the MMAs of tile i+1
  constexpr uint32_t TMEM_COLS = NACC * SET_COLS;
  constexpr int EPI_BUF = 32 * 128;                          // one 32 x 32 fp32 block, 128-byte rows, chunks swizzled by the row
  constexpr int CONV0 = 2 + NEPI;                            // fp32x3: first of the four converter warps

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* epi_stage = smem + STAGES * STAGE_BYTES;          // [NEPI warps][EPI_BUF]
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(epi_stage + NEPI * EPI_BUF);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* conv_bar = empty_bar + STAGES;                   // lo tiles of the stage written (fp32x3)
  uint64_t* tmem_full_bar = conv_bar + STAGES;               // [2] accumulator complete
  uint64_t* tmem_empty_bar = tmem_full_bar + 2;              // [2] accumulator read out by the four epilogue warps
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty_bar + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nkb = (ep.K + BK - 1) / BK;
  // persistent: this CTA walks the output tiles blockIdx.x, blockIdx.x + gridDim.x, ... (column tile fastest)
  // (A gathered per sample: row tiles do not cross samples, ep.a_tps of them per sample)
  const int tiles_n = (ep.N + BN - 1) / BN;
  const bool per_sample = ep.a_rows > 0 && ep.a_gran == 0;
  const int n_tiles = tiles_n * (per_sample ? (ep.M / ep.a_rows) * ep.a_tps : (ep.M + TC_BM - 1) / TC_BM);

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_b) : "memory");
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < STAGES; ++i) { mbar_init(full_bar + i, 1); mbar_init(empty_bar + i, 1); mbar_init(conv_bar + i, NCONV > 0 ? NCONV : 1); }
      mbar_init(tmem_full_bar, 1); mbar_init(tmem_full_bar + 1, 1);
      mbar_init(tmem_empty_bar, NEPI); mbar_init(tmem_empty_bar + 1, NEPI);
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

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      int it = 0;                                    // running stage counter: the ring does not restart per tile
      for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int mt = tile / tiles_n, n0 = (tile % tiles_n) * BN;
        int m0 = mt * TC_BM, samp = 0;
        // gathered A: the boxes of this tile (rows inside a sample, sample number, height index)
        int nbox = 0, box_r[8], box_s[8], box_h[8], rows_loaded = TC_BM;
        if (per_sample) {
          const int b = mt / ep.a_tps;
          m0 = (mt % ep.a_tps) * TC_BM;              // row inside the sample
          samp = ep.a_idx ? ep.a_idx[b] : b;
        } else if (ep.a_rows > 0) {
          rows_loaded = min(TC_BM, ep.M - m0);
          for (int row = m0, left = rows_loaded; left > 0;) {
            const int b = row / ep.a_rows, r = row % ep.a_rows, h = min(left, ep.a_rows - r);
            box_r[nbox] = r; box_s[nbox] = ep.a_idx ? ep.a_idx[b] : b; box_h[nbox] = h; ++nbox;
            row += h; left -= h;
          }
        }
        for (int kb = 0; kb < nkb; ++kb, ++it) {
          const int st = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1;
          mbar_wait(empty_bar + st, ph ^ 1);
          uint8_t* sa = smem + st * STAGE_BYTES;
          const uint32_t lo_bytes = (SPLIT && ep.b_lo) ? (uint32_t)B_BYTES : 0u;
          if (nbox > 0) {
            mbar_expect_tx(full_bar + st, (uint32_t)(rows_loaded * BK * 4 + B_BYTES) + lo_bytes);
            int off = 0;
            for (int j = 0; j < nbox; ++j) {
              tma_load_3d(sa + off, &gmaps.m[box_h[j] / ep.a_gran - 1], kb * BK, box_r[j], box_s[j], full_bar + st);
              off += box_h[j] * BK * 4;
            }
          } else {
            mbar_expect_tx(full_bar + st, RAW_BYTES + lo_bytes);
            if (per_sample) tma_load_3d(sa, &tma_a, kb * BK, m0, samp, full_bar + st);
            else tma_load_2d(sa, &tma_a, kb * BK, m0, full_bar + st);
          }
          tma_load_2d(sa + A_BYTES, &tma_b, kb * BK, n0, full_bar + st);
          if (SPLIT && ep.b_lo) tma_load_2d(sa + RAW_BYTES + A_BYTES, &tma_b_lo, kb * BK, n0, full_bar + st);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (lane == 0) {
      const uint32_t idesc = idesc_tf32(BN);
      int it = 0, i = 0;
      for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++i) {
        const int acc = i % NACC;
        mbar_wait(tmem_empty_bar + acc, ((i / NACC) & 1) ^ 1);    // the epilogue has drained this accumulator set
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)acc * SET_COLS;
        for (int kb = 0; kb < nkb; ++kb, ++it) {
          const int st = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1;
          mbar_wait((SPLIT ? conv_bar : full_bar) + st, ph);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + st * STAGE_BYTES);
          const uint64_t adesc = BK == 32 ? desc_k_sw128(sa) : desc_k_sw64(sa);
          const uint64_t bdesc = BK == 32 ? desc_k_sw128(sa + A_BYTES) : desc_k_sw64(sa + A_BYTES);
          constexpr uint64_t LO = RAW_BYTES >> 4;    // the lo tiles sit RAW_BYTES above their sources
#pragma unroll
          for (int k = 0; k < BK / 8; ++k) {         // UMMA_K = 8 tf32 = 32 bytes = 2 descriptor units
            umma_tf32(tmem_d, adesc + 2 * k, bdesc + 2 * k, idesc, (kb > 0 || k > 0) ? 1u : 0u);
            if (SPLIT) {
              umma_tf32(tmem_d + ACC_COLS, adesc + LO + 2 * k, bdesc + 2 * k, idesc, (kb > 0 || k > 0) ? 1u : 0u);
              umma_tf32(tmem_d + ACC_COLS, adesc + 2 * k, bdesc + LO + 2 * k, idesc, 1u);
            }
          }
          umma_commit(empty_bar + st);               // frees the smem stage once these MMAs have read it
        }
        umma_commit(tmem_full_bar + acc);            // accumulator complete
      }
    }
  } else if (SPLIT && warp >= CONV0) {
    // ===== converters (fp32x3): lo tile of every landed stage =====
    {
      const int t = threadIdx.x - CONV0 * 32;
      int it = 0;
      for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        for (int kb = 0; kb < nkb; ++kb, ++it) {
          const int st = it % STAGES;
          mbar_wait(full_bar + st, (it / STAGES) & 1);
          const uint32_t sa = smem_u32(smem + st * STAGE_BYTES);
          split_lo_tile(sa, sa + RAW_BYTES, ep.b_lo ? A_BYTES : RAW_BYTES, t, NCONV * 32);
          fence_async_smem();                        // generic-proxy writes -> visible to the tensor core
          __syncwarp();
          if (lane == 0) mbar_arrive(conv_bar + st);
        }
      }
    }
  } else {
    // ===== epilogue: TMEM -> registers -> staging -> global, one accumulator behind the MMA issuer =====
    // The warps of the epilogue run a dependent instruction stream at ~13 cycles per instruction (1.5 warps per
    // scheduler): what counts is the NUMBER of instructions per 32 x 32 block.  The epilogue class is a template
    // parameter (no per-element tests of absent terms), whole blocks take unguarded vector loads, and tf32 (whose main
    // loop is 3x shorter) splits the column blocks of a lane quarter over two warps.
    const int q = warp & 3;                        // TMEM lane quarter this warp may access
    const int ew = warp - 2;                       // epilogue warp index
    const int half = ew >> 2;                      // NEPI == 8: even / odd column blocks
    const uint32_t my_stage = smem_u32(epi_stage + ew * EPI_BUF);
    int i = 0;
#pragma unroll 1
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++i) {
      const int mt = tile / tiles_n, n0 = (tile % tiles_n) * BN;
      int m0 = mt * TC_BM, m_lim = ep.M;
      if (ep.a_rows > 0 && ep.a_gran == 0) {
        const int b = mt / ep.a_tps, r0 = (mt % ep.a_tps) * TC_BM;
        m0 = b * ep.a_rows + r0;
        m_lim = b * ep.a_rows + min(ep.a_rows, r0 + TC_BM);
      }
      const int acc = i % NACC;
      mbar_wait(tmem_full_bar + acc, (i / NACC) & 1);
      tc_fence_after();
      tc_epilogue_tile<BN, SPLIT, EPI, NEPI, ACC_COLS>(ep, m0, n0, tmem_base + (uint32_t)acc * SET_COLS, my_stage, q, half, lane, m_lim);
      // this warp is done reading the accumulator: hand it back to the MMA issuer
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tmem_empty_bar + acc);
    }
  }
  // ===== teardown =====
  tc_fence_before();
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

cudaError_t tc_set_smem_once(const void* func, int bytes, std::atomic<unsigned long long>& mask) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  const unsigned long long bit = 1ull << (dev & 63);
  if (mask.load(std::memory_order_acquire) & bit) return cudaSuccess;
  e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e != cudaSuccess) return e;
  mask.fetch_or(bit, std::memory_order_release);
  return cudaSuccess;
}

int tc_current_num_sms() {
  static std::atomic<int> cache[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  int n = cache[dev & 63].load(std::memory_order_relaxed);
  if (n > 0) return n;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
  cache[dev & 63].store(n, std::memory_order_relaxed);
  return n;
}

namespace {
struct MapKey {
  const void* base; unsigned long long inner, outer; long long ld; unsigned bi, bo; int sw;
  bool operator==(const MapKey& o) const {
    return base == o.base && inner == o.inner && outer == o.outer && ld == o.ld && bi == o.bi && bo == o.bo && sw == o.sw;
  }
};
struct MapSlot { bool used; MapKey key; CUtensorMap map; };
constexpr int MAP_CACHE = 256;
}  // namespace

bool tc_make_map(CUtensorMap* out, const float* base, unsigned long long inner, unsigned long long outer, long long ld,
                 unsigned box_inner, unsigned box_outer, int swizzle) {
  static thread_local MapSlot cache[MAP_CACHE];      // direct-mapped; device addresses are unique across devices (UVA)
  const MapKey key{base, inner, outer, ld, box_inner, box_outer, swizzle};
  unsigned long long h = reinterpret_cast<uintptr_t>(base) >> 4;
  h = h * 0x9E3779B97F4A7C15ull + inner * 1315423911ull + outer * 2654435761ull + (unsigned long long)ld * 97 + box_inner * 31 + box_outer * 7 + swizzle;
  MapSlot& s = cache[(h >> 20) % MAP_CACHE];
  if (s.used && s.key == key) { *out = s.map; return true; }
  PFN_tmapEncodeTiled enc = get_encode();
  if (!enc) return false;
  cuuint64_t gdim[2] = {(cuuint64_t)inner, (cuuint64_t)outer};
  cuuint64_t gstride[1] = {(cuuint64_t)ld * sizeof(float)};
  cuuint32_t box[2] = {(cuuint32_t)box_inner, (cuuint32_t)box_outer};
  cuuint32_t estr[2] = {1, 1};
  if (enc(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), gdim, gstride, box, estr,
          CU_TENSOR_MAP_INTERLEAVE_NONE, (CUtensorMapSwizzle)swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
    return false;
  s.used = true; s.key = key; s.map = *out;
  return true;
}

bool tc_make_map3(CUtensorMap* out, const float* base, unsigned long long inner, unsigned long long rows,
                  unsigned long long samples, long long ld, long long ld2, unsigned box_inner, unsigned box_rows, int swizzle) {
  PFN_tmapEncodeTiled enc = get_encode();
  if (!enc) return false;
  cuuint64_t gdim[3] = {(cuuint64_t)inner, (cuuint64_t)rows, (cuuint64_t)samples};
  cuuint64_t gstride[2] = {(cuuint64_t)ld * sizeof(float), (cuuint64_t)ld2 * sizeof(float)};
  cuuint32_t box[3] = {(cuuint32_t)box_inner, (cuuint32_t)box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  return enc(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(base), gdim, gstride, box, estr,
             CU_TENSOR_MAP_INTERLEAVE_NONE, (CUtensorMapSwizzle)swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

bool gemm_tc_supported(const GemmOp& op) {
  if (op.a_sk != 1 || op.b_sk != 1) return false;                  // both operands K-major
  if (op.a_mod > 0 || op.b_kmod > 0) return false;
  if (op.atomic_c || op.k_per_split) return false;
  if (op.a_rows > 0 && (op.M % op.a_rows || (op.a_ld2 & 3) || op.a_nsamp < 1)) return false;
  if (op.rowdot && !op.yref) return false;
  if (op.C == nullptr && op.rowsum == nullptr) return false;
  if (op.C && ((op.ldc & 3) || (reinterpret_cast<uintptr_t>(op.C) & 15))) return false;
  if ((op.a_sm & 3) || (op.b_sn & 3)) return false;                // TMA: 16-byte row pitch
  if ((reinterpret_cast<uintptr_t>(op.A) & 15) || (reinterpret_cast<uintptr_t>(op.B) & 15)) return false;
  if (op.scale && ((op.ld_sc & 3) || (reinterpret_cast<uintptr_t>(op.scale) & 15))) return false;
  if (op.add_pre && ((op.ld_ap & 3) || (reinterpret_cast<uintptr_t>(op.add_pre) & 15))) return false;
  if (op.add_post && ((op.ld_apo & 3) || (reinterpret_cast<uintptr_t>(op.add_post) & 15))) return false;
  if (op.bias && (reinterpret_cast<uintptr_t>(op.bias) & 15)) return false;
  if (op.s_out && ((op.ld_s & 3) || (reinterpret_cast<uintptr_t>(op.s_out) & 15))) return false;
  if ((op.rowsum || op.rowdot) && (!op.yref || (op.ld_y & 3) || (reinterpret_cast<uintptr_t>(op.yref) & 15))) return false;
  if (op.act && !op.C) return false;
  return op.M > 0 && op.N > 0 && op.K > 0;
}

template <int BN, int BK, int STAGES, bool SPLIT, int EPI, int NEPI, int NCONV>
static cudaError_t launch_tc1(const GemmOp& op, cudaStream_t st) {
  CUtensorMap ma, mb;
  const int sw_in = BK == 32 ? (int)CU_TENSOR_MAP_SWIZZLE_128B : (int)CU_TENSOR_MAP_SWIZZLE_64B;
  // A[M,K] rows of a_sm floats; B operand is stored [N][K] (element (k,n) at n*b_sn + k)
  TcGatherMaps gm;
  int gran = 0;
  if (op.a_rows > 0) {
    if (!tc_make_map3(&ma, op.A, op.K, op.a_rows, op.a_nsamp, op.a_sm, op.a_ld2, BK, TC_BM, sw_in)) return cudaErrorInvalidValue;
    for (int g = 64; g >= 16 && gran == 0; g >>= 1)
      if (op.a_rows % g == 0) gran = g;
    if (op.a_rows % TC_BM == 0) gran = 0;            // tiles never cross a sample: one box each
    for (int j = 0; gran > 0 && j < TC_BM / gran; ++j)
      if (!tc_make_map3(&gm.m[j], op.A, op.K, op.a_rows, op.a_nsamp, op.a_sm, op.a_ld2, BK, (unsigned)((j + 1) * gran), sw_in)) return cudaErrorInvalidValue;
  } else if (!tc_make_map(&ma, op.A, op.K, op.M, op.a_sm, BK, TC_BM, sw_in)) return cudaErrorInvalidValue;
  if (!tc_make_map(&mb, op.B, op.K, op.N, op.b_sn, BK, BN, sw_in)) return cudaErrorInvalidValue;
  CUtensorMap mbl = mb;
  const bool b_lo = SPLIT && op.B_lo != nullptr && (reinterpret_cast<uintptr_t>(op.B_lo) & 15) == 0;
  if (b_lo && !tc_make_map(&mbl, op.B_lo, op.K, op.N, op.b_sn, BK, BN, sw_in)) return cudaErrorInvalidValue;
  TcEpilogue ep;
  ep.a_rows = op.a_rows; ep.a_tps = op.a_rows > 0 ? (op.a_rows + TC_BM - 1) / TC_BM : 0; ep.a_idx = op.a_idx;
  ep.rowdot = op.rowdot; ep.pre_first = op.pre_first;
  ep.a_gran = gran; ep.b_lo = b_lo ? 1 : 0;
  ep.M = op.M; ep.N = op.N; ep.K = op.K; ep.C = op.C; ep.ldc = op.ldc; ep.sdim = op.sdim > 0 ? op.sdim : 1;
  ep.scale = op.scale; ep.ld_sc = op.ld_sc; ep.add_pre = op.add_pre; ep.ld_ap = op.ld_ap;
  ep.add_post = op.add_post; ep.ld_apo = op.ld_apo; ep.bias = op.bias; ep.act = op.act;
  ep.s_out = op.act ? op.s_out : nullptr; ep.ld_s = op.ld_s;
  ep.yref = (op.rowsum || op.rowdot) ? op.yref : nullptr; ep.ld_y = op.ld_y; ep.rowsum = op.rowsum; ep.ld_rs = op.ld_rs;
  ep.round_tf32 = SPLIT ? 0 : op.round_tf32;
  constexpr int RAW = (TC_BM + BN) * BK * 4;
  constexpr int SMEM = STAGES * (SPLIT ? 2 * RAW : RAW) + NEPI * 32 * 128 + 256 + 1024;
  static_assert(SMEM <= 232448, "shared memory budget of one CTA");
  static std::atomic<unsigned long long> attr_mask{0};
  cudaError_t e = tc_set_smem_once(reinterpret_cast<const void*>(gemm_tc_kernel<BN, BK, STAGES, SPLIT, EPI, NEPI, NCONV>), SMEM, attr_mask);
  if (e != cudaSuccess) return e;
  // persistent CTAs, one per SM: each walks its share of the 128 x BN output tiles
  const int num_sms = tc_current_num_sms();
  const long long m_tiles = (op.a_rows > 0 && gran == 0) ? (long long)(op.M / op.a_rows) * ((op.a_rows + TC_BM - 1) / TC_BM) : (op.M + TC_BM - 1) / TC_BM;
  const long long n_tiles = (long long)((op.N + BN - 1) / BN) * m_tiles;
  const int grid = (int)(n_tiles < num_sms ? n_tiles : num_sms);
  gemm_tc_kernel<BN, BK, STAGES, SPLIT, EPI, NEPI, NCONV><<<grid, (2 + NEPI + (SPLIT ? NCONV : 0)) * 32, SMEM, st>>>(ma, mb, ep, gm, mbl);
  return cudaGetLastError();
}

template <int BN, int BK, int STAGES, bool SPLIT>
static cudaError_t launch_tc(const GemmOp& op, cudaStream_t st) {
  constexpr int NC = SPLIT ? TC_NCONV : 0;
  // fp32x3 with a short reduction (the width-100 layers of the full-Jacobian network: K = 100 against 17 KB rows): the
  // epilogue, not the main loop, sets the pace: eight epilogue warps and four converter warps (448 threads)
  constexpr bool FITS8 = STAGES * 2 * (TC_BM + BN) * BK * 4 + 8 * 32 * 128 + 1280 <= 232448;
  const bool short_k = SPLIT && FITS8 && op.K <= 256;
  if (op.bias || op.act) return launch_tc1<BN, BK, STAGES, SPLIT, 2, SPLIT ? 4 : 8, NC>(op, st);
  if (op.rowdot || op.pre_first) {      // the first layer without its Jacobian (plan.cu): class 1 plus row dot products
    if (short_k) return launch_tc1<BN, BK, STAGES, SPLIT, 3, 8, SPLIT ? 4 : 0>(op, st);
    return launch_tc1<BN, BK, STAGES, SPLIT, 3, SPLIT ? 4 : 8, NC>(op, st);
  }
  if (op.rowsum || op.scale || op.add_pre || op.add_post || (op.round_tf32 && !SPLIT)) {
    if (short_k) return launch_tc1<BN, BK, STAGES, SPLIT, 1, 8, SPLIT ? 4 : 0>(op, st);
    return launch_tc1<BN, BK, STAGES, SPLIT, 1, SPLIT ? 4 : 8, NC>(op, st);
  }
  if (short_k) return launch_tc1<BN, BK, STAGES, SPLIT, 0, 8, SPLIT ? 4 : 0>(op, st);
  return launch_tc1<BN, BK, STAGES, SPLIT, 0, SPLIT ? 4 : 8, NC>(op, st);
}

cudaError_t launch_gemm_tc(const GemmOp& op, int split, cudaStream_t st) {
  if (!gemm_tc_supported(op)) return cudaErrorNotSupported;
  static const int variant = getenv("DINO_TC_VARIANT") ? atoi(getenv("DINO_TC_VARIANT")) : 0;   // read once per process
  static const int cg2 = getenv("DINO_TC_CG2") ? atoi(getenv("DINO_TC_CG2")) : 3;              // CTA-pair kernel for bit 0: tf32, bit 1: fp32x3
  if (((split ? 2 : 1) & cg2) && op.a_rows == 0 && !op.rowdot && !op.pre_first && gemm_tc2_wanted(op)) return launch_gemm_tc2(op, split, st);
  const int num_sms = tc_current_num_sms();
  const long long m_tiles = (op.M + TC_BM - 1) / TC_BM;
  // widest tile that still gives every SM work (small-M contractions: the network's own forward / backward)
  int bn = 256;
  if (op.N <= 128 || m_tiles * ((op.N + 255) / 256) * 2 <= num_sms) bn = 128;
  if (op.N <= 64 || (bn == 128 && m_tiles * ((op.N + 127) / 128) * 2 <= num_sms)) bn = 64;
  if (split) {
    if (bn == 64) return launch_tc<64, 32, 4, true>(op, st);
    if (bn == 128) return launch_tc<128, 32, 3, true>(op, st);
    if (variant == 1) return launch_tc<256, 32, 2, true>(op, st);
    if (variant == 2) return launch_tc<256, 16, 4, true>(op, st);
    return launch_tc<128, 32, 3, true>(op, st);        // measured best of the three at [25600, 1024] x [1024, 1024] (profiles/r02b)
  }
  if (bn == 64) return launch_tc<64, 32, 4, false>(op, st);
  if (bn == 128) return launch_tc<128, 32, 4, false>(op, st);
  return launch_tc<256, 32, 4, false>(op, st);
}

}  // namespace dino
