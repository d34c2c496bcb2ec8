// Fused MLP forward / backward for SMALL networks (every BASELINE reduced-basis config with widths
// <= ~128): one CTA carries a tile of TS samples through ALL layers, activations stay in shared
// memory, weights stream from L2 (they total a few hundred KB).  This replaces 2L+ dependent
// latency-bound GEMM launches (measured 20 us each, profiles/r01a_summary.md) by two launches.
// Exact fp32 FFMA arithmetic, same formulas as the layer-by-layer path (plan.cu header).
//
// Large networks keep the layer-by-layer GEMM path: there each weight must be reused across many
// samples per load, which is what the tiled engines do.
#include "tc_common.cuh"

namespace dino {

constexpr int TS = 2;            // samples per CTA
constexpr int MT = 256;          // threads per CTA

static inline int cdiv_i(int a, int b) { return (a + b - 1) / b; }

// Asynchronous copy of one layer's kernel block (a flat run of r4(n_in*n_out) floats, 16-byte aligned in the
// parameter arena) into shared memory: cp.async, no register staging, one commit group per layer.
__device__ __forceinline__ void stage_weights_async(float* dst, const float* __restrict__ src, int floats) {
  const int vecs = (floats + 3) >> 2;
  for (int e = threadIdx.x; e < vecs; e += MT) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(dst + 4 * e);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(src + 4 * e) : "memory");
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
}
__device__ __forceinline__ void wait_weights_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void wait_weights_but_one() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }

// ---- forward: one layer for the CTA's TS samples --------------------------------------------------
// NC threads span the columns, the remaining MT/NC thread groups split the reduction index k; the
// partial sums meet in shared memory.  (The layer is a latency chain of L2 weight loads; splitting k
// shortens it by MT/NC.)
template <int NC>
__device__ __forceinline__ void fwd_layer(const MlpLayerDev& ly, const float* __restrict__ params, const float* W,
                                          const float* zin, int ld_zin, float* zout, const float* zskip, int ld_skip,
                                          float* __restrict__ ws, int b0, int nb, float* red) {
  constexpr int KG = MT / NC;
  const int tid = threadIdx.x;
  const int cl = tid % NC, kg = tid / NC;
  for (int c0 = 0; c0 < ly.n_out; c0 += NC) {
    const int c = c0 + cl;
    float acc[TS];
#pragma unroll
    for (int t = 0; t < TS; ++t) acc[t] = 0.f;
    if (c < ly.n_out) {
#pragma unroll 4
      for (int k = kg; k < ly.n_in; k += KG) {
        const float w = W[k * ly.n_out + c];                 // shared memory, lanes = consecutive columns
#pragma unroll
        for (int t = 0; t < TS; ++t) acc[t] = fmaf(zin[t * ld_zin + k], w, acc[t]);
      }
    }
#pragma unroll
    for (int t = 0; t < TS; ++t) red[(kg * TS + t) * NC + cl] = acc[t];
    __syncthreads();
    for (int e = tid; e < TS * NC; e += MT) {
      const int t = e / NC, cc = c0 + e % NC;
      if (cc >= ly.n_out) continue;
      float v = ly.boff >= 0 ? __ldg(params + ly.boff + cc) : 0.f;
#pragma unroll
      for (int g = 0; g < KG; ++g) v += red[(g * TS + t) * NC + e % NC];
      if (ly.act) {
        if (t < nb) ws[ly.sg_off + (long long)(b0 + t) * ly.ld_out + cc] = sigmoid_f(v);
        v = softplus_f(v);
      }
      if (zskip != nullptr) v += zskip[t * ld_skip + cc];
      zout[t * ly.ld_out + cc] = v;
      if (t < nb) ws[ly.z_off + (long long)(b0 + t) * ly.ld_out + cc] = v;
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(MT) mlp_forward_fused_kernel(const __grid_constant__ MlpNetDev net,
                                                               const float* __restrict__ params, float* __restrict__ ws,
                                                               const float* __restrict__ m, const int* __restrict__ idx,
                                                               int B, int w_floats) {
  extern __shared__ float sm[];
  __shared__ float red[MT * TS];
  const int b0 = blockIdx.x * TS;
  const int nb = min(TS, B - b0);
  // shared layout: weight tiles [2][w_floats] | z_0 [TS, ld_in] | the output of every layer [TS, ld_out]
  float* Wbuf[2] = {sm, sm + w_floats};
  float* z0 = sm + 2 * w_floats;
  auto wsize = [&](int l) { return (net.layer[l].n_in * net.layer[l].n_out + 3) & ~3; };
  stage_weights_async(Wbuf[0], params + net.layer[0].koff, wsize(0));
  for (int e = threadIdx.x; e < TS * net.ld_in; e += MT) {
    const int t = e / net.ld_in, k = e % net.ld_in;
    float v = 0.f;
    if (t < nb && k < net.d_in) {
      const long long row = idx ? idx[b0 + t] : (b0 + t);
      v = __ldg(m + row * net.d_in + k);
      ws[net.z0_off + (long long)(b0 + t) * net.ld_in + k] = v;
    }
    z0[e] = v;
  }
  int off = TS * net.ld_in;
  const float* zin = z0;
  int ld_zin = net.ld_in;
#pragma unroll 1
  for (int l = 0; l < net.L; ++l) {
    const MlpLayerDev& ly = net.layer[l];
    // prefetch the next layer's weights into the other buffer (its previous reader, layer l-1, has finished:
    // every fwd_layer ends with a block-wide barrier), then wait for this layer's tile
    if (l + 1 < net.L) { stage_weights_async(Wbuf[(l + 1) & 1], params + net.layer[l + 1].koff, wsize(l + 1)); wait_weights_but_one(); }
    else wait_weights_all();
    __syncthreads();
    const float* W = Wbuf[l & 1];
    float* zout = z0 + off;
    const float* zskip = nullptr;
    int ld_skip = 0;
    if (ly.skip_from >= 0) {
      // input of layer skip_from: z_0 or the output of layer skip_from-1
      int so = 0, j = 0, ld = net.ld_in;
      for (; j < ly.skip_from; ++j) { so += TS * ld; ld = net.layer[j].ld_out; }
      zskip = z0 + so; ld_skip = ld;
    }
    if (ly.n_out <= 64) fwd_layer<64>(ly, params, W, zin, ld_zin, zout, zskip, ld_skip, ws, b0, nb, red);
    else if (ly.n_out <= 128) fwd_layer<128>(ly, params, W, zin, ld_zin, zout, zskip, ld_skip, ws, b0, nb, red);
    else fwd_layer<256>(ly, params, W, zin, ld_zin, zout, zskip, ld_skip, ws, b0, nb, red);
    zin = zout; ld_zin = ly.ld_out;
    off += TS * ly.ld_out;
  }
}

// ---- backward ---------------------------------------------------------------------------------------
template <int NC>
__device__ __forceinline__ void bwd_dz_layer(const MlpLayerDev& ly, const float* Ws, int ldw, const float* da,
                                             int ld_da, float* dz_out, int ld_dz, const float* dz_skip, float* red) {
  constexpr int NG = MT / NC;                      // thread groups splitting the reduction index n
  const int tid = threadIdx.x;
  const int kl = tid % NC, ng = tid / NC;
  for (int k0 = 0; k0 < ly.n_in; k0 += NC) {
    const int k = k0 + kl;
    float acc[TS];
#pragma unroll
    for (int t = 0; t < TS; ++t) acc[t] = 0.f;
    if (k < ly.n_in) {
      const float* wrow = Ws + k * ldw;
#pragma unroll 4
      for (int n = ng; n < ly.n_out; n += NG) {
        const float w = wrow[n];
#pragma unroll
        for (int t = 0; t < TS; ++t) acc[t] = fmaf(da[t * ld_da + n], w, acc[t]);
      }
    }
#pragma unroll
    for (int t = 0; t < TS; ++t) red[(ng * TS + t) * NC + kl] = acc[t];
    __syncthreads();
    for (int e = tid; e < TS * NC; e += MT) {
      const int t = e / NC, kk = k0 + e % NC;
      if (kk >= ly.n_in) continue;
      float v = 0.f;
#pragma unroll
      for (int g = 0; g < NG; ++g) v += red[(g * TS + t) * NC + e % NC];
      if (dz_skip != nullptr) v += dz_skip[t * ld_dz + kk];
      dz_out[t * ld_dz + kk] = v;
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(MT) mlp_backward_fused_kernel(const __grid_constant__ MlpNetDev net,
                                                                const float* __restrict__ params, float* __restrict__ ws,
                                                                float* __restrict__ grad, int B, int with_rowsum,
                                                                int w_floats) {
  extern __shared__ float sm[];
  __shared__ float red[MT * TS];
  const int b0 = blockIdx.x * TS;
  const int nb = min(TS, B - b0);
  const int ld_dz = net.ld_dz;
  // shared layout: W tiles [2][w_floats] | dz[3][TS, ld_dz] | da [TS, ld_dz]
  float* Wbuf[2] = {sm, sm + w_floats};
  float* dzb = sm + 2 * w_floats;
  if (net.L - 1 > net.lowest_trainable)
    stage_weights_async(Wbuf[(net.L - 1) & 1], params + net.layer[net.L - 1].koff,
                        (net.layer[net.L - 1].n_in * net.layer[net.L - 1].n_out + 3) & ~3);
  float* da = dzb + 3 * TS * ld_dz;
  // sigmoids and row sums of every layer for this tile's samples, fetched up front with all loads in flight (inside the
  // layer loop each of them would be an exposed L2 round trip): [L][TS][ld_dz] each
  float* sg_s = da + TS * ld_dz;
  float* rs_s = sg_s + net.L * TS * ld_dz;
  for (int l = net.lowest_trainable; l < net.L; ++l) {
    const MlpLayerDev& ly = net.layer[l];
    if (!ly.act) continue;
    const bool has_rs = with_rowsum && ly.rs_off >= 0;
    for (int e = threadIdx.x; e < TS * ly.n_out; e += MT) {
      const int t = e / ly.n_out, n = e % ly.n_out;
      float sv = 0.f, rv = 0.f;
      if (t < nb) {
        const long long row = (long long)(b0 + t) * ly.ld_out + n;
        sv = ws[ly.sg_off + row];
        if (has_rs) rv = ws[ly.rs_off + row];
      }
      sg_s[(l * TS + t) * ld_dz + n] = sv;
      rs_s[(l * TS + t) * ld_dz + n] = rv;
    }
  }
  auto dz_buf = [&](int l) { return dzb + (l % 3) * TS * ld_dz; };
  {
    float* top = dz_buf(net.L);
    for (int e = threadIdx.x; e < TS * ld_dz; e += MT) {
      const int t = e / ld_dz;
      top[e] = t < nb ? ws[net.dz_top_off + (long long)(b0 + t) * ld_dz + (e % ld_dz)] : 0.f;
    }
  }
  __syncthreads();
#pragma unroll 1
  for (int l = net.L - 1; l >= net.lowest_trainable; --l) {
    const MlpLayerDev& ly = net.layer[l];
    const float* g = dz_buf(l + 1);
    // da_l = dz_{l+1} * sg + rowsum * (1 - sg)  (softplus) | dz_{l+1} (linear); also stored for the weight gradient
    for (int e = threadIdx.x; e < TS * ly.n_out; e += MT) {
      const int t = e / ly.n_out, n = e % ly.n_out;
      float v = g[t * ld_dz + n];
      if (t < nb) {
        const long long row = (long long)(b0 + t) * ly.ld_out + n;
        if (ly.act) {
          const float sv = sg_s[(l * TS + t) * ld_dz + n];
          const float rs = rs_s[(l * TS + t) * ld_dz + n];
          v = fmaf(rs, 1.f - sv, v * sv);
        }
        if (ly.trainable) ws[ly.da_off + row] = v;
      } else {
        v = 0.f;
      }
      da[t * ld_dz + n] = v;
    }
    const bool need_dz = l > net.lowest_trainable;
    const int ldw = ly.n_out;                        // natural pitch of the asynchronously staged tile
    const float* Ws = Wbuf[l & 1];
    // prefetch the weights of the layer below (needed only if that layer still propagates dz)
    if (l - 1 > net.lowest_trainable) {
      stage_weights_async(Wbuf[(l - 1) & 1], params + net.layer[l - 1].koff, (net.layer[l - 1].n_in * net.layer[l - 1].n_out + 3) & ~3);
      wait_weights_but_one();
    } else {
      wait_weights_all();
    }
    __syncthreads();
    if (ly.trainable && ly.boff >= 0) {
      for (int n = threadIdx.x; n < ly.n_out; n += MT) {
        float sacc = 0.f;
#pragma unroll
        for (int t = 0; t < TS; ++t) sacc += da[t * ld_dz + n];
        atomicAdd(grad + ly.boff + n, sacc);
      }
    }
    if (need_dz) {
      const float* dz_skip = ly.lands_skip >= 0 ? dz_buf(ly.lands_skip + 1) : nullptr;
      float* out = dz_buf(l);
      if (ly.n_in <= 64) bwd_dz_layer<64>(ly, Ws, ldw, da, ld_dz, out, ld_dz, dz_skip, red);
      else if (ly.n_in <= 128) bwd_dz_layer<128>(ly, Ws, ldw, da, ld_dz, out, ld_dz, dz_skip, red);
      else bwd_dz_layer<256>(ly, Ws, ldw, da, ld_dz, out, ld_dz, dz_skip, red);
    }
    __syncthreads();
  }
}

// ---- host ---------------------------------------------------------------------------------------------
static int fwd_w_floats(const MlpNetDev& net) {
  int w = 0;
  for (int l = 0; l < net.L; ++l) {
    const int f = (net.layer[l].n_in * net.layer[l].n_out + 3) & ~3;
    w = f > w ? f : w;
  }
  return w;
}
static size_t fwd_smem(const MlpNetDev& net) {
  size_t f = (size_t)TS * net.ld_in + 2 * (size_t)fwd_w_floats(net);
  for (int l = 0; l < net.L; ++l) f += (size_t)TS * net.layer[l].ld_out;
  return f * sizeof(float);
}
static int bwd_w_floats(const MlpNetDev& net) {
  int w = 0;
  for (int l = net.lowest_trainable + 1; l < net.L; ++l) {
    const int f = (net.layer[l].n_in * net.layer[l].n_out + 3) & ~3;
    w = f > w ? f : w;
  }
  return w > 4 ? w : 4;
}
static size_t bwd_smem(const MlpNetDev& net) {
  return (2 * (size_t)bwd_w_floats(net) + (4 + 2 * (size_t)net.L) * TS * net.ld_dz) * sizeof(float);
}

bool mlp_fused_supported(const MlpNetDev& net) {
  if (net.L < 1 || net.L > MLP_FUSED_MAX_LAYERS) return false;
  long long params = 0;
  for (int l = 0; l < net.L; ++l) {
    params += (long long)net.layer[l].n_in * net.layer[l].n_out;
    if (net.layer[l].n_out > 512) return false;
  }
  if (params > 256 * 1024) return false;
  return fwd_smem(net) <= 160 * 1024 && bwd_smem(net) <= 160 * 1024;
}

cudaError_t launch_mlp_forward_fused(const MlpNetDev& net, const float* params, float* ws, const float* m,
                                     const int* idx, int B, cudaStream_t st) {
  const size_t smem = fwd_smem(net);
  // the attribute is per device: raise it once per device to the largest size mlp_fused_supported() admits
  static std::atomic<unsigned long long> mask{0};
  cudaError_t e = tc_set_smem_once(reinterpret_cast<const void*>(mlp_forward_fused_kernel), 160 * 1024, mask);
  if (e != cudaSuccess) return e;
  mlp_forward_fused_kernel<<<cdiv_i(B, TS), MT, smem, st>>>(net, params, ws, m, idx, B, fwd_w_floats(net));
  return cudaGetLastError();
}

cudaError_t launch_mlp_backward_fused(const MlpNetDev& net, const float* params, float* ws, float* grad, int B,
                                      int with_rowsum, cudaStream_t st) {
  const size_t smem = bwd_smem(net);
  static std::atomic<unsigned long long> mask{0};
  cudaError_t e = tc_set_smem_once(reinterpret_cast<const void*>(mlp_backward_fused_kernel), 160 * 1024, mask);
  if (e != cudaSuccess) return e;
  mlp_backward_fused_kernel<<<cdiv_i(B, TS), MT, smem, st>>>(net, params, ws, grad, B, with_rowsum, bwd_w_floats(net));
  return cudaGetLastError();
}

}  // namespace dino
