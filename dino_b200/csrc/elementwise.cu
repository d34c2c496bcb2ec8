// Bandwidth-bound kernels of the H1 step: gathers, the two loss kernels (which also emit the
// loss gradients in place), the activation-derivative merge of the MLP backward, the sketch
// open/close contractions and the fused Keras-form Adam / SGD updates.
#include "common.cuh"

namespace dino {

static inline int cdiv(long long a, long long b) { return (int)((a + b - 1) / b); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ---------------------------------------------------------------------------------------------
__global__ void gather_rows_kernel(const float* __restrict__ src, const int* __restrict__ idx, int rows,
                                   int cols, long long ld_src, float* __restrict__ dst, long long ld_dst) {
  const long long total = (long long)rows * cols;
  for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total;
       e += (long long)gridDim.x * blockDim.x) {
    const int r = (int)(e / cols), c = (int)(e % cols);
    const long long sr = idx ? idx[r] : r;
    dst[r * ld_dst + c] = __ldg(src + sr * ld_src + c);
  }
}
cudaError_t launch_gather_rows(const float* src, const int* idx, int rows, int cols, long long ld_src,
                               float* dst, long long ld_dst, cudaStream_t st) {
  if (rows <= 0 || cols <= 0) return cudaSuccess;
  const long long total = (long long)rows * cols;
  gather_rows_kernel<<<min(cdiv(total, 256), 148 * 8), 256, 0, st>>>(src, idx, rows, cols, ld_src, dst, ld_dst);
  return cudaGetLastError();
}

cudaError_t launch_copy2d(const float* src, long long ld_src, float* dst, long long ld_dst, int rows,
                          int cols, cudaStream_t st) {
  return launch_gather_rows(src, nullptr, rows, cols, ld_src, dst, ld_dst, st);
}

__global__ void fill_identity_kernel(float* dst, int n, long long ld) {
  const long long total = (long long)n * ld;
  for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total;
       e += (long long)gridDim.x * blockDim.x) {
    const long long r = e / ld, c = e % ld;
    dst[e] = (r == c) ? 1.f : 0.f;
  }
}
cudaError_t launch_fill_identity(float* dst, int n, long long ld, cudaStream_t st) {
  fill_identity_kernel<<<min(cdiv((long long)n * ld, 256), 148 * 8), 256, 0, st>>>(dst, n, ld);
  return cudaGetLastError();
}

// out[(b*sdim+i), n] = seed[b*sample_stride + i*ld_seed + n] * s[b, n]
__global__ void scale_seed_kernel(const float* __restrict__ seed, long long sample_stride, long long ld_seed,
                                  const float* __restrict__ s, long long ld_s, int B, int sdim, int n,
                                  float* __restrict__ out, long long ld_out, int round_out) {
  const long long total = (long long)B * sdim * n;
  for (long long e = blockIdx.x * (long long)blockDim.x + threadIdx.x; e < total;
       e += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(e % n);
    const long long row = e / n;
    const int i = (int)(row % sdim);
    const long long b = row / sdim;
    float v = __ldg(seed + b * sample_stride + i * ld_seed + c);
    if (s) v *= s[b * ld_s + c];
    out[row * ld_out + c] = round_out ? round_tf32_f(v) : v;
  }
}
cudaError_t launch_scale_seed(const float* seed, long long seed_sample_stride, long long ld_seed,
                              const float* s, long long ld_s, int B, int sdim, int n, float* out,
                              long long ld_out, cudaStream_t st, int round_out) {
  const long long total = (long long)B * sdim * n;
  if (total <= 0) return cudaSuccess;
  scale_seed_kernel<<<min(cdiv(total, 256), 148 * 8), 256, 0, st>>>(seed, seed_sample_stride, ld_seed, s, ld_s,
                                                                  B, sdim, n, out, ld_out, round_out);
  return cudaGetLastError();
}

// dst[(b*cols + j), r] = src[idx[b]][r][j]      (U [*, d_Q, k] -> U^T rows, one sample per z-block)
__global__ void transpose_batched_kernel(const float* __restrict__ src, int rows, int cols,
                                         float* __restrict__ dst, long long ld_dst, const int* __restrict__ idx) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z;
  const long long sb = idx ? idx[b] : b;
  const float* s = src + sb * (long long)rows * cols;
  const int r0 = blockIdx.y * 32, c0 = blockIdx.x * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int r = r0 + i, c = c0 + threadIdx.x;
    tile[i][threadIdx.x] = (r < rows && c < cols) ? __ldg(s + (long long)r * cols + c) : 0.f;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int c = c0 + i, r = r0 + threadIdx.x;
    if (c < cols && r < rows) dst[((long long)b * cols + c) * ld_dst + r] = tile[threadIdx.x][i];
  }
}
cudaError_t launch_transpose_batched(const float* src, int B, int rows, int cols, float* dst,
                                     long long ld_dst, const int* idx, cudaStream_t st) {
  if (B <= 0) return cudaSuccess;
  dim3 grid(cdiv(cols, 32), cdiv(rows, 32), B), block(32, 8);
  transpose_batched_kernel<<<grid, block, 0, st>>>(src, rows, cols, dst, ld_dst, idx);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// normalized_mse on q (trainingUtilities.py:25-29): one warp per sample.
__global__ void loss_q_kernel(const float* __restrict__ qhat, long long ld_qh, const float* __restrict__ q,
                              const float* __restrict__ qn2, const int* __restrict__ idx, int B, int n,
                              float coef, float* __restrict__ dq, long long ld_dq, float* __restrict__ stats) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= B) return;
  const long long sb = idx ? idx[warp] : warp;
  const float* qt = q + sb * n;
  const float* qh = qhat + (long long)warp * ld_qh;
  float e = 0.f;
  for (int c = lane; c < n; c += 32) {
    const float d = qh[c] - __ldg(qt + c);
    e = fmaf(d, d, e);
  }
  e = warp_sum(e);
  const float nq = __ldg(qn2 + sb);
  if (dq != nullptr) {
    const float g = coef / nq;
    for (int c = lane; c < n; c += 32) dq[(long long)warp * ld_dq + c] = g * (qh[c] - __ldg(qt + c));
  }
  if (lane == 0) {
    atomicAdd(stats + ST_L2_SUM, e / nq);
    atomicAdd(stats + ST_EQ_SUM, e);
    atomicAdd(stats + ST_NQ_SUM, nq);
    atomicAdd(stats + ST_COUNT, 1.f);
  }
}
cudaError_t launch_loss_q(const float* qhat, long long ld_qh, const float* q, const float* qn2, const int* idx,
                          int B, int n, float coef, float* dq, long long ld_dq, float* stats, cudaStream_t st) {
  if (B <= 0) return cudaSuccess;
  loss_q_kernel<<<cdiv((long long)B * 32, 128), 128, 0, st>>>(qhat, ld_qh, q, qn2, idx, B, n, coef, dq, ld_dq, stats);
  return cudaGetLastError();
}

// normalized_mse_matrix on the Jacobian (trainingUtilities.py:31-37): one warp per Jacobian row.
// Reads the prediction row and the target row once; overwrites the prediction with
// d loss / d J_hat = coef * (J_hat - J) / ||J_b||_F^2 when write_grad.
__global__ void __launch_bounds__(256) loss_j_kernel(float* __restrict__ jhat, long long ld_jh,
                                                     const float* __restrict__ jt, const float* __restrict__ jn2,
                                                     const int* __restrict__ idx, int B, int sdim, int n, long long ld_jt,
                                                     float coef, int write_grad, float* __restrict__ stats) {
  __shared__ float red[3][8];
  const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long rows = (long long)B * sdim;
  float h1 = 0.f, rr = 0.f, cnt = 0.f;
  for (long long row = (long long)blockIdx.x * 8 + wib; row < rows; row += (long long)gridDim.x * 8) {
    const int b = (int)(row / sdim), i = (int)(row % sdim);
    const long long sb = idx ? idx[b] : b;
    const float* t = jt + (sb * sdim + i) * ld_jt;
    float* p = jhat + row * ld_jh;
    const float nj = __ldg(jn2 + sb);
    const float g = coef / nj;
    float e = 0.f, nr = 0.f;
    int c = lane;
    // four independent 128-byte requests per operand in flight per warp (rows of the full Jacobian are 17 KB long and
    // not 16-byte aligned, so the loads stay scalar; what was missing is memory-level parallelism)
    for (; c + 96 < n; c += 128) {
      float tv[4], pv[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) { tv[u] = __ldg(t + c + 32 * u); pv[u] = p[c + 32 * u]; }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const float d = pv[u] - tv[u];
        e = fmaf(d, d, e);
        nr = fmaf(tv[u], tv[u], nr);
        if (write_grad) p[c + 32 * u] = g * d;
      }
    }
    for (; c < n; c += 32) {
      const float tv = __ldg(t + c);
      const float d = p[c] - tv;
      e = fmaf(d, d, e);
      nr = fmaf(tv, tv, nr);
      if (write_grad) p[c] = g * d;
    }
    e = warp_sum(e);
    nr = warp_sum(nr);
    h1 += e / nj;
    rr += e / nr;
    cnt += 1.f;
  }
  if (lane == 0) { red[0][wib] = h1; red[1][wib] = rr; red[2][wib] = cnt; }
  __syncthreads();
  if (threadIdx.x < 3) {
    float v = 0.f;
    for (int w = 0; w < 8; ++w) v += red[threadIdx.x][w];
    const int slot = threadIdx.x == 0 ? ST_H1_SUM : (threadIdx.x == 1 ? ST_JROW_SUM : ST_JROWS);
    atomicAdd(stats + slot, v);
  }
}
cudaError_t launch_loss_j(float* jhat, long long ld_jh, const float* jt, long long ld_jt, const float* jn2, const int* idx, int B,
                          int sdim, int n, float coef, int write_grad, float* stats, cudaStream_t st) {
  const long long rows = (long long)B * sdim;
  if (rows <= 0) return cudaSuccess;
  loss_j_kernel<<<min(cdiv(rows, 8), 148 * 16), 256, 0, st>>>(jhat, ld_jh, jt, jn2, idx, B, sdim, n, ld_jt > 0 ? ld_jt : n, coef,
                                                            write_grad, stats);
  return cudaGetLastError();
}

// Weight gradient of a linear top layer under the identity seed (FULL Jacobian): the chain state above it is I for every
// sample, so dW[k, i] += sum_b dR[b*s + i, k]: a sum over samples, no contraction
__global__ void __launch_bounds__(256) identity_seed_wgrad_kernel(const float* __restrict__ dR, long long ld, int B, int sdim, int n_in,
                                                                  float* __restrict__ dW) {
  __shared__ float tile[32][33];
  const int i0 = blockIdx.x * 32, k0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;     // 8 rows of 32
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (int b = blockIdx.z; b < B; b += gridDim.z)
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = i0 + ty + 8 * u, k = k0 + tx;
      if (i < sdim && k < n_in) acc[u] += dR[((long long)b * sdim + i) * ld + k];
    }
#pragma unroll
  for (int u = 0; u < 4; ++u) tile[ty + 8 * u][tx] = acc[u];
  __syncthreads();
#pragma unroll
  for (int u = 0; u < 4; ++u) {
    const int k = k0 + ty + 8 * u, i = i0 + tx;
    if (i < sdim && k < n_in) atomicAdd(dW + (long long)k * sdim + i, tile[tx][ty + 8 * u]);
  }
}
cudaError_t launch_identity_seed_wgrad(const float* dR, long long ld, int B, int sdim, int n_in, float* dW, cudaStream_t st) {
  if (B <= 0) return cudaSuccess;
  const int gz = B < 8 ? B : 8;
  identity_seed_wgrad_kernel<<<dim3(cdiv(sdim, 32), cdiv(n_in, 32), gz), 256, 0, st>>>(dR, ld, B, sdim, n_in, dW);
  return cudaGetLastError();
}

// z = act(acc + bias) [+ add_post], s_out = act'(pre-activation): the epilogue of a network GEMM whose reduction was split over
// CTAs (plan.cu: small-batch layers with a long reduction), in place on the accumulated sums
__global__ void __launch_bounds__(256) bias_act_kernel(float* __restrict__ z, long long ld, const float* __restrict__ bias, int B, int n,
                                                       int act, float* __restrict__ s_out, long long ld_s,
                                                       const float* __restrict__ add_post, long long ld_apo) {
  const long long total = (long long)B * n;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
    const int b = (int)(i / n), c = (int)(i % n);
    float v = z[(long long)b * ld + c];
    if (bias) v += __ldg(bias + c);
    if (act) {
      const float e = expf(-fabsf(v)), r1 = 1.f / (1.f + e);
      if (s_out) s_out[(long long)b * ld_s + c] = v >= 0.f ? r1 : e * r1;
      v = fmaxf(v, 0.f) + log1pf(e);
    }
    if (add_post) v += add_post[(long long)b * ld_apo + c];
    z[(long long)b * ld + c] = v;
  }
}
cudaError_t launch_bias_act(float* z, long long ld, const float* bias, int B, int n, int act, float* s_out, long long ld_s,
                            const float* add_post, long long ld_apo, cudaStream_t st) {
  const long long total = (long long)B * n;
  if (total <= 0) return cudaSuccess;
  bias_act_kernel<<<min(cdiv(total, 256), 148 * 8), 256, 0, st>>>(z, ld, bias, B, n, act, s_out, ld_s, add_post, ld_apo);
  return cudaGetLastError();
}

// ---- the first layer of a full-Jacobian network without its Jacobian (plan.cu: run_implicit_layer0) -----------------
// Qc[row, :] = c_b Q[row, :] and negc[b, :] = -c_b with c_b = coef / ||J_b||^2: the per-sample weight of the H1 term
__global__ void __launch_bounds__(256) implicit_prep_kernel(const float* __restrict__ Q, long long ld, int n, const float* __restrict__ jn2,
                                                            const int* __restrict__ idx, int B, int sdim, float coef,
                                                            float* __restrict__ Qc, float* __restrict__ negc) {
  const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long rows = (long long)B * sdim;
  for (long long row = (long long)blockIdx.x * 8 + wib; row < rows; row += (long long)gridDim.x * 8) {
    const int b = (int)(row / sdim);
    const long long sb = idx ? idx[b] : b;
    const float c = coef / __ldg(jn2 + sb);
    for (int j = lane; j < ld; j += 32) Qc[row * ld + j] = j < n ? c * Q[row * ld + j] : 0.f;
    if (row % sdim == 0)
      for (int j = lane; j < ld; j += 32) negc[(long long)b * ld + j] = -c;
  }
}
cudaError_t launch_implicit_prep(const float* Q, long long ld, int n, const float* jn2, const int* idx, int B, int sdim, float coef,
                                 float* Qc, float* negc, cudaStream_t st) {
  const long long rows = (long long)B * sdim;
  if (rows <= 0) return cudaSuccess;
  implicit_prep_kernel<<<min(cdiv(rows, 8), 148 * 16), 256, 0, st>>>(Q, ld, n, jn2, idx, B, sdim, coef, Qc, negc);
  return cudaGetLastError();
}

// ||Jhat_row - J_row||^2 = ddot / c_b - tdot + ||J_row||^2 with the two row dot products the contractions left behind
// (ddot = c_b <Q M - T, Q>_row, tdot = <T, Q>_row, T = J W, M = W^T W); the statistics loss_j_kernel would have written
__global__ void __launch_bounds__(256) implicit_loss_kernel(const float* __restrict__ ddot, const float* __restrict__ tdot,
                                                            const float* __restrict__ jn2, const float* __restrict__ jrn2,
                                                            const int* __restrict__ idx, int B, int sdim, float coef,
                                                            float* __restrict__ stats) {
  __shared__ float red[3][8];
  const int wib = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long rows = (long long)B * sdim;
  float h1 = 0.f, rr = 0.f, cnt = 0.f;
  for (long long row = (long long)blockIdx.x * 256 + threadIdx.x; row < rows; row += (long long)gridDim.x * 256) {
    const int b = (int)(row / sdim), i = (int)(row % sdim);
    const long long sb = idx ? idx[b] : b;
    const float nj = __ldg(jn2 + sb), nr = __ldg(jrn2 + sb * sdim + i);
    const float e = ddot[row] * (nj / coef) - tdot[row] + nr;
    h1 += e / nj;
    rr += e / nr;
    cnt += 1.f;
  }
  h1 = warp_sum(h1); rr = warp_sum(rr); cnt = warp_sum(cnt);
  if (lane == 0) { red[0][wib] = h1; red[1][wib] = rr; red[2][wib] = cnt; }
  __syncthreads();
  if (threadIdx.x < 3) {
    float v = 0.f;
    for (int w = 0; w < 8; ++w) v += red[threadIdx.x][w];
    const int slot = threadIdx.x == 0 ? ST_H1_SUM : (threadIdx.x == 1 ? ST_JROW_SUM : ST_JROWS);
    atomicAdd(stats + slot, v);
  }
}
cudaError_t launch_implicit_loss(const float* ddot, const float* tdot, const float* jn2, const float* jrn2, const int* idx, int B,
                                 int sdim, float coef, float* stats, cudaStream_t st) {
  const long long rows = (long long)B * sdim;
  if (rows <= 0) return cudaSuccess;
  implicit_loss_kernel<<<min(cdiv(rows, 256), 148 * 4), 256, 0, st>>>(ddot, tdot, jn2, jrn2, idx, B, sdim, coef, stats);
  return cudaGetLastError();
}

// da = dz * s + rowsum * (1 - s)   and   dbias[n] = sum_b da[b,n]
// (s == nullptr: linear layer, da = dz; rowsum == nullptr: no Jacobian term)
__global__ void __launch_bounds__(1024) mlp_da_kernel(const float* __restrict__ dz, long long ld_dz,
                                                      const float* __restrict__ s, long long ld_s,
                                                      const float* __restrict__ rowsum, long long ld_rs, int B, int n,
                                                      float* __restrict__ da, long long ld_da,
                                                      float* __restrict__ dbias) {
  __shared__ float red[32][33];
  const int c = blockIdx.x * 32 + threadIdx.x;
  float acc = 0.f;
  if (c < n) {
#pragma unroll 2
    for (int b = threadIdx.y; b < B; b += 32) {
      float v = dz[(long long)b * ld_dz + c];
      if (s != nullptr) {
        const float sv = s[(long long)b * ld_s + c];
        const float rs = rowsum != nullptr ? rowsum[(long long)b * ld_rs + c] : 0.f;
        v = fmaf(rs, 1.f - sv, v * sv);
      }
      if (da != nullptr) da[(long long)b * ld_da + c] = v;
      acc += v;
    }
  }
  red[threadIdx.y][threadIdx.x] = acc;
  __syncthreads();
  if (threadIdx.y == 0 && c < n && dbias != nullptr) {
    float v = 0.f;
#pragma unroll 8
    for (int r = 0; r < 32; ++r) v += red[r][threadIdx.x];
    dbias[c] = v;
  }
}
cudaError_t launch_mlp_da(const float* dz, long long ld_dz, const float* s, long long ld_s, const float* rowsum,
                          long long ld_rs, int B, int n, float* da, long long ld_da, float* dbias, cudaStream_t st) {
  if (n <= 0) return cudaSuccess;
  mlp_da_kernel<<<cdiv(n, 32), dim3(32, 32), 0, st>>>(dz, ld_dz, s, ld_s, rowsum, ld_rs, B, n, da, ld_da, dbias);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// Weight staging for the TMA-fed tensor-core engine (row pitches must be multiples of 16 bytes and
// both GEMM operands K-major): one launch covers every layer; blockIdx.z selects the layer.
__global__ void prep_weights_kernel(const float* __restrict__ params, float* __restrict__ ws, const PrepTable t) {
  __shared__ float tile[32][33];
  const PrepEntry e = t.e[blockIdx.z];
  const int k0 = blockIdx.y * 32, n0 = blockIdx.x * 32;
  if (k0 >= e.n_in || n0 >= e.n_out) return;
  const float* W = params + e.koff;
  const bool lo = e.wp_lo_off >= 0;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int k = k0 + i, n = n0 + threadIdx.x;
    const float w = (k < e.n_in && n < e.n_out) ? W[(long long)k * e.n_out + n] : 0.f;
    const float v = t.mode == 0 ? round_tf32_f(w) : w;
    tile[i][threadIdx.x] = v;
    if (k < e.n_in && n < e.ld_wp) {
      ws[e.wp_off + (long long)k * e.ld_wp + n] = v;
      if (lo) ws[e.wp_lo_off + (long long)k * e.ld_wp + n] = v - __uint_as_float(__float_as_uint(v) & 0xffffe000u);
    }
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int n = n0 + i, k = k0 + threadIdx.x;
    if (n < e.n_out && k < e.ld_wtp) {
      const float v = tile[threadIdx.x][i];
      ws[e.wtp_off + (long long)n * e.ld_wtp + k] = v;
      if (lo) ws[e.wtp_lo_off + (long long)n * e.ld_wtp + k] = v - __uint_as_float(__float_as_uint(v) & 0xffffe000u);
    }
  }
}
cudaError_t launch_prep_weights(const float* params, float* ws, const PrepTable& t, int max_in, int max_out,
                                cudaStream_t st) {
  if (t.n <= 0) return cudaSuccess;
  dim3 grid(cdiv(max_out, 32), cdiv(max_in, 32), t.n), block(32, 8);
  prep_weights_kernel<<<grid, block, 0, st>>>(params, ws, t);
  return cudaGetLastError();
}

// d sigma row-sum and per-sample scaling after a tensor-core data-gradient GEMM (its epilogue has no
// cross-row reduction): one CTA per (sample, 64 columns), threads own columns, so no atomics.
__global__ void __launch_bounds__(256) rowsum_scale_kernel(float* __restrict__ dq, long long ld_dq,
                                                           const float* __restrict__ q, long long ld_q,
                                                           const float* __restrict__ scale, long long ld_sc,
                                                           const float* __restrict__ add_post, long long ld_apo,
                                                           float* __restrict__ rowsum, long long ld_rs, int sdim, int n,
                                                           int write_out, int round_out) {
  __shared__ float red[8][33];
  const int c = blockIdx.x * 32 + threadIdx.x;
  const long long b = blockIdx.y;
  float acc = 0.f;
  if (c < n) {
    const float sc = scale ? scale[b * ld_sc + c] : 1.f;
#pragma unroll 4
    for (int i = threadIdx.y; i < sdim; i += 8) {
      const long long row = b * sdim + i;
      const float v = dq[row * ld_dq + c];
      acc = fmaf(v, q[row * ld_q + c], acc);
      if (write_out) {
        const float o = add_post ? fmaf(v, sc, add_post[row * ld_apo + c]) : v * sc;
        dq[row * ld_dq + c] = round_out ? round_tf32_f(o) : o;
      }
    }
  }
  red[threadIdx.y][threadIdx.x] = acc;
  __syncthreads();
  if (threadIdx.y == 0 && c < n && rowsum) {
    float v = 0.f;
#pragma unroll
    for (int r = 0; r < 8; ++r) v += red[r][threadIdx.x];
    rowsum[b * ld_rs + c] = v;
  }
}
cudaError_t launch_rowsum_scale(float* dq, long long ld_dq, const float* q, long long ld_q, const float* scale,
                                long long ld_sc, const float* add_post, long long ld_apo, float* rowsum,
                                long long ld_rs, int B, int sdim, int n, int write_out, int round_tf32, cudaStream_t st) {
  if (B <= 0 || n <= 0) return cudaSuccess;
  rowsum_scale_kernel<<<dim3(cdiv(n, 32), B), dim3(32, 8), 0, st>>>(dq, ld_dq, q, ld_q, scale, ld_sc, add_post, ld_apo,
                                                                    rowsum, ld_rs, sdim, n, write_out, round_tf32);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// Sketch closure  S[b] = R0[b] V[b]   ([k,dM] x [dM,k]) and its adjoint dR0[b] = dS[b] V[b]^T.
// One CTA per (sample, 8 rows of S); V tiles staged through shared memory.
__global__ void __launch_bounds__(256) sketch_close_kernel(const float* __restrict__ r0, long long ld_r0,
                                                           const float* __restrict__ V, const int* __restrict__ idx,
                                                           int k, int dM, float* __restrict__ S) {
  extern __shared__ float sm[];
  float* rt = sm;              // [8][64]
  const int b = blockIdx.y, i0 = blockIdx.x * 8;
  const long long sb = idx ? idx[b] : b;
  const float* Vb = V + sb * (long long)dM * k;
  const int tid = threadIdx.x;
  // thread -> (row i = tid / 32, column group j = tid % 32 + 32*t)
  const int li = tid >> 5, lane = tid & 31;
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (int d0 = 0; d0 < dM; d0 += 64) {
    for (int e = tid; e < 8 * 64; e += 256) {
      const int i = e / 64, d = e % 64;
      rt[e] = (i0 + i < k && d0 + d < dM) ? r0[((long long)b * k + i0 + i) * ld_r0 + d0 + d] : 0.f;
    }
    __syncthreads();
    const int dmax = min(64, dM - d0);
    for (int d = 0; d < dmax; ++d) {
      const float rv = rt[li * 64 + d];
      const float* vrow = Vb + (long long)(d0 + d) * k;
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const int j = lane + 32 * t;
        if (j < k) acc[t] = fmaf(rv, __ldg(vrow + j), acc[t]);
      }
    }
    __syncthreads();
  }
  if (i0 + li < k) {
#pragma unroll
    for (int t = 0; t < 4; ++t) {
      const int j = lane + 32 * t;
      if (j < k) S[((long long)b * k + i0 + li) * k + j] = acc[t];
    }
  }
}
cudaError_t launch_sketch_close(const float* r0, long long ld_r0, const float* V, const int* idx, int B, int k,
                                int dM, float* S, cudaStream_t st) {
  if (B <= 0) return cudaSuccess;
  if (k > 128) return cudaErrorInvalidValue;
  sketch_close_kernel<<<dim3(cdiv(k, 8), B), 256, 8 * 64 * sizeof(float), st>>>(r0, ld_r0, V, idx, k, dM, S);
  return cudaGetLastError();
}

__global__ void __launch_bounds__(256) sketch_open_kernel(const float* __restrict__ dS, const float* __restrict__ V,
                                                          const int* __restrict__ idx, int k, int dM,
                                                          float* __restrict__ dr0, long long ld_r0) {
  extern __shared__ float sm[];   // dS[b] : k*k
  const int b = blockIdx.y;
  const long long sb = idx ? idx[b] : b;
  const float* Vb = V + sb * (long long)dM * k;
  for (int e = threadIdx.x; e < k * k; e += 256) sm[e] = dS[(long long)b * k * k + e];
  __syncthreads();
  const int d = blockIdx.x * 256 + threadIdx.x;
  if (d >= dM) return;
  const float* vrow = Vb + (long long)d * k;
  for (int i = 0; i < k; ++i) {
    float acc = 0.f;
    for (int j = 0; j < k; ++j) acc = fmaf(sm[i * k + j], __ldg(vrow + j), acc);
    dr0[((long long)b * k + i) * ld_r0 + d] = acc;
  }
}
cudaError_t launch_sketch_open(const float* dS, const float* V, const int* idx, int B, int k, int dM, float* dr0,
                               long long ld_r0, cudaStream_t st) {
  if (B <= 0) return cudaSuccess;
  if (k > 128) return cudaErrorInvalidValue;
  const size_t smem = (size_t)k * k * sizeof(float);
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(sketch_open_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  sketch_open_kernel<<<dim3(cdiv(dM, 256), B), 256, smem, st>>>(dS, V, idx, k, dM, dr0, ld_r0);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// Keras-form Adam (TF ResourceApplyAdam): theta -= lr_t * m / (sqrt(v) + eps), lr_t carries the
// bias correction and eps sits OUTSIDE it.
__global__ void adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                            float* __restrict__ v, long long n, float lr_t, float b1, float b2, float eps) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  const long long n4 = n >> 2;
  float4* p4 = reinterpret_cast<float4*>(p);
  const float4* g4 = reinterpret_cast<const float4*>(g);
  float4* m4 = reinterpret_cast<float4*>(m);
  float4* v4 = reinterpret_cast<float4*>(v);
  const float ob1 = 1.f - b1, ob2 = 1.f - b2;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 P = p4[i], G = g4[i], M = m4[i], V = v4[i];
#define DINO_ADAM1(c)                                   \
    M.c = b1 * M.c + ob1 * G.c;                         \
    V.c = b2 * V.c + ob2 * G.c * G.c;                   \
    P.c -= lr_t * M.c / (sqrtf(V.c) + eps);
    DINO_ADAM1(x) DINO_ADAM1(y) DINO_ADAM1(z) DINO_ADAM1(w)
#undef DINO_ADAM1
    p4[i] = P; m4[i] = M; v4[i] = V;
  }
  for (long long i = (n4 << 2) + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += stride) {
    const float G = g[i];
    const float M = b1 * m[i] + ob1 * G;
    const float V = b2 * v[i] + ob2 * G * G;
    m[i] = M; v[i] = V;
    p[i] -= lr_t * M / (sqrtf(V) + eps);
  }
}
cudaError_t launch_adam(float* p, const float* g, float* m, float* v, long long n, float lr_t, float b1, float b2,
                        float eps, cudaStream_t st) {
  if (n <= 0) return cudaSuccess;
  adam_kernel<<<min(cdiv(n, 1024), 148 * 8), 256, 0, st>>>(p, g, m, v, n, lr_t, b1, b2, eps);
  return cudaGetLastError();
}

__global__ void sgd_kernel(float* __restrict__ p, const float* __restrict__ g, long long n, float lr) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    p[i] -= lr * g[i];
}
cudaError_t launch_sgd(float* p, const float* g, long long n, float lr, cudaStream_t st) {
  if (n <= 0) return cudaSuccess;
  sgd_kernel<<<min(cdiv(n, 256), 148 * 8), 256, 0, st>>>(p, g, n, lr);
  return cudaGetLastError();
}

// ---- device-side cursors: let a CUDA graph of the training step be replayed with no host arguments ----
// idx_stage[i] = idx_all[*pos + i]
__global__ void stage_batch_kernel(const int* __restrict__ idx_all, const int* __restrict__ pos, int* __restrict__ idx_stage, int B) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < B) idx_stage[i] = idx_all[*pos + i];
}
cudaError_t launch_stage_batch(const int* idx_all, const int* pos, int* idx_stage, int B, cudaStream_t st) {
  if (B <= 0) return cudaSuccess;
  stage_batch_kernel<<<cdiv(B, 256), 256, 0, st>>>(idx_all, pos, idx_stage, B);
  return cudaGetLastError();
}
// hist[*step] = stage; stage = 0; *pos += advance; *step += 1; *t += 1
__global__ void roll_stats_kernel(float* stage, float* hist, int* pos, int advance, int* step, long long* t) {
  const int i = threadIdx.x;
  const int s = *step;
  if (i < ST_SLOTS) { hist[(long long)s * ST_SLOTS + i] = stage[i]; stage[i] = 0.f; }
  __syncthreads();
  if (i == 0) { *pos += advance; *step = s + 1; *t += 1; }
}
cudaError_t launch_roll_stats(float* stage, float* hist, int* pos, int advance, int* step, long long* t, cudaStream_t st) {
  roll_stats_kernel<<<1, 32, 0, st>>>(stage, hist, pos, advance, step, t);
  return cudaGetLastError();
}
// Adam whose step number lives on the device: t = *t_done + 1
__global__ void adam_dev_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                float* __restrict__ v, long long n, float lr, float b1, float b2, float eps,
                                const long long* __restrict__ t_done) {
  // the bias-corrected step size lr sqrt(1 - b2^t) / (1 - b1^t), once per block.  1 - b^t is formed as -expm1(t log b):
  // no cancellation for small t, no drift of powf for large t; it agrees with dino_adam's host double to ~3e-7 relative
  // for every t (tests/test_gpu_parity.py pins both against the oracle's double over 1 .. 10^5 steps)
  __shared__ float lr_t_s;
  if (threadIdx.x == 0) {
    const float t = (float)(*t_done + 1);
    lr_t_s = lr * sqrtf(-expm1f(t * logf(b2))) / (-expm1f(t * logf(b1)));
  }
  __syncthreads();
  const float lr_t = lr_t_s;
  const float ob1 = 1.f - b1, ob2 = 1.f - b2;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float G = g[i];
    const float M = b1 * m[i] + ob1 * G;
    const float V = b2 * v[i] + ob2 * G * G;
    m[i] = M; v[i] = V;
    p[i] -= lr_t * M / (sqrtf(V) + eps);
  }
}
cudaError_t launch_adam_dev(float* p, const float* g, float* m, float* v, long long n, float lr, float b1, float b2,
                            float eps, const long long* t_done, cudaStream_t st) {
  if (n <= 0) return cudaSuccess;
  adam_dev_kernel<<<min(cdiv(n, 256), 148 * 4), 256, 0, st>>>(p, g, m, v, n, lr, b1, b2, eps, t_done);
  return cudaGetLastError();
}

// out[i] = sum_j x[i,j]^2 : one CTA per row (rows are Jacobians of 2.5k .. 1.7M floats, or q rows)
__global__ void __launch_bounds__(256) sqnorms_kernel(const float* __restrict__ x, long long len,
                                                      float* __restrict__ out) {
  __shared__ float red[8];
  const float* r = x + (long long)blockIdx.x * len;
  float acc = 0.f;
  for (long long j = threadIdx.x; j < len; j += 256) {
    const float v = __ldg(r + j);
    acc = fmaf(v, v, acc);
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float v = 0.f;
    for (int w = 0; w < 8; ++w) v += red[w];
    out[blockIdx.x] = v;
  }
}
cudaError_t launch_sqnorms(const float* x, long long n, long long len, float* out, cudaStream_t st) {
  if (n <= 0) return cudaSuccess;
  if (n > 2147483647LL) return cudaErrorInvalidValue;
  sqnorms_kernel<<<(unsigned)n, 256, 0, st>>>(x, len, out);
  return cudaGetLastError();
}

}  // namespace dino
