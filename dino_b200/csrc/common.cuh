// Shared declarations for the dino_b200 CUDA library (sm_100a only).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

namespace dino {

// One dense contraction  C[M,N] = A[M,K] * B[K,N]  with the fused epilogue every stage of the
// H1 step needs.  Operand element (m,k) of A lives at A[row(m)*a_sm + k*a_sk]; element (k,n) of
// B at B[krow(k)*b_sk + n*b_sn].  Rows of C are grouped into samples of `sdim` consecutive rows
// (the seed rows of one sample's Jacobian); per-sample vectors are indexed by b = m / sdim.
//
//   acc      = sum_k A(m,k) B(k,n)
//   rowsum[b,n] += acc * yref[m,n]                 (optional; the d sigma term)
//   v        = acc + bias[n]                       (optional)
//   v        = act ? softplus(v) : v ; s_out[m,n] = act ? sigmoid(pre-activation) : -
//   v        = (v + add_pre[m,n]) * scale[b,n] + add_post[m,n]      (each optional)
//   C[m,n]   = v            or   atomicAdd(C[m,n], v)  when atomic_c (split-K weight gradients)
struct GemmOp {
  const float* A; long long a_sm, a_sk; int a_mod;   // a_mod>0: row(m) = m % a_mod (shared seed)
  const float* B; long long b_sk, b_sn; int b_kmod;  // b_kmod>0: krow(k) = k % b_kmod
  int M, N, K;
  int k_per_split;                                   // K range handled by one blockIdx.z
  float* C; long long ldc; int atomic_c;
  int sdim;
  const float* bias;
  int act; float* s_out; long long ld_s;
  const float* yref; long long ld_y; float* rowsum; long long ld_rs;
  const float* add_pre; long long ld_ap;
  const float* scale; long long ld_sc;
  const float* add_post; long long ld_apo;
  int round_tf32;                                    // round the stored result to nearest TF32 (it feeds a TF32 MMA,
                                                     // which would otherwise TRUNCATE it: a systematic bias per layer)
  // A gathered per sample from a dataset (tensor-core engines only): A row m = sample m / a_rows, row m % a_rows of
  // that sample, and the sample sits at a_idx[.] (null: its own number) in an array [a_nsamp, a_rows, *] with row pitch
  // a_sm (a_sk for the MN-major operand of a weight gradient) and sample pitch a_ld2 floats.  a_rows == 0: plain A.
  const int* a_idx; int a_rows; long long a_ld2; long long a_nsamp;
  float* rowdot;                                     // rowdot[m] += sum_n v * yref[m,n]  (v as the row sums see it)
  int pre_first;                                     // add_pre joins BEFORE the row sums / rowdot see the value
  float alpha;                                       // weight gradients: C += alpha * X^T Y
  const float* B_lo;                                 // fp32x3 on the tcgen05 engines: B - trunc_tf32(B), same layout as B (staged
                                                     // weights): the converter warps then split A only
};

inline GemmOp gemm_op_zero() {
  GemmOp o;
  o.A = nullptr; o.a_sm = 0; o.a_sk = 0; o.a_mod = 0;
  o.B = nullptr; o.b_sk = 0; o.b_sn = 0; o.b_kmod = 0;
  o.M = o.N = o.K = 0; o.k_per_split = 0;
  o.C = nullptr; o.ldc = 0; o.atomic_c = 0; o.sdim = 1;
  o.bias = nullptr; o.act = 0; o.s_out = nullptr; o.ld_s = 0;
  o.yref = nullptr; o.ld_y = 0; o.rowsum = nullptr; o.ld_rs = 0;
  o.add_pre = nullptr; o.ld_ap = 0; o.scale = nullptr; o.ld_sc = 0;
  o.add_post = nullptr; o.ld_apo = 0; o.round_tf32 = 0;
  o.a_idx = nullptr; o.a_rows = 0; o.a_ld2 = 0; o.a_nsamp = 0; o.rowdot = nullptr; o.pre_first = 0; o.alpha = 1.f; o.B_lo = nullptr;
  return o;
}

// fp32 SIMT engine (exact-parity mode).  Returns a cudaError_t from the launch.
cudaError_t launch_gemm_simt(const GemmOp& op, int num_sms, cudaStream_t st);
// several independent split-K weight-gradient contractions in one launch (passed as a kernel parameter)
constexpr int GEMM_GROUP_MAX = 24;
struct GemmGroup { int n; int z_start[GEMM_GROUP_MAX + 1]; GemmOp ops[GEMM_GROUP_MAX]; };
cudaError_t launch_gemm_simt_grouped(GemmGroup& g, int num_sms, cudaStream_t st);
// the same group on the tensor cores (wgrad_tc.cu): X^T Y with MN-major TF32 UMMA operands, widths <= 128
bool wgrad_tc_supported(const GemmOp& op);
cudaError_t launch_wgrad_tc(const GemmGroup& g, int split, int num_sms, cudaStream_t st);

// Fused small-network MLP kernels (mlp_fused.cu): one CTA runs every layer for a tile of samples.
struct MlpLayerDev {
  long long koff, boff;        // parameter arena offsets (boff < 0: no bias)
  long long z_off, sg_off, rs_off, da_off;   // workspace offsets (floats); < 0 when absent
  int n_in, n_out, ld_out, act, trainable, skip_from, lands_skip;
};
constexpr int MLP_FUSED_MAX_LAYERS = 24;
struct MlpNetDev {
  int L, d_in, ld_in, lowest_trainable;
  long long z0_off;            // gathered input [B, ld_in]
  long long dz_top_off;        // dz_L = d loss / d q_hat [B, ld_dz]
  int ld_dz;
  MlpLayerDev layer[MLP_FUSED_MAX_LAYERS];
};
bool mlp_fused_supported(const MlpNetDev& net);
cudaError_t launch_mlp_forward_fused(const MlpNetDev& net, const float* params, float* ws, const float* m,
                                     const int* idx, int B, cudaStream_t st);
cudaError_t launch_mlp_backward_fused(const MlpNetDev& net, const float* params, float* ws, float* grad, int B,
                                      int with_rowsum, cudaStream_t st);
// tcgen05 TF32 engine (gemm_tc.cu): both operands K-major (a_sk == b_sk == 1), 16-byte row pitches,
// every epilogue term of GemmOp except atomic_c / split-K.  gemm_tc_supported() says whether an op qualifies.
bool gemm_tc_supported(const GemmOp& op);
cudaError_t launch_gemm_tc(const GemmOp& op, int split, cudaStream_t st);
// the same contraction on CTA pairs (gemm_tc2.cu: tcgen05 cta_group::2, 256 x 256 tiles per pair)
bool gemm_tc2_wanted(const GemmOp& op);
cudaError_t launch_gemm_tc2(const GemmOp& op, int split, cudaStream_t st);   // split: fp32x3 (three MMAs on hi/lo operand splits)

// weight staging for the TMA-fed engine: padded copy Wp[n_in][r4(n_out)] and transpose WTp[n_out][r4(n_in)]
// mode 0 (tf32): both copies rounded to nearest TF32.  mode 1 (fp32x3): exact copies, plus (when wp_lo_off >= 0) the low
// parts w - trunc_tf32(w) of both for the fused chain kernel, whose weight tiles are not split on chip
struct PrepEntry { long long koff, wp_off, wtp_off, wp_lo_off, wtp_lo_off; int n_in, n_out, ld_wp, ld_wtp; };
struct PrepTable { int n, mode; PrepEntry e[32]; };
cudaError_t launch_prep_weights(const float* params, float* ws, const PrepTable& t, int max_in, int max_out,
                                cudaStream_t st);
// rowsum[b,n] = sum_i dq[b*s+i,n]*q[b*s+i,n];  dq[row,n] = dq[row,n]*scale[b,n] + add_post[row,n]   (in place)
cudaError_t launch_rowsum_scale(float* dq, long long ld_dq, const float* q, long long ld_q, const float* scale,
                                long long ld_sc, const float* add_post, long long ld_apo, float* rowsum,
                                long long ld_rs, int B, int sdim, int n, int write_out, int round_tf32, cudaStream_t st);

// ---- elementwise / reduction kernels (elementwise.cu) ---------------------------------------
cudaError_t launch_gather_rows(const float* src, const int* idx, int rows, int cols, long long ld_src,
                               float* dst, long long ld_dst, cudaStream_t st);
cudaError_t launch_fill_identity(float* dst, int n, long long ld, cudaStream_t st);
cudaError_t launch_scale_seed(const float* seed, long long seed_sample_stride, long long ld_seed,
                              const float* s, long long ld_s, int B, int sdim, int n,
                              float* out, long long ld_out, cudaStream_t st, int round_out = 0);
cudaError_t launch_identity_seed_wgrad(const float* dR, long long ld, int B, int sdim, int n_in, float* dW, cudaStream_t st);
cudaError_t launch_transpose_batched(const float* src, int B, int rows, int cols, float* dst,
                                     long long ld_dst, const int* idx, cudaStream_t st);
cudaError_t launch_loss_q(const float* qhat, long long ld_qh, const float* q, const float* qn2,
                          const int* idx, int B, int n, float coef, float* dq, long long ld_dq,
                          float* stats, cudaStream_t st);
cudaError_t launch_loss_j(float* jhat, long long ld_jh, const float* jt, long long ld_jt, const float* jn2,
                          const int* idx, int B, int sdim, int n, float coef, int write_grad,
                          float* stats, cudaStream_t st);     // ld_jt: row pitch of the target (0: n)
cudaError_t launch_bias_act(float* z, long long ld, const float* bias, int B, int n, int act, float* s_out, long long ld_s,
                            const float* add_post, long long ld_apo, cudaStream_t st);
cudaError_t launch_implicit_prep(const float* Q, long long ld, int n, const float* jn2, const int* idx, int B, int sdim, float coef,
                                 float* Qc, float* negc, cudaStream_t st);
cudaError_t launch_implicit_loss(const float* ddot, const float* tdot, const float* jn2, const float* jrn2, const int* idx, int B,
                                 int sdim, float coef, float* stats, cudaStream_t st);
cudaError_t launch_mlp_da(const float* dz, long long ld_dz, const float* s, long long ld_s,
                          const float* rowsum, long long ld_rs, int B, int n, float* da,
                          long long ld_da, float* dbias, cudaStream_t st);
cudaError_t launch_sketch_close(const float* r0, long long ld_r0, const float* V, const int* idx,
                                int B, int k, int dM, float* S, cudaStream_t st);
cudaError_t launch_sketch_open(const float* dS, const float* V, const int* idx, int B, int k, int dM,
                               float* dr0, long long ld_r0, cudaStream_t st);
cudaError_t launch_adam(float* p, const float* g, float* m, float* v, long long n, float lr_t,
                        float b1, float b2, float eps, cudaStream_t st);
cudaError_t launch_sgd(float* p, const float* g, long long n, float lr, cudaStream_t st);
cudaError_t launch_stage_batch(const int* idx_all, const int* pos, int* idx_stage, int B, cudaStream_t st);
cudaError_t launch_roll_stats(float* stage, float* hist, int* pos, int advance, int* step, long long* t, cudaStream_t st);
cudaError_t launch_adam_dev(float* p, const float* g, float* m, float* v, long long n, float lr, float b1, float b2,
                            float eps, const long long* t_done, cudaStream_t st);
cudaError_t launch_sqnorms(const float* x, long long n, long long len, float* out, cudaStream_t st);
cudaError_t launch_copy2d(const float* src, long long ld_src, float* dst, long long ld_dst, int rows,
                          int cols, cudaStream_t st);

}  // namespace dino
#ifdef __cplusplus
#include <string>
namespace dino {
int set_last_error(const std::string& msg);   // sets dino_last_error(), returns 1   (plan.cu)
void count_launch();                           // dino_launch_count() += 1            (plan.cu)
}
#endif
namespace dino {

// Fused tcgen05 chain kernel (chain_fused.cu)
constexpr int CHAIN_FUSED_MAX_LAYERS = 16;
// q_off / dr_off: row-major chain states Q_l and dR_{l+1} [B*s, ld_out] (also the operands of the weight-gradient launch)
struct ChainLayerDev { long long q_off, sg_off, rs_off, dr_off; int n_in, n_out, ld_out, act, skip_from, lands_skip; };
// srow: chain rows reserved per sample inside a CTA tile (s rounded up to a warp when the tile has room, so a warp never straddles two samples)
struct ChainNetDev { int L, s, G, srow, d_in, ld_in, top_materialised, has_skip; long long r0_off; ChainLayerDev layer[CHAIN_FUSED_MAX_LAYERS]; };
// One contraction + epilogue of the fused kernel, in execution order (built on the host: the device code
// carries no layer logic).  Offsets are into the workspace, -1 = absent.
struct ChainStep {
  int K, N;                       // X [rows, K] times the step's weight tile -> [rows, N]
  int mode;                       // 0 plain, 1 loss (last forward step), 2 row-sum (backward through a softplus layer)
  int add_first;                  // residual: 1 -> (acc + add) * scale, 0 -> acc * scale + add
  int write_x;                    // the result is the next contraction's operand
  int sc_ld, out_ld, add_ld, q_ld, rs_ld;
  long long sc_off;               // per-sample scale vectors [B, sc_ld]
  long long out_off;              // row-major destination rows
  long long add_off;              // residual rows
  long long q_off;                // Q_l rows (row-sum operand)
  long long rs_off;               // row-sum destination [B, rs_ld]
};
struct ChainSteps { int n_fwd, n_all; ChainStep st[2 * CHAIN_FUSED_MAX_LAYERS]; };
// fwd / bwd: weight tiles per layer; out[t]: row-major destination of step t's result (TMA store from the operand tile)
struct ChainMaps { CUtensorMap fwd[CHAIN_FUSED_MAX_LAYERS]; CUtensorMap bwd[CHAIN_FUSED_MAX_LAYERS]; CUtensorMap fwd_lo[CHAIN_FUSED_MAX_LAYERS]; CUtensorMap bwd_lo[CHAIN_FUSED_MAX_LAYERS];
                   CUtensorMap out[2 * CHAIN_FUSED_MAX_LAYERS]; CUtensorMap top; };
bool chain_fused_supported(const ChainNetDev& net);
bool chain_fused_encode_maps(ChainMaps* maps, const ChainNetDev& net, const ChainSteps& steps, long long max_rows, const float* ws,
                             const long long* wp_off, const long long* ld_wp, const long long* wtp_off, const long long* ld_wtp,
                             const long long* wp_lo_off, const long long* wtp_lo_off);
void chain_fused_build_steps(ChainSteps* steps, const ChainNetDev& net);
cudaError_t launch_chain_fused(const ChainMaps& maps, const ChainNetDev& net, const ChainSteps& steps, float* ws, const float* jt, const float* jn2,
                               const int* idx, const float* seed, int B, float coef, int do_loss, int do_backward,
                               float* stats, const MlpNetDev& mlp, const float* params, const float* m_in, const float* q_tgt,
                               const float* qn2, float coef_q, int mlp_mode, int split, cudaStream_t st);

cudaError_t chain_fused_read_trace(unsigned long long* out);

// statistics slots accumulated by the loss kernels (floats)
enum { ST_L2_SUM = 0, ST_H1_SUM = 1, ST_EQ_SUM = 2, ST_NQ_SUM = 3, ST_JROW_SUM = 4, ST_COUNT = 5,
       ST_JROWS = 6, ST_SLOTS = 8 };

__device__ __forceinline__ float softplus_f(float x) {
  return fmaxf(x, 0.f) + log1pf(expf(-fabsf(x)));
}
__device__ __forceinline__ float round_tf32_f(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}
__device__ __forceinline__ float sigmoid_f(float x) {
  return 1.f / (1.f + expf(-x));
}

}  // namespace dino
