/* dino_b200 - C ABI of the B200-native DINO derivative-informed (H1) training step.
 *
 * The reference (tomoleary/dino) has no FFI layer: its hot path is TensorFlow 2.7 executing the
 * Keras graph that dino/construction builds.  Each entry point below replaces one piece of that
 * graph execution and cites the reference call site it stands in for.  The Python host layer
 * (dino_b200/model.py, a duck-typed Keras Model) binds these with ctypes; see INTEGRATION.md.
 *
 * Conventions
 *   - plain C: pointers, sizes, no torch / C++ types.  Every pointer named d_* / listed as
 *     "device" is a CUDA device pointer owned by the caller.  The library never allocates or
 *     frees device memory: the caller supplies one workspace (dino_plan_workspace_bytes).
 *   - all tensors fp32, row-major, dense (the one exception: DinoBatch.ld_J).  Kernels are launched on
 *     `stream` (a cudaStream_t passed as void*), asynchronously; no entry point synchronises the host.
 *     dino_h1_fwd_bwd of a tensor-core plan also uses one plan-owned side stream, forked from and joined
 *     back into `stream` by events inside the call (capturable in a CUDA graph; DINO_B200_NO_SIDE_STREAM=1
 *     keeps everything on `stream`): work enqueued on `stream` after the call sees all of its results.
 *   - return 0 on success, non-zero on error; dino_last_error() gives a thread-local message.
 *   - a plan is not thread-safe; use one plan per (device, network, host thread).
 *   - built for sm_100a only; there is no CPU or other-architecture fallback.
 */
#ifndef DINO_B200_H
#define DINO_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct DinoPlan DinoPlan;
typedef void* dino_stream_t; /* cudaStream_t */

enum { DINO_ACT_LINEAR = 0, DINO_ACT_SOFTPLUS = 1 };
/* which second model output the network is equipped with (jacobianConstruction.py:102,128,155) */
enum { DINO_JAC_NONE = 0, DINO_JAC_FULL = 1, DINO_JAC_OUT_REDUCED = 2, DINO_JAC_SKETCHED = 3 };
/* arithmetic of the contractions (Jacobian chain, data gradients, weight gradients):
 *   FP32    SIMT FFMA, no tensor cores                                   parity bar 1e-4 (reference: Keras floatx fp32)
 *   TF32    tcgen05 kind::tf32, one MMA per product (10-bit mantissas)   reported variant, bars 5e-3 / 1e-2
 *   FP32X3  tcgen05 kind::tf32, three MMAs on hi/lo operand splits       parity bar 1e-4; the default of the host layer */
enum { DINO_PREC_FP32 = 0, DINO_PREC_TF32 = 1, DINO_PREC_FP32X3 = 2 };

/* One Keras Dense layer: y = act(x W + b) [+ input of layer skip_from]; W is [n_in, n_out]
 * (Keras kernel layout).  Mirrors tf.keras.layers.Dense as used at neuralNetworks.py:34-61,
 * 83-87 and the low-rank residual block of neuralNetworks.py:93-117,158-160 (skip_from). */
typedef struct {
  int32_t n_out;
  int32_t act;       /* DINO_ACT_* */
  int32_t has_bias;  /* 0/1 */
  int32_t trainable; /* 0/1 : frozen layers get no gradient (neuralNetworks.py:66-71) */
  int32_t skip_from; /* -1, or index j<=this of the layer whose INPUT is added to this output */
} DinoLayerSpec;

typedef struct {
  int32_t d_in;      /* network input width (d_M, or r_in when trained in reduced coordinates) */
  int32_t n_layers;
  const DinoLayerSpec* layers;
  int32_t jac_mode;  /* DINO_JAC_* */
  int32_t seed_dim;  /* rows of the Jacobian output: d_out (FULL), r_out (OUT_REDUCED), k (SKETCHED) */
  int32_t max_batch; /* largest minibatch the workspace is sized for */
  int32_t precision; /* DINO_PREC_* */
} DinoNetSpec;

/* One minibatch, addressed inside device-resident datasets.  Sample b of the batch is dataset
 * row idx[b] (idx == NULL: rows 0..B-1).  Mirrors the feed of Keras fit/evaluate
 * (trainingUtilities.py:105-122): inputs [m] or [m,U,V]; targets [q, J] or [q, diag(sigma)]. */
typedef struct {
  const float* m;    /* [n, d_in] */
  const float* q;    /* [n, d_out] */
  const float* J;    /* [n, seed_dim, d_in]  (SKETCHED: [n, k, k]); NULL when jac_mode NONE */
  const float* q_sqnorm; /* [n] sum_j q^2      (dino_sample_sqnorms) */
  const float* J_sqnorm; /* [n] ||J||_F^2      (dino_sample_sqnorms) */
  const int32_t* idx;    /* [B] or NULL */
  const float* U;    /* SKETCHED: [n, d_out, k] (jacobianConstruction.py:115) else NULL */
  const float* V;    /* SKETCHED: [n, d_in, k]  (jacobianConstruction.py:116) else NULL */
  const float* seed; /* OUT_REDUCED: Phi^T [seed_dim, d_out] (jacobianConstruction.py:181) else NULL */
  /* optional (zero / NULL: dense rows, not given).  With all three of ld_J % 4 == 0, J_row_sqnorm and a plan whose first
   * layer qualifies (FULL Jacobian of an unreduced input, d_in >= 1024, tensor-core precision), dino_h1_fwd_bwd never
   * forms J_hat [B, d_out, d_in]: see DESIGN.md, "first layer without its Jacobian". */
  int64_t ld_J;                /* row pitch of J in floats (>= d_in; padding columns must hold zeros) */
  const float* J_row_sqnorm;   /* [n * seed_dim] squared norm of every row of J (dino_sample_sqnorms over the rows) */
  int64_t n_samples;           /* rows n of the dataset arrays idx addresses (0: unknown) */
} DinoBatch;

/* statistics a step accumulates (+=) into the caller's float[DINO_STAT_SLOTS] device array;
 * Keras' loss and the four metrics of trainingUtilities.py:25-51 derive from them:
 *   l2 loss = L2_SUM/COUNT, h1 loss = H1_SUM/COUNT,
 *   l2_accuracy(q) = 1-sqrt(L2_SUM/COUNT), f_accuracy_matrix(q) = 1-sqrt(EQ_SUM/NQ_SUM),
 *   l2_accuracy(J) = 1-sqrt(JROW_SUM/JROWS), f_accuracy_matrix(J) = 1-sqrt(H1_SUM/COUNT) */
enum { DINO_STAT_L2_SUM = 0, DINO_STAT_H1_SUM = 1, DINO_STAT_EQ_SUM = 2, DINO_STAT_NQ_SUM = 3,
       DINO_STAT_JROW_SUM = 4, DINO_STAT_COUNT = 5, DINO_STAT_JROWS = 6, DINO_STAT_SLOTS = 8 };

/* Build the execution plan of a network.  Replaces the Keras functional-graph construction of
 * projected_dense / generic_dense / projected_resnet (neuralNetworks.py:21,77,122) plus
 * equip_model_with_{full,output_reduced,sketched}_jacobian (jacobianConstruction.py:102-185). */
int dino_plan_create(const DinoNetSpec* spec, int device, DinoPlan** out);
void dino_plan_destroy(DinoPlan* plan);

/* Flat parameter arena: layers in graph order, kernel [n_in,n_out] row-major then bias [n_out],
 * each block padded to a multiple of 4 floats.  This is the order of layer.get_weights()
 * (jacobianTrainingDriver.py:327-328).  bias_off[l] = -1 for use_bias=False layers. */
int64_t dino_plan_param_count(const DinoPlan* plan);
int dino_plan_layer_offsets(const DinoPlan* plan, int64_t* kernel_off, int64_t* bias_off);

/* Workspace (activations, chain intermediates, split-K partials): size query, then bind. */
size_t dino_plan_workspace_bytes(const DinoPlan* plan);
int dino_plan_bind_workspace(DinoPlan* plan, void* d_workspace, size_t bytes, dino_stream_t stream);

/* q_hat = net(m): Model.predict (inferenceOracle.py:114) / first model output. d_q_out [B,d_out]. */
int dino_forward(DinoPlan* plan, const float* d_params, const DinoBatch* batch, int32_t B,
                 float* d_q_out, dino_stream_t stream);

/* Second model output: J_hat [B,seed_dim,d_in] (FULL: tape.batch_jacobian, jacobianConstruction.py
 * :145-149; OUT_REDUCED: Phi^T J_hat, :181) or U^T J_hat V [B,k,k] (SKETCHED, :119-123).
 * Also the array form of InferenceOracle.J (inferenceOracle.py:127-139). d_q_out may be NULL. */
int dino_jacobian(DinoPlan* plan, const float* d_params, const DinoBatch* batch, int32_t B,
                  float* d_q_out, float* d_J_out, dino_stream_t stream);

/* Jacobian-vector product J_hat(m) mhat [B, d_out] by one forward-mode sweep through the network (the array form of
 * InferenceOracle.Jvp, inferenceOracle.py:96-98,116-125, which contracts the full batch Jacobian with mhat).
 * d_mhat is [mhat_rows, d_in] dense with mhat_rows = 1 (one direction for all samples) or B.  The transposed product
 * J_hat^T qhat (InferenceOracle.Jtvp, :79-82,105-114) is dino_jacobian of a DINO_JAC_OUT_REDUCED plan with
 * seed_dim = 1 and batch.seed = qhat: a reverse chain of a single row. */
int dino_jvp(DinoPlan* plan, const float* d_params, const DinoBatch* batch, int32_t B, const float* d_mhat,
             int32_t mhat_rows, float* d_out, dino_stream_t stream);

/* One training step minus the optimiser: forward, Jacobian, w_l2*normalized_mse +
 * w_h1*normalized_mse_matrix, and d loss / d theta by the double backward.  Replaces what
 * session.run(train_op) computes inside network.fit (trainingUtilities.py:135-138).
 * Losses are scaled by inv_global_batch (= 1/B on one GPU; 1/B_global when the batch is sharded
 * over ranks so that a SUM all-reduce of d_grad gives the global-mean gradient).
 * d_grad [param_count] is overwritten; d_stats [DINO_STAT_SLOTS] is accumulated into. */
int dino_h1_fwd_bwd(DinoPlan* plan, const float* d_params, const DinoBatch* batch, int32_t B,
                    float w_l2, float w_h1, float inv_global_batch, float* d_grad, float* d_stats,
                    dino_stream_t stream);

/* Forward + losses/metrics only: one batch of Model.evaluate (trainingUtilities.py:125,169). */
int dino_h1_eval(DinoPlan* plan, const float* d_params, const DinoBatch* batch, int32_t B,
                 float* d_stats, dino_stream_t stream);

/* Keras Adam (trainingUtilities.py:90; TF ResourceApplyAdam): with t the 1-based step,
 * lr_t = lr*sqrt(1-b2^t)/(1-b1^t) is computed on the host in double precision;
 * m = b1 m+(1-b1) g; v = b2 v+(1-b2) g^2; p -= lr_t m/(sqrt(v)+eps).  Frozen parameters have
 * g == 0 and m == v == 0 and therefore do not move. */
int dino_adam(float* d_params, const float* d_grad, float* d_m, float* d_v, int64_t count,
              float lr, float beta1, float beta2, float eps, int64_t t, dino_stream_t stream);
/* Device-side cursors, so that a whole training step can be captured once in a CUDA graph and replayed
 * without host-supplied arguments (the Keras fit loop, trainingUtilities.py:135, feeds slices of a
 * shuffled index array; here that slicing happens on the device):
 *   dino_stage_batch : d_idx_stage[i] = d_idx_all[*d_pos + i], i < B        (this step's sample indices)
 *   dino_adam_dev    : dino_adam with t = *d_steps_done + 1 read on the device
 *   dino_roll_stats  : d_stats_hist[*d_step] = d_stats_stage; d_stats_stage = 0; *d_pos += advance;
 *                      *d_step += 1; *d_steps_done += 1                     (end of step) */
int dino_stage_batch(const int32_t* d_idx_all, const int32_t* d_pos, int32_t* d_idx_stage, int32_t B,
                     dino_stream_t stream);
int dino_adam_dev(float* d_params, const float* d_grad, float* d_m, float* d_v, int64_t count, float lr,
                  float beta1, float beta2, float eps, const int64_t* d_steps_done, dino_stream_t stream);
int dino_roll_stats(float* d_stats_stage, float* d_stats_hist, int32_t* d_pos, int32_t advance, int32_t* d_step,
                    int64_t* d_steps_done, dino_stream_t stream);
/* Keras SGD (trainingUtilities.py:92): p -= lr g. */
int dino_sgd(float* d_params, const float* d_grad, int64_t count, float lr, dino_stream_t stream);

/* out[i] = sum_j x[i, j]^2 over n rows of length len: the denominators of
 * normalized_mse / normalized_mse_matrix (trainingUtilities.py:28,36), computed once per
 * dataset because the targets do not change during fit. */
int dino_sample_sqnorms(const float* d_x, int64_t n, int64_t len, float* d_out, dino_stream_t stream);

/* No counterpart in the reference.  q_hat = net(m) like dino_forward, for callers that have no
 * device memory to hand over as a workspace and are not in a position to obtain any: code that
 * runs between the nodes of somebody else's solver loop (e.g. the inverse-problem iterations
 * that drive InferenceOracle-style evaluations) and owns nothing but its input and output
 * buffers; processes in which another framework's caching pool has already reserved the whole
 * HBM, so a cudaMalloc here would fail; and latency-critical loops in which the implicit
 * device-wide synchronisation of cudaMalloc/cudaFree is not acceptable.  The call performs no
 * cudaMalloc/cudaFree and calls no allocating library routine: activations live in a fixed
 * 3 MiB array inside the library image (held by every GPU that loads the library until the
 * process exits).  Limits: 1 <= B <= DINO_NOALLOC_MAX_BATCH, every layer width <=
 * DINO_NOALLOC_MAX_WIDTH (d_in is not limited); larger problems are rejected with an error and
 * must use dino_forward.  `spec` is the same description dino_plan_create takes (jac_mode,
 * seed_dim, max_batch, precision are ignored); d_params is the flat arena in the layout of
 * dino_plan_layer_offsets; d_m is [B, d_in] dense, d_q_out [B, d_out] dense.
 * The first call on each device looks up the array's device address (cudaGetSymbolAddress) with
 * that device current; make that first call before starting any stream capture.  All calls on a
 * device share the one array: two calls whose kernels may overlap on the GPU (different streams,
 * or different host threads) race on it - serialise them.  Nothing in the array needs to be zero
 * on entry and its contents after a call are unspecified. */
#define DINO_NOALLOC_MAX_BATCH 256
#define DINO_NOALLOC_MAX_WIDTH 1024
int dino_forward_noalloc(const DinoNetSpec* spec, const float* d_params, const float* d_m, int32_t B,
                         float* d_q_out, dino_stream_t stream);

/* The contraction primitive of the Jacobian chain on its own: C[M,N] = A[M,K] * Bt[N,K]^T, fp32 in and
 * out (what a TF MatMul inside tape.batch_jacobian computes, jacobianConstruction.py:149).  engine =
 * DINO_PREC_FP32 (SIMT FFMA) or DINO_PREC_TF32 / DINO_PREC_FP32X3 (tcgen05 + TMA; need lda, ldb, ldc multiples of 4 and
 * 16-byte aligned bases).  Used by the parity tests and the kernel micro-benchmarks. */
int dino_gemm_nt(int engine, const float* d_A, int64_t lda, const float* d_Bt, int64_t ldb, float* d_C,
                 int64_t ldc, int32_t M, int32_t N, int32_t K, dino_stream_t stream);

/* The split-K weight-gradient contraction on its own: C[n_in, n_out] += XT[n_in, rows] YT[n_out, rows]^T, i.e.
 * dW += X^T Y with both factors given TRANSPOSED (reduction index contiguous), as the fused chain kernel
 * stores them (what the second-order graph of network.fit reduces over the batch, trainingUtilities.py:135).
 * engine as above; the TF32 engine needs n_in, n_out <= 128, ldx, ldy multiples of 4, 16-byte aligned bases. */
int dino_gemm_nt_acc(int engine, const float* d_XT, int64_t ldx, const float* d_YT, int64_t ldy, float* d_C,
                     int32_t rows, int32_t n_in, int32_t n_out, dino_stream_t stream);

/* The same contraction with the factors given ROW-major, as the chain keeps them: C[n_in, n_out] += X[rows, n_in]^T
 * Y[rows, n_out] (pitches ldx, ldy floats, multiples of 4).  The TF32 engine reads them as MN-major tensor-core
 * operands (no transposed copies).  Same reference line as dino_gemm_nt_acc (trainingUtilities.py:135). */
int dino_gemm_tn_acc(int engine, const float* d_X, int64_t ldx, const float* d_Y, int64_t ldy, float* d_C,
                     int32_t rows, int32_t n_in, int32_t n_out, dino_stream_t stream);

/* Measurement aid: time the Jacobian-chain stage of dino_h1_fwd_bwd with CUDA events on the launching stream: the
 * fused chain kernel, or - for networks wider than 128 - the chain contractions of the layer-by-layer path (forward
 * chain, H1 loss kernel, data-gradient chain).  enable = 1 starts (totals reset), enable = 0 stops; the running totals
 * are returned.  Use with eager steps only (events are not recorded inside a replayed graph).  No counterpart in the
 * reference. */
int dino_plan_profile(DinoPlan* plan, int enable, double* out_ms_sum, int64_t* out_count);

/* Debug aid: %globaltimer stamps [step][phase] of CTA 0 of the fused chain kernel's last launch (all zero unless
 * the library was built with -DDINO_CF_TRACE).  Synchronises the device.  No counterpart in the reference. */
int dino_debug_chain_trace(uint64_t* out_64x8);

/* number of kernels this library has launched in the calling process (bench.py gpu_launches) */
int64_t dino_launch_count(void);

const char* dino_last_error(void);
const char* dino_version(void);

#ifdef __cplusplus
}
#endif
#endif /* DINO_B200_H */
