// Execution plan and C ABI of the H1 training step (see include/dino_b200.h).
//
// The step runs the per-sample Jacobian in REVERSE mode, seeded on the output side with s rows
// (identity, Phi^T or U_b^T), because s = min(d_in_seed, d_out_seed) for every configuration the
// reference trains.  With z_0 = m, a_l = z_l W_l + b_l, sg_l = act'(a_l):
//
//   MLP forward      l = 0..L-1   z_{l+1} = act(a_l) [+ z_skip]
//   chain forward    l = L-1..0   Q_l = R_{l+1} * sg_l ;  R_l = Q_l W_l^T [+ Q_skip]      R_L = seed
//   losses                        dq, G = d loss / d R_0  (in place over R_0)
//   chain backward   l = 0..L-1   dQ_l = dR_l W_l ; dW_l += dR_l^T Q_l ; rs_l = sum_rows dQ_l*Q_l ;
//                                 dR_{l+1} = dQ_l * sg_l [+ dR_skip]
//   MLP backward     l = L-1..0   da_l = dz_{l+1}*sg_l + rs_l*(1-sg_l) ; dW_l += z_l^T da_l ;
//                                 db_l = sum_b da_l ; dz_l = da_l W_l^T [+ dz_skip]
//
// (softplus'' = sg(1-sg), which is why only the post-scaling Q_l has to be kept.)  The formulas
// are checked against autograd double-backward in tests/test_pipeline_model.py.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <string>
#include <vector>

#include "../../include/dino_b200.h"
#include "common.cuh"

using namespace dino;

static thread_local std::string g_err;
static std::atomic<long long> g_launches{0};

static int fail(const std::string& msg) {
  g_err = msg;
  return 1;
}

namespace dino {
// used by translation units that hold entry points of their own (noalloc.cu)
int set_last_error(const std::string& msg) { return fail(msg); }
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }
}  // namespace dino

#define DINO_CUDA(expr)                                                                      \
  do {                                                                                       \
    cudaError_t _e = (expr);                                                                 \
    if (_e != cudaSuccess)                                                                   \
      return fail(std::string(#expr) + ": " + cudaGetErrorString(_e));                       \
  } while (0)
#define DINO_LAUNCH(expr)                                                                    \
  do {                                                                                       \
    g_launches.fetch_add(1, std::memory_order_relaxed);                                      \
    DINO_CUDA(expr);                                                                         \
  } while (0)

static inline long long r4(long long n) { return (n + 3) & ~3LL; }
// Row pitch of a workspace array (floats).  A multiple of 8 floats = 32 bytes: a row then starts on a DRAM sector, so
// the 16-byte pieces of a row never split a sector (a pitch of 100 or 4228 floats - multiples of 4 only - made every
// other row start in the middle of one, and half-sector writes are read-modify-write).  Long rows (the full-Jacobian
// output, 4225 wide) get whole 128-byte lines.
static inline long long rpitch(long long n) { return n >= 1024 ? (n + 31) & ~31LL : (n + 7) & ~7LL; }

struct LayerPlan {
  int n_in, n_out, act, has_bias, trainable, skip_from;
  long long koff, boff;      // offsets in the parameter arena (floats)
  long long z_off, sg_off, q_off, rs_off;   // workspace offsets (floats); -1 when absent
  long long wp_off, wtp_off, ld_wp, ld_wtp; // tensor-core modes: padded W [n_in][ld_wp] and W^T [n_out][ld_wtp]
  long long wp_lo_off, wtp_lo_off;          // fp32x3 + fused chain kernel: their low parts
  long long dr_off;          // dR_{l+1} = gradient wrt the chain state above this layer [B*s, ld_out]; -1 for the top layer
  long long da_off;          // da_l [B, ld_out] (fused MLP backward keeps every layer's for the grouped weight gradient)
  long long ld_out;          // rpitch(n_out)
  int lands_skip;            // index l2 of the skip layer whose source is THIS layer's input, or -1
};

struct DinoPlan {
  int device, num_sms;
  int d_in, d_out, L, jac_mode, s, max_batch, precision;
  bool tc;                   // contractions on the tcgen05 engine (tf32 / fp32x3)
  int split;                 // 1: fp32x3 (three MMAs on hi/lo operand splits)
  std::vector<LayerPlan> layers;
  long long n_params;
  bool seed_shared;          // FULL / OUT_REDUCED
  bool top_materialised;     // Q_{L-1} stored in the workspace (else it is the shared seed itself)
  bool top_identity;         // tensor-core plans, FULL Jacobian, linear top layer: identity seed, never stored: the first chain
                             // step is W_L^T (* sigma') per sample, the top weight gradient a sum over samples
  // workspace offsets (floats)
  long long off_ident, off_z0, off_r0, off_sk, off_dz[3], off_da, off_rs_begin, rs_floats;
  long long ld_in, ld_dz, ld_ident;
  long long ws_floats;
  float* ws;
  PrepTable prep;            // TF32 mode: weight staging table
  int max_in, max_out;
  MlpNetDev mlp;             // descriptor of the fused small-network MLP kernels
  bool mlp_fused;
  ChainNetDev chain;         // descriptor + weight tensor maps of the fused tcgen05 chain kernel
  ChainMaps chain_maps;
  ChainSteps chain_steps;
  bool mlp_in_chain;         // the fused chain kernel also runs the MLP forward and the output loss of its samples
  bool chain_fused;
  // first layer without its Jacobian (run_implicit_layer0): workspace of the scaled chain state, J W, the two Gram
  // matrices and the per-row dot products
  bool implicit0;
  long long off_qc, off_tb, off_negc, off_mw, off_sb, off_tdot, off_ddot, imp_zero_floats;
  // optional per-kernel timing (dino_plan_profile): CUDA events around the Jacobian-chain stage of each step
  // side stream: the chain's weight gradients run concurrently with the MLP backward (fork/join by events)
  cudaStream_t side;
  cudaEvent_t ev_fork, ev_join, ev_prep;
  bool profile;
  cudaEvent_t ev0, ev1;
  double prof_ms; long long prof_n; bool ev_pending;
};

static cudaStream_t S(dino_stream_t s) { return reinterpret_cast<cudaStream_t>(s); }

extern "C" const char* dino_last_error(void) { return g_err.c_str(); }
extern "C" const char* dino_version(void) { return "dino_b200 0.1 (sm_100a)"; }
extern "C" int64_t dino_launch_count(void) { return g_launches.load(); }

extern "C" int dino_plan_create(const DinoNetSpec* spec, int device, DinoPlan** out) {
  if (!spec || !out) return fail("dino_plan_create: null argument");
  if (spec->n_layers < 1 || spec->d_in < 1 || spec->max_batch < 1) return fail("dino_plan_create: bad sizes");
  if (spec->jac_mode < DINO_JAC_NONE || spec->jac_mode > DINO_JAC_SKETCHED) return fail("dino_plan_create: bad jac_mode");
  if (spec->precision != DINO_PREC_FP32 && spec->precision != DINO_PREC_TF32 && spec->precision != DINO_PREC_FP32X3)
    return fail("dino_plan_create: bad precision");
  int ndev = 0;
  DINO_CUDA(cudaGetDeviceCount(&ndev));
  if (device < 0 || device >= ndev) return fail("dino_plan_create: no such CUDA device");
  cudaDeviceProp prop;
  DINO_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) return fail("dino_plan_create: this library is built for sm_100a (B200) only");

  DinoPlan* p = new DinoPlan();
  p->device = device;
  p->num_sms = prop.multiProcessorCount;
  p->d_in = spec->d_in;
  p->L = spec->n_layers;
  p->jac_mode = spec->jac_mode;
  p->max_batch = spec->max_batch;
  p->precision = spec->precision;
  p->tc = spec->precision != DINO_PREC_FP32;
  p->split = spec->precision == DINO_PREC_FP32X3 ? 1 : 0;
  p->ws = nullptr;
  p->side = nullptr; p->ev_fork = p->ev_join = p->ev_prep = nullptr;
  p->profile = false; p->prof_ms = 0.0; p->prof_n = 0; p->ev_pending = false; p->ev0 = p->ev1 = nullptr;
  long long poff = 0;
  int n_in = spec->d_in;
  for (int l = 0; l < p->L; ++l) {
    const DinoLayerSpec& ls = spec->layers[l];
    LayerPlan lp;
    lp.n_in = n_in; lp.n_out = ls.n_out; lp.act = ls.act; lp.has_bias = ls.has_bias;
    lp.trainable = ls.trainable; lp.skip_from = ls.skip_from; lp.lands_skip = -1;
    if (ls.n_out < 1) { delete p; return fail("dino_plan_create: layer width < 1"); }
    if (ls.act != DINO_ACT_LINEAR && ls.act != DINO_ACT_SOFTPLUS) { delete p; return fail("dino_plan_create: bad activation"); }
    if (ls.skip_from >= 0) {
      // the low-rank residual block of neuralNetworks.py:93-117: in(softplus) -> out(linear) + block input
      if (ls.skip_from != l - 1 || l < 1 || ls.act != DINO_ACT_LINEAR || p->layers[l - 1].n_in != ls.n_out) {
        delete p;
        return fail("dino_plan_create: skip_from must name the previous layer of a two-layer residual block "
                    "with a linear, width-preserving output");
      }
    }
    lp.koff = poff; poff += r4((long long)lp.n_in * lp.n_out);
    if (lp.has_bias) { lp.boff = poff; poff += r4(lp.n_out); } else lp.boff = -1;
    lp.ld_out = rpitch(lp.n_out);
    p->layers.push_back(lp);
    n_in = ls.n_out;
  }
  for (int l = 0; l < p->L; ++l)
    if (p->layers[l].skip_from >= 0) p->layers[p->layers[l].skip_from].lands_skip = l;
  p->n_params = poff;
  p->d_out = n_in;
  if (p->jac_mode == DINO_JAC_NONE) p->s = 0;
  else if (p->jac_mode == DINO_JAC_FULL) p->s = p->d_out;
  else p->s = spec->seed_dim;
  if (p->jac_mode != DINO_JAC_NONE && p->s < 1) { delete p; return fail("dino_plan_create: seed_dim < 1"); }
  if (p->jac_mode == DINO_JAC_FULL && spec->seed_dim != 0 && spec->seed_dim != p->d_out) {
    delete p;
    return fail("dino_plan_create: FULL jacobian needs seed_dim == d_out");
  }
  if (p->jac_mode == DINO_JAC_SKETCHED && p->s > 128) { delete p; return fail("dino_plan_create: sketch rank > 128"); }
  p->seed_shared = (p->jac_mode == DINO_JAC_FULL || p->jac_mode == DINO_JAC_OUT_REDUCED);
  const LayerPlan& last = p->layers[p->L - 1];
  bool wide = p->d_in > 128;          // (narrow networks go to the fused chain kernel, which has its own top-layer handling)
  for (auto& lp : p->layers) wide = wide || lp.n_out > 128;
  p->top_identity = p->tc && wide && p->jac_mode == DINO_JAC_FULL && last.act == DINO_ACT_LINEAR && last.skip_from < 0 &&
                    last.lands_skip < 0 && getenv("DINO_B200_NO_TOP_IDENTITY") == nullptr;
  p->top_materialised = p->jac_mode != DINO_JAC_NONE && !p->top_identity &&
                        (!p->seed_shared || last.act == DINO_ACT_SOFTPLUS || last.skip_from >= 0 ||
                         p->tc /* TMA-addressable operand for the tensor-core weight gradient */);

  // ---- workspace layout ------------------------------------------------------------------------
  const long long Bm = p->max_batch, Ms = Bm * (long long)p->s;
  long long off = 0;
  auto take = [&](long long n) { long long o = off; off += (n + 31) & ~31LL; return o; };     // arrays start on 128-byte lines
  p->ld_in = rpitch(p->d_in);
  p->ld_ident = rpitch(p->d_out);
  p->off_ident = (p->jac_mode == DINO_JAC_FULL) ? take((long long)p->d_out * p->ld_ident) : -1;
  p->off_z0 = take(Bm * p->ld_in);
  long long max_w = p->ld_in;
  for (auto& lp : p->layers) {
    lp.z_off = take(Bm * lp.ld_out);
    lp.sg_off = lp.act == DINO_ACT_SOFTPLUS ? take(Bm * lp.ld_out) : -1;
    if (lp.ld_out > max_w) max_w = lp.ld_out;
  }
  p->off_rs_begin = off;
  for (auto& lp : p->layers)
    lp.rs_off = (p->jac_mode != DINO_JAC_NONE && lp.act == DINO_ACT_SOFTPLUS) ? take(Bm * lp.ld_out) : -1;
  p->rs_floats = off - p->off_rs_begin;
  p->ld_dz = max_w;
  for (int i = 0; i < 3; ++i) p->off_dz[i] = take(Bm * max_w);
  p->off_da = take(Bm * max_w);
  p->off_r0 = p->off_sk = -1;
  for (auto& lp : p->layers) { lp.q_off = -1; lp.dr_off = -1; lp.da_off = -1; }
  if (p->jac_mode != DINO_JAC_NONE) {
    for (int l = 0; l < p->L; ++l) {
      LayerPlan& lp = p->layers[l];
      if (l < p->L - 1 || p->top_materialised) lp.q_off = take(Ms * lp.ld_out);
    }
    p->off_r0 = take(Ms * p->ld_in);
    for (int l = 0; l < p->L - 1; ++l) p->layers[l].dr_off = take(Ms * p->layers[l].ld_out);
    if (p->jac_mode == DINO_JAC_SKETCHED) p->off_sk = take(Bm * (long long)p->s * p->s);
  }
  // fused small-network MLP path
  {
    MlpNetDev& n = p->mlp;
    memset(&n, 0, sizeof(n));
    n.L = p->L; n.d_in = p->d_in; n.ld_in = (int)p->ld_in; n.z0_off = p->off_z0;
    n.dz_top_off = p->off_dz[p->L % 3]; n.ld_dz = (int)p->ld_dz;
    n.lowest_trainable = p->L;
    for (int l = 0; l < p->L; ++l)
      if (p->layers[l].trainable) { n.lowest_trainable = l; break; }
    p->mlp_fused = p->L <= MLP_FUSED_MAX_LAYERS;
    if (p->mlp_fused) {
      for (int l = 0; l < p->L; ++l) {
        LayerPlan& lp = p->layers[l];
        lp.da_off = take(Bm * lp.ld_out);
        MlpLayerDev& d = n.layer[l];
        d.koff = lp.koff; d.boff = lp.boff; d.z_off = lp.z_off; d.sg_off = lp.sg_off; d.rs_off = lp.rs_off;
        d.da_off = lp.da_off; d.n_in = lp.n_in; d.n_out = lp.n_out; d.ld_out = (int)lp.ld_out;
        d.act = lp.act == DINO_ACT_SOFTPLUS; d.trainable = lp.trainable; d.skip_from = lp.skip_from;
        d.lands_skip = lp.lands_skip;
      }
      p->mlp_fused = mlp_fused_supported(n);
    }
    if (!p->mlp_fused)       // layer-by-layer backward: every softplus layer keeps its da, so that dW_l += z_l^T da_l can join the
      for (auto& lp : p->layers)   // grouped weight-gradient launch at the end of the step
        lp.da_off = lp.act == DINO_ACT_SOFTPLUS ? take(Bm * lp.ld_out) : -1;
  }
  p->prep.n = 0;
  p->max_in = p->max_out = 1;
  p->prep.mode = p->split;
  for (auto& lp : p->layers) { lp.wp_off = lp.wtp_off = lp.wp_lo_off = lp.wtp_lo_off = -1; lp.ld_wp = rpitch(lp.n_out); lp.ld_wtp = rpitch(lp.n_in); }
  if (p->tc && p->jac_mode != DINO_JAC_NONE) {
    if (p->L > 32) { delete p; return fail("dino_plan_create: the tensor-core modes support at most 32 layers"); }
    for (auto& lp : p->layers) {
      lp.wp_off = take((long long)lp.n_in * lp.ld_wp);
      lp.wtp_off = take((long long)lp.n_out * lp.ld_wtp);
      PrepEntry& e = p->prep.e[p->prep.n++];
      e.koff = lp.koff; e.wp_off = lp.wp_off; e.wtp_off = lp.wtp_off; e.wp_lo_off = e.wtp_lo_off = -1;
      e.n_in = lp.n_in; e.n_out = lp.n_out; e.ld_wp = (int)lp.ld_wp; e.ld_wtp = (int)lp.ld_wtp;
      if (lp.n_in > p->max_in) p->max_in = lp.n_in;
      if (lp.n_out > p->max_out) p->max_out = lp.n_out;
    }
  }
  p->chain_fused = false;
  p->mlp_in_chain = false;
  if (p->tc && (p->jac_mode == DINO_JAC_FULL || p->jac_mode == DINO_JAC_OUT_REDUCED) &&
      p->L <= CHAIN_FUSED_MAX_LAYERS && getenv("DINO_B200_NO_FUSED_CHAIN") == nullptr) {
    ChainNetDev& c = p->chain;
    memset(&c, 0, sizeof(c));
    c.L = p->L; c.s = p->s; c.G = p->s <= 128 ? 128 / p->s : 0; if (c.G > 16) c.G = 16;
    c.srow = (c.G * ((p->s + 31) & ~31) <= 128) ? ((p->s + 31) & ~31) : p->s;
    c.d_in = p->d_in; c.ld_in = (int)p->ld_in; c.top_materialised = p->top_materialised ? 1 : 0; c.r0_off = p->off_r0;
    for (int l = 0; l < p->L; ++l) {
      const LayerPlan& lp = p->layers[l];
      ChainLayerDev& d = c.layer[l];
      d.q_off = lp.q_off; d.sg_off = lp.sg_off; d.rs_off = lp.rs_off; d.dr_off = lp.dr_off;
      d.n_in = lp.n_in; d.n_out = lp.n_out; d.ld_out = (int)lp.ld_out; d.act = lp.act == DINO_ACT_SOFTPLUS;
      d.skip_from = lp.skip_from; d.lands_skip = lp.lands_skip;
      if (lp.skip_from >= 0) c.has_skip = 1;
    }
    p->chain_fused = chain_fused_supported(c);
    if (p->chain_fused) chain_fused_build_steps(&p->chain_steps, c);
    p->mlp_in_chain = p->chain_fused && p->mlp_fused && !c.has_skip && p->L <= MLP_FUSED_MAX_LAYERS &&
                      getenv("DINO_B200_NO_MLP_IN_CHAIN") == nullptr;
    if (p->chain_fused && p->split) {
      // the fused kernel takes its weight tiles ready-made: it needs their low parts from memory as well
      for (int l = 0; l < p->L; ++l) {
        LayerPlan& lp = p->layers[l];
        lp.wp_lo_off = take((long long)lp.n_in * lp.ld_wp);
        lp.wtp_lo_off = take((long long)lp.n_out * lp.ld_wtp);
        p->prep.e[l].wp_lo_off = lp.wp_lo_off; p->prep.e[l].wtp_lo_off = lp.wtp_lo_off;
      }
    }
  }
  if (p->tc && p->split && !p->chain_fused && p->prep.n > 0 && getenv("DINO_B200_NO_WLO") == nullptr) {
    // layer-by-layer fp32x3: the low parts of the staged weights are written once per step by prep_weights_kernel, and the
    // contraction kernels fetch them by TMA instead of splitting the weight tile of every stage on chip
    for (int l = 0; l < p->L; ++l) {
      LayerPlan& lp = p->layers[l];
      lp.wp_lo_off = take((long long)lp.n_in * lp.ld_wp);
      lp.wtp_lo_off = take((long long)lp.n_out * lp.ld_wtp);
      p->prep.e[l].wp_lo_off = lp.wp_lo_off; p->prep.e[l].wtp_lo_off = lp.wtp_lo_off;
    }
  }
  // first layer without its Jacobian: FULL Jacobian of a wide (unreduced) input through a plain softplus layer
  p->implicit0 = false;
  p->off_qc = p->off_tb = p->off_negc = p->off_mw = p->off_sb = p->off_tdot = p->off_ddot = -1; p->imp_zero_floats = 0;
  {
    const LayerPlan& l0 = p->layers[0];
    static const int min_din = getenv("DINO_B200_IMPLICIT_MIN") ? atoi(getenv("DINO_B200_IMPLICIT_MIN")) : 1024;   // read once per process
    if (p->tc && !p->chain_fused && p->jac_mode == DINO_JAC_FULL && p->L >= 2 && l0.act == DINO_ACT_SOFTPLUS &&
        l0.skip_from < 0 && l0.lands_skip < 0 && p->layers[1].skip_from < 0 && (l0.n_out & 3) == 0 && p->d_in >= min_din &&
        getenv("DINO_B200_NO_IMPLICIT0") == nullptr) {
      p->implicit0 = true;
      const long long n1 = l0.n_out;
      p->off_qc = take(Ms * l0.ld_out);
      p->off_tb = take(Ms * l0.ld_out);
      p->off_negc = take(Bm * l0.ld_out);
      p->off_mw = take(n1 * n1);                 // the four below are cleared together at the start of a step
      p->off_sb = take(n1 * n1);
      p->off_tdot = take(Ms);
      p->off_ddot = take(Ms);
      p->imp_zero_floats = off - p->off_mw;
    }
  }
  p->ws_floats = off;
  if (p->chain_fused && getenv("DINO_B200_NO_SIDE_STREAM") == nullptr) {
    DINO_CUDA(cudaSetDevice(device));
    DINO_CUDA(cudaStreamCreateWithFlags(&p->side, cudaStreamNonBlocking));
    DINO_CUDA(cudaEventCreateWithFlags(&p->ev_fork, cudaEventDisableTiming));
    DINO_CUDA(cudaEventCreateWithFlags(&p->ev_join, cudaEventDisableTiming));
    DINO_CUDA(cudaEventCreateWithFlags(&p->ev_prep, cudaEventDisableTiming));
  } else if (p->tc && p->jac_mode != DINO_JAC_NONE && getenv("DINO_B200_NO_SIDE_STREAM") == nullptr) {
    // layer-by-layer tensor-core plans: the network's backward pass (some twenty short launches on a few SMs each) runs
    // on a HIGH-PRIORITY side stream next to the grouped weight gradients of the chain, which fill the GPU for a
    // millisecond: its CTAs take the SMs the big launch frees first
    int least = 0, greatest = 0;
    DINO_CUDA(cudaSetDevice(device));
    DINO_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
    DINO_CUDA(cudaStreamCreateWithPriority(&p->side, cudaStreamNonBlocking, greatest));
    DINO_CUDA(cudaEventCreateWithFlags(&p->ev_fork, cudaEventDisableTiming));
    DINO_CUDA(cudaEventCreateWithFlags(&p->ev_join, cudaEventDisableTiming));
    DINO_CUDA(cudaEventCreateWithFlags(&p->ev_prep, cudaEventDisableTiming));
  }
  *out = p;
  return 0;
}

extern "C" void dino_plan_destroy(DinoPlan* plan) {
  if (plan && plan->ev0) { cudaEventDestroy(plan->ev0); cudaEventDestroy(plan->ev1); }
  if (plan && plan->side) { cudaEventDestroy(plan->ev_fork); cudaEventDestroy(plan->ev_join); cudaEventDestroy(plan->ev_prep); cudaStreamDestroy(plan->side); }
  delete plan;
}

extern "C" int64_t dino_plan_param_count(const DinoPlan* plan) { return plan ? plan->n_params : -1; }

extern "C" int dino_plan_layer_offsets(const DinoPlan* plan, int64_t* kernel_off, int64_t* bias_off) {
  if (!plan || !kernel_off || !bias_off) return fail("dino_plan_layer_offsets: null argument");
  for (int l = 0; l < plan->L; ++l) {
    kernel_off[l] = plan->layers[l].koff;
    bias_off[l] = plan->layers[l].boff;
  }
  return 0;
}

extern "C" size_t dino_plan_workspace_bytes(const DinoPlan* plan) {
  return plan ? (size_t)plan->ws_floats * sizeof(float) : 0;
}

extern "C" int dino_plan_bind_workspace(DinoPlan* plan, void* d_ws, size_t bytes, dino_stream_t stream) {
  if (!plan || !d_ws) return fail("dino_plan_bind_workspace: null argument");
  if (bytes < (size_t)plan->ws_floats * sizeof(float)) return fail("dino_plan_bind_workspace: workspace too small");
  if ((reinterpret_cast<uintptr_t>(d_ws) & 255) != 0) return fail("dino_plan_bind_workspace: workspace must be 256-byte aligned");
  plan->ws = static_cast<float*>(d_ws);
  if (plan->chain_fused) {
    long long wp[CHAIN_FUSED_MAX_LAYERS], lwp[CHAIN_FUSED_MAX_LAYERS], wtp[CHAIN_FUSED_MAX_LAYERS], lwtp[CHAIN_FUSED_MAX_LAYERS];
    long long wpl[CHAIN_FUSED_MAX_LAYERS], wtpl[CHAIN_FUSED_MAX_LAYERS];
    for (int l = 0; l < plan->L; ++l) {
      wp[l] = plan->layers[l].wp_off; lwp[l] = plan->layers[l].ld_wp;
      wtp[l] = plan->layers[l].wtp_off; lwtp[l] = plan->layers[l].ld_wtp;
      wpl[l] = plan->layers[l].wp_lo_off; wtpl[l] = plan->layers[l].wtp_lo_off;
    }
    if (!chain_fused_encode_maps(&plan->chain_maps, plan->chain, plan->chain_steps, (long long)plan->max_batch * plan->s, plan->ws, wp, lwp, wtp, lwtp,
                                 wpl, wtpl))
      return fail("dino_plan_bind_workspace: cuTensorMapEncodeTiled failed for the fused chain kernel");
  }
  if (plan->off_ident >= 0)
    DINO_LAUNCH(launch_fill_identity(plan->ws + plan->off_ident, plan->d_out, plan->ld_ident, S(stream)));
  return 0;
}

// ---------------------------------------------------------------------------------------------
static int check_call(const DinoPlan* p, const float* params, const DinoBatch* b, int B, bool need_targets) {
  if (!p || !params || !b) return fail("null argument");
  if (!p->ws) return fail("workspace not bound (dino_plan_bind_workspace)");
  int cur = -1;
  if (cudaGetDevice(&cur) != cudaSuccess || cur != p->device)
    return fail("the plan belongs to CUDA device " + std::to_string(p->device) + " but device " + std::to_string(cur) + " is current");
  if (B < 1 || B > p->max_batch) return fail("batch size outside [1, max_batch]");
  if (!b->m) return fail("batch.m is null");
  if (p->jac_mode == DINO_JAC_SKETCHED && (!b->U || !b->V)) return fail("sketched Jacobian needs batch.U and batch.V");
  if (p->jac_mode == DINO_JAC_OUT_REDUCED && !b->seed) return fail("output-reduced Jacobian needs batch.seed (Phi^T)");
  if (need_targets) {
    if (!b->q || !b->q_sqnorm) return fail("batch.q / batch.q_sqnorm is null");
    if (p->jac_mode != DINO_JAC_NONE && (!b->J || !b->J_sqnorm)) return fail("batch.J / batch.J_sqnorm is null");
    if (b->ld_J != 0 && (p->jac_mode == DINO_JAC_SKETCHED || p->chain_fused || b->ld_J < p->d_in))
      return fail("batch.ld_J: a row pitch is only taken for the FULL / OUT_REDUCED Jacobian targets of the layer-by-layer path, and must be >= d_in");
  }
  return 0;
}

// force_split: the network's own (small) GEMMs keep fp32-grade products in the tf32 mode too
static int gemm(const DinoPlan* p, const GemmOp& op, cudaStream_t st, bool force_split = false) {
  if (p->tc && gemm_tc_supported(op)) {
    DINO_LAUNCH(launch_gemm_tc(op, force_split ? 1 : p->split, st));
    return 0;
  }
  DINO_LAUNCH(launch_gemm_simt(op, p->num_sms, st));
  return 0;
}

static int gemm_simt_only(const DinoPlan* p, const GemmOp& op, cudaStream_t st) {
  DINO_LAUNCH(launch_gemm_simt(op, p->num_sms, st));
  return 0;
}

// One GEMM of the network itself (M = minibatch rows) on the tensor-core plans.  Three regimes:
//   tiny (a few MFLOP: the width-100 layers)            -> the FFMA engine: one short launch beats the tcgen05 pipeline's set-up
//   few row tiles (M <= 256, K >= 128)                       -> the reduction is split over CTAs (the weight-gradient kernel
//       with both factors K-major, fp32 atomics into a cleared block), bias / activation / residual in a second pass:
//       a [256, 1024] x [1024, 1024] layer has 16 output tiles for 148 SMs, a [128, 4225] x [4225, 100] one has ONE
//   otherwise                                           -> gemm_tc_kernel with its fused epilogue
// op describes the fused form (K-major A and B, bias / act / s_out / add_post).
static int mlp_gemm(DinoPlan* p, const GemmOp& op, const GemmOp& op_simt, cudaStream_t st) {
  const double flops = 2.0 * op.M * op.N * op.K;
  if (flops < 2.0e7) return gemm_simt_only(p, op_simt, st);
  static const int min_k = getenv("DINO_MLP_SPLIT_K") ? atoi(getenv("DINO_MLP_SPLIT_K")) : 128;     // read once per process
  if (op.M <= 256 && op.K >= min_k && op.C && (op.ldc & 3) == 0 && !op.add_pre && !op.scale) {
    GemmGroup g;
    g.n = 1;
    GemmOp& w = g.ops[0];
    w = gemm_op_zero();
    w.A = op.A; w.a_sm = op.a_sm; w.a_sk = 1; w.B = op.B; w.b_sn = op.b_sn; w.b_sk = 1;
    w.M = op.M; w.N = op.N; w.K = op.K; w.C = op.C; w.ldc = op.ldc; w.atomic_c = 1;
    if (wgrad_tc_supported(w)) {
      DINO_CUDA(cudaMemsetAsync(op.C, 0, ((size_t)(op.M - 1) * op.ldc + op.N) * sizeof(float), st));
      DINO_LAUNCH(launch_wgrad_tc(g, 1, p->num_sms, st));
      if (op.bias || op.act || op.add_post)
        DINO_LAUNCH(launch_bias_act(op.C, op.ldc, op.bias, op.M, op.N, op.act, op.s_out, op.ld_s, op.add_post, op.ld_apo, st));
      return 0;
    }
  }
  return gemm(p, op, st, true);
}

static int run_mlp_forward(DinoPlan* p, const float* params, const DinoBatch* b, int B, cudaStream_t st) {
  float* ws = p->ws;
  if (p->mlp_fused) {
    DINO_LAUNCH(launch_mlp_forward_fused(p->mlp, params, ws, b->m, b->idx, B, st));
    return 0;
  }
  // wide networks: the network's own GEMMs run on the tcgen05 engine too (always with fp32-grade products); they read
  // the staged weight copies, so those are written first
  const bool tc_mlp = p->tc && p->prep.n > 0 && !p->chain_fused;
  if (tc_mlp) DINO_LAUNCH(launch_prep_weights(params, ws, p->prep, p->max_in, p->max_out, st));
  DINO_LAUNCH(launch_gather_rows(b->m, b->idx, B, p->d_in, p->d_in, ws + p->off_z0, p->ld_in, st));
  const float* zin = ws + p->off_z0;
  long long ld_zin = p->ld_in;
  for (int l = 0; l < p->L; ++l) {
    const LayerPlan& lp = p->layers[l];
    GemmOp op = gemm_op_zero();
    op.A = zin; op.a_sm = ld_zin; op.a_sk = 1;
    op.B = params + lp.koff; op.b_sk = lp.n_out; op.b_sn = 1;
    op.M = B; op.N = lp.n_out; op.K = lp.n_in;
    op.C = ws + lp.z_off; op.ldc = lp.ld_out;
    op.bias = lp.has_bias ? params + lp.boff : nullptr;
    op.act = lp.act == DINO_ACT_SOFTPLUS;
    op.s_out = lp.sg_off >= 0 ? ws + lp.sg_off : nullptr; op.ld_s = lp.ld_out;
    if (lp.skip_from >= 0) {
      const int j = lp.skip_from;   // add the INPUT of layer j
      op.add_post = j == 0 ? ws + p->off_z0 : ws + p->layers[j - 1].z_off;
      op.ld_apo = j == 0 ? p->ld_in : p->layers[j - 1].ld_out;
    }
    if (tc_mlp) {
      GemmOp opt = op;
      opt.B = ws + lp.wtp_off; opt.b_sk = 1; opt.b_sn = lp.ld_wtp;             // K-major W^T for the tcgen05 engine
      if (p->split && lp.wtp_lo_off >= 0) opt.B_lo = ws + lp.wtp_lo_off;
      if (mlp_gemm(p, opt, op, st)) return 1;
    } else if (gemm_simt_only(p, op, st)) return 1;
    zin = ws + lp.z_off; ld_zin = lp.ld_out;
  }
  return 0;
}

// Top operand of the chain: Q_{L-1}.  Returns pointer / ld / row-modulus.
struct TopQ { const float* ptr; long long ld; int mod; };

static int run_chain_forward(DinoPlan* p, const float* params, const DinoBatch* b, int B, cudaStream_t st, TopQ* top_out,
                             bool skip_first = false) {
  float* ws = p->ws;
  const int s = p->s, L = p->L;
  const int M = B * s;
  const LayerPlan& last = p->layers[L - 1];
  TopQ top;
  const float* seed; long long ld_seed, seed_stride;
  if (p->jac_mode == DINO_JAC_FULL) { seed = ws + p->off_ident; ld_seed = p->ld_ident; seed_stride = 0; }
  else if (p->jac_mode == DINO_JAC_OUT_REDUCED) { seed = b->seed; ld_seed = p->d_out; seed_stride = 0; }
  else { seed = nullptr; ld_seed = 0; seed_stride = 0; }
  if (p->jac_mode == DINO_JAC_SKETCHED) {
    // Q_{L-1}[b] = U_b^T (* sg_{L-1})
    DINO_LAUNCH(launch_transpose_batched(b->U, B, p->d_out, s, ws + last.q_off, last.ld_out, b->idx, st));
    if (last.act == DINO_ACT_SOFTPLUS)
      DINO_LAUNCH(launch_scale_seed(ws + last.q_off, (long long)s * last.ld_out, last.ld_out, ws + last.sg_off,
                                    last.ld_out, B, s, p->d_out, ws + last.q_off, last.ld_out, st));
    top = {ws + last.q_off, last.ld_out, 0};
  } else if (p->top_materialised) {
    DINO_LAUNCH(launch_scale_seed(seed, seed_stride, ld_seed, last.sg_off >= 0 ? ws + last.sg_off : nullptr,
                                  last.ld_out, B, s, p->d_out, ws + last.q_off, last.ld_out, st));
    top = {ws + last.q_off, last.ld_out, 0};
  } else {
    top = {seed, ld_seed, s};
  }
  *top_out = top;
  const bool tc = p->tc;
  // (the staged weight copies were written ahead of the MLP forward; with the fused MLP kernels, here)
  if (tc && (p->mlp_fused || p->chain_fused)) DINO_LAUNCH(launch_prep_weights(params, ws, p->prep, p->max_in, p->max_out, st));
  for (int l = L - 1; l >= (skip_first ? 1 : 0); --l) {
    const LayerPlan& lp = p->layers[l];
    if (l == L - 1 && p->top_identity) {
      // I * W_L^T: row i of every sample is row i of the staged W_L^T, times sigma' of the layer below
      const LayerPlan* below = l > 0 ? &p->layers[l - 1] : nullptr;
      const bool sc = below && below->act == DINO_ACT_SOFTPLUS;
      DINO_LAUNCH(launch_scale_seed(ws + lp.wtp_off, 0, lp.ld_wtp, sc ? ws + below->sg_off : nullptr, sc ? below->ld_out : 0, B, s,
                                    lp.n_in, below ? ws + below->q_off : ws + p->off_r0, below ? below->ld_out : p->ld_in, st,
                                    (below && !p->split) ? 1 : 0));
      continue;
    }
    GemmOp op = gemm_op_zero();
    if (l == L - 1) { op.A = top.ptr; op.a_sm = top.ld; op.a_mod = top.mod; }
    else { op.A = ws + lp.q_off; op.a_sm = lp.ld_out; }
    op.a_sk = 1;
    // W_l^T: element (k = out index, n = in index) = W_l[n][k]; the TF32 engine reads the 16-byte-pitch copy
    if (tc) { op.B = ws + lp.wp_off; op.b_sk = 1; op.b_sn = lp.ld_wp; if (p->split && lp.wp_lo_off >= 0) op.B_lo = ws + lp.wp_lo_off; }
    else { op.B = params + lp.koff; op.b_sk = 1; op.b_sn = lp.n_out; }
    op.M = M; op.N = lp.n_in; op.K = lp.n_out; op.sdim = s;
    if (lp.lands_skip >= 0) {
      const LayerPlan& sk = p->layers[lp.lands_skip];
      op.add_pre = ws + sk.q_off; op.ld_ap = sk.ld_out;             // Q of the (linear) skip layer = R above it
    }
    if (l > 0) {
      const LayerPlan& below = p->layers[l - 1];
      if (below.act == DINO_ACT_SOFTPLUS) { op.scale = ws + below.sg_off; op.ld_sc = below.ld_out; }
      op.C = ws + below.q_off; op.ldc = below.ld_out;
      op.round_tf32 = (tc && !p->split) ? 1 : 0;
    } else {
      op.C = ws + p->off_r0; op.ldc = p->ld_in;
    }
    if (gemm(p, op, st)) return 1;
  }
  return 0;
}

static int push_wgrad(DinoPlan* p, GemmGroup& g, const GemmOp& op, cudaStream_t st);

// dino_plan_profile: CUDA events around the Jacobian-chain stage of a step (the fused chain kernel, or the chain
// contractions of the layer-by-layer path: forward chain, H1 loss, data-gradient chain)
static int prof_begin(DinoPlan* p, cudaStream_t st) {
  if (!p->profile) return 0;
  if (p->ev_pending && cudaEventSynchronize(p->ev1) == cudaSuccess) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, p->ev0, p->ev1) == cudaSuccess) { p->prof_ms += ms; p->prof_n += 1; }
  }
  DINO_CUDA(cudaEventRecord(p->ev0, st));
  return 0;
}
static int prof_end(DinoPlan* p, cudaStream_t st) {
  if (!p->profile) return 0;
  DINO_CUDA(cudaEventRecord(p->ev1, st));
  p->ev_pending = true;
  return 0;
}

// Fused tcgen05 path: prep weights, then ONE kernel runs chain forward [+ H1 loss [+ chain backward]].
// mlp_mode: 0 = the MLP forward (and the output loss) ran in their own launches; 1 = run the forward inside the chain
// kernel; 2 = forward + output-loss statistics; 3 = forward + statistics + d loss / d q_hat (coef_q = 2 w_l2 / B_global)
static int run_chain_fused(DinoPlan* p, const float* params, const DinoBatch* b, int B, float coef, int do_loss,
                           int do_backward, float* stats, cudaStream_t st, TopQ* top_out, bool weights_ready = false,
                           int mlp_mode = 0, float coef_q = 0.f) {
  float* ws = p->ws;
  const LayerPlan& last = p->layers[p->L - 1];
  const float* seed = p->jac_mode == DINO_JAC_OUT_REDUCED ? b->seed : nullptr;
  if (p->top_materialised) *top_out = {ws + last.q_off, last.ld_out, 0};
  else if (p->jac_mode == DINO_JAC_FULL) *top_out = {ws + p->off_ident, p->ld_ident, p->s};
  else *top_out = {b->seed, (long long)p->d_out, p->s};
  if (!weights_ready) DINO_LAUNCH(launch_prep_weights(params, ws, p->prep, p->max_in, p->max_out, st));
  if (prof_begin(p, st)) return 1;
  DINO_LAUNCH(launch_chain_fused(p->chain_maps, p->chain, p->chain_steps, ws, b->J, b->J_sqnorm, b->idx, seed, B, coef, do_loss,
                                 do_backward, stats, p->mlp, params, b->m, b->q, b->q_sqnorm, coef_q, mlp_mode, p->split, st));
  return prof_end(p, st);
}

// weight gradients of the chain: dW_l += dR_l^T Q_l for every trainable layer (joins the grouped launch)
static int push_chain_wgrads(DinoPlan* p, int B, float* grad, const TopQ& top, GemmGroup& wg, cudaStream_t st) {
  float* ws = p->ws;
  const int L = p->L, M = B * p->s;
  for (int l = 0; l < L; ++l) {
    const LayerPlan& lp = p->layers[l];
    if (!lp.trainable) continue;
    GemmOp op = gemm_op_zero();
    // dW_l += dR_l^T Q_l with both factors row-major [rows, n]: MN-major operands for the tensor-core engine,
    // plain strides for the SIMT engine
    op.A = l == 0 ? ws + p->off_r0 : ws + p->layers[l - 1].dr_off;
    op.a_sm = 1; op.a_sk = l == 0 ? p->ld_in : p->layers[l - 1].ld_out;
    if (l == L - 1) { op.B = top.ptr; op.b_sk = top.ld; op.b_kmod = top.mod; }
    else { op.B = ws + lp.q_off; op.b_sk = lp.ld_out; }
    op.b_sn = 1;
    op.M = lp.n_in; op.N = lp.n_out; op.K = M;
    op.C = grad + lp.koff; op.ldc = lp.n_out; op.atomic_c = 1;
    if (push_wgrad(p, wg, op, st)) return 1;
  }
  return 0;
}

static int run_losses(DinoPlan* p, const DinoBatch* b, int B, float w_l2, float w_h1, float inv_gb, bool grads,
                      float* stats, cudaStream_t st, bool q_only = false) {
  float* ws = p->ws;
  const LayerPlan& last = p->layers[p->L - 1];
  float* dq = grads ? ws + p->off_dz[p->L % 3] : nullptr;
  DINO_LAUNCH(launch_loss_q(ws + last.z_off, last.ld_out, b->q, b->q_sqnorm, b->idx, B, p->d_out,
                            2.f * w_l2 * inv_gb, dq, p->ld_dz, stats, st));
  if (p->jac_mode == DINO_JAC_NONE || q_only) return 0;
  const float cj = 2.f * w_h1 * inv_gb;
  if (p->jac_mode == DINO_JAC_SKETCHED) {
    const int k = p->s;
    DINO_LAUNCH(launch_sketch_close(ws + p->off_r0, p->ld_in, b->V, b->idx, B, k, p->d_in, ws + p->off_sk, st));
    DINO_LAUNCH(launch_loss_j(ws + p->off_sk, k, b->J, 0, b->J_sqnorm, b->idx, B, k, k, cj, grads ? 1 : 0, stats, st));
    if (grads)
      DINO_LAUNCH(launch_sketch_open(ws + p->off_sk, b->V, b->idx, B, k, p->d_in, ws + p->off_r0, p->ld_in, st));
  } else {
    DINO_LAUNCH(launch_loss_j(ws + p->off_r0, p->ld_in, b->J, b->ld_J, b->J_sqnorm, b->idx, B, p->s, p->d_in, cj,
                              grads ? 1 : 0, stats, st));
  }
  return 0;
}

// Appends one weight-gradient contraction to the pending group; flushes the group when it is full.
static int flush_wgrad(DinoPlan* p, GemmGroup& g, cudaStream_t st);
static int push_wgrad(DinoPlan* p, GemmGroup& g, const GemmOp& op, cudaStream_t st) {
  if (g.n == GEMM_GROUP_MAX && flush_wgrad(p, g, st)) return 1;
  g.ops[g.n++] = op;
  return 0;
}
static int flush_wgrad(DinoPlan* p, GemmGroup& g, cudaStream_t st) {
  if (g.n > 0 && p->tc) {
    // tensor-core weight gradients for every op that qualifies; the rest stays on the SIMT engine
    GemmGroup tc, rest;
    tc.n = rest.n = 0;
    for (int i = 0; i < g.n; ++i) {
      if (wgrad_tc_supported(g.ops[i])) tc.ops[tc.n++] = g.ops[i];
      else if (g.ops[i].a_sm == 1 && g.ops[i].b_sn == 1) rest.ops[rest.n++] = g.ops[i];
      else DINO_LAUNCH(launch_gemm_simt(g.ops[i], p->num_sms, st));
    }
    if (tc.n > 0) DINO_LAUNCH(launch_wgrad_tc(tc, p->split, p->num_sms, st));
    if (rest.n > 0) DINO_LAUNCH(launch_gemm_simt_grouped(rest, p->num_sms, st));
  } else if (g.n > 0) {
    DINO_LAUNCH(launch_gemm_simt_grouped(g, p->num_sms, st));
  }
  g.n = 0;
  return 0;
}

// ---- first layer without its Jacobian ------------------------------------------------------------------------
// FULL Jacobian of an unreduced input (d_in in the thousands, first layer W [d_in, n1], chain state Q [B*s, n1] above it):
//   J_hat = Q W^T  [B*s, d_in]   is the only d_in-wide array of the step; the explicit path writes it, reads it with the
// target, writes G = c_b (J_hat - J) over it and reads G twice more (data gradient G W, weight gradient G^T Q).  Both
// products are linear in G, so with T = J W and M = W^T W (n1 x n1):
//   G W   = c_b (Q M - T)                       -> contraction over d_in with the TARGET as operand, then an n1-wide one
//   G^T Q = W (sum_b c_b Q_b^T Q_b) - J^T (c_b Q)
//   ||J_hat - J||^2_row = <Q M, Q>_row - 2 <T, Q>_row + ||J_row||^2
// and J_hat never exists: the target is read twice (by TMA, gathered per sample through a 3-D tensor map) instead of
// the six passes over [B*s, d_in].  The arithmetic is the explicit path's to rounding: there J_hat carries the
// contraction's error (1e-6 |J_hat| in fp32x3) into the difference, here T and Q M carry it.
struct ImplicitCtx { const DinoBatch* b; float cj; float* stats; };

static long long implicit_nsamp(const DinoBatch* b, int B) {
  return b->idx ? (b->n_samples > 0 ? (long long)b->n_samples : (1LL << 22)) : (long long)B;
}
static GemmOp implicit_t1_op(const DinoPlan* p, const DinoBatch* b, int B) {
  const LayerPlan& l0 = p->layers[0];
  float* ws = p->ws;
  GemmOp op = gemm_op_zero();
  op.A = b->J; op.a_sm = b->ld_J; op.a_sk = 1;
  op.a_rows = p->s; op.a_ld2 = (long long)p->s * b->ld_J; op.a_idx = b->idx; op.a_nsamp = implicit_nsamp(b, B);
  op.B = ws + l0.wtp_off; op.b_sk = 1; op.b_sn = l0.ld_wtp;
  if (p->split && l0.wtp_lo_off >= 0) op.B_lo = ws + l0.wtp_lo_off;
  op.M = B * p->s; op.N = l0.n_out; op.K = l0.n_in; op.sdim = p->s;
  op.scale = ws + p->off_negc; op.ld_sc = l0.ld_out;
  op.yref = ws + l0.q_off; op.ld_y = l0.ld_out; op.rowdot = ws + p->off_tdot;
  op.C = ws + p->off_tb; op.ldc = l0.ld_out;
  return op;
}
static GemmOp implicit_t2_op(const DinoPlan* p, const DinoBatch* b, int B, float* grad) {
  const LayerPlan& l0 = p->layers[0];
  float* ws = p->ws;
  GemmOp op = gemm_op_zero();
  op.A = b->J; op.a_sm = 1; op.a_sk = b->ld_J;
  op.a_rows = p->s; op.a_ld2 = (long long)p->s * b->ld_J; op.a_idx = b->idx; op.a_nsamp = implicit_nsamp(b, B);
  op.B = ws + p->off_qc; op.b_sk = l0.ld_out; op.b_sn = 1;
  op.M = l0.n_in; op.N = l0.n_out; op.K = B * p->s;
  op.C = grad + l0.koff; op.ldc = l0.n_out; op.atomic_c = 1; op.alpha = -1.f;
  return op;
}
// this call can take the implicit first layer: plan, target layout and the engines' alignment rules
static bool implicit_applicable(const DinoPlan* p, const DinoBatch* b, int B, float w_h1, float* grad) {
  if (!p->implicit0 || w_h1 == 0.f || !b->J_row_sqnorm || b->ld_J < p->d_in || (b->ld_J & 3)) return false;
  return gemm_tc_supported(implicit_t1_op(p, b, B)) && wgrad_tc_supported(implicit_t2_op(p, b, B, grad));
}

static int flush_wgrad(DinoPlan* p, GemmGroup& g, cudaStream_t st);
static int push_wgrad(DinoPlan* p, GemmGroup& g, const GemmOp& op, cudaStream_t st);

// layer 0 of the backward chain in that form: leaves dR_1 (the state gradient above layer 0), rs_0, the H1 statistics,
// W S in the gradient of W (plain store: it runs before every accumulation into that block) and J^T (c Q) pending in wg
static int run_implicit_layer0(DinoPlan* p, const float* params, int B, float* grad, GemmGroup& wg, const ImplicitCtx& cx,
                               cudaStream_t st) {
  float* ws = p->ws;
  const DinoBatch* b = cx.b;
  const LayerPlan& l0 = p->layers[0];
  const int s = p->s, M = B * s, n1 = l0.n_out;
  const float* Q = ws + l0.q_off;
  const float* W = params + l0.koff;
  DINO_CUDA(cudaMemsetAsync(ws + p->off_mw, 0, (size_t)p->imp_zero_floats * sizeof(float), st));
  DINO_LAUNCH(launch_implicit_prep(Q, l0.ld_out, n1, b->J_sqnorm, b->idx, B, s, cx.cj, ws + p->off_qc, ws + p->off_negc, st));
  {   // M = W^T W and S = (c Q)^T Q: two small Gram matrices in one grouped launch
    GemmGroup g2;
    g2.n = 0;
    GemmOp op = gemm_op_zero();
    op.A = W; op.a_sm = 1; op.a_sk = n1; op.B = W; op.b_sk = n1; op.b_sn = 1;
    op.M = n1; op.N = n1; op.K = l0.n_in; op.C = ws + p->off_mw; op.ldc = n1; op.atomic_c = 1;
    g2.ops[g2.n++] = op;
    if (l0.trainable) {
      op = gemm_op_zero();
      op.A = ws + p->off_qc; op.a_sm = 1; op.a_sk = l0.ld_out; op.B = Q; op.b_sk = l0.ld_out; op.b_sn = 1;
      op.M = n1; op.N = n1; op.K = M; op.C = ws + p->off_sb; op.ldc = n1; op.atomic_c = 1;
      g2.ops[g2.n++] = op;
    }
    if (flush_wgrad(p, g2, st)) return 1;
  }
  // T = J W, stored as -c_b T; tdot[row] = <T, Q>_row
  if (gemm(p, implicit_t1_op(p, b, B), st)) return 1;
  {   // dQ_0 = c_b (Q M - T): row sums for d sigma, ddot[row] = <dQ_0, Q>_row, then * sg_0 -> dR_1
    GemmOp op = gemm_op_zero();
    op.A = ws + p->off_qc; op.a_sm = l0.ld_out; op.a_sk = 1;
    op.B = ws + p->off_mw; op.b_sk = 1; op.b_sn = n1;              // symmetric
    op.M = M; op.N = n1; op.K = n1; op.sdim = s;
    op.add_pre = ws + p->off_tb; op.ld_ap = l0.ld_out; op.pre_first = 1;
    op.yref = Q; op.ld_y = l0.ld_out;
    op.rowsum = ws + l0.rs_off; op.ld_rs = l0.ld_out;
    op.rowdot = ws + p->off_ddot;
    op.scale = ws + l0.sg_off; op.ld_sc = l0.ld_out;
    op.C = ws + l0.dr_off; op.ldc = l0.ld_out;
    op.round_tf32 = p->split ? 0 : 1;
    if (!gemm_tc_supported(op)) return fail("implicit first layer: contraction not supported by the tcgen05 engine");
    if (gemm(p, op, st)) return 1;
  }
  DINO_LAUNCH(launch_implicit_loss(ws + p->off_ddot, ws + p->off_tdot, b->J_sqnorm, b->J_row_sqnorm, b->idx, B, s, cx.cj,
                                   cx.stats, st));
  if (l0.trainable) {
    GemmOp op = gemm_op_zero();          // dW_0 = W S ...
    op.A = W; op.a_sm = n1; op.a_sk = 1;
    op.B = ws + p->off_sb; op.b_sk = 1; op.b_sn = n1;              // symmetric
    op.M = l0.n_in; op.N = n1; op.K = n1;
    op.C = grad + l0.koff; op.ldc = n1;
    if (!gemm_tc_supported(op)) return fail("implicit first layer: contraction not supported by the tcgen05 engine");
    if (gemm(p, op, st)) return 1;
    if (push_wgrad(p, wg, implicit_t2_op(p, b, B, grad), st)) return 1;      // ... - J^T (c Q), with the step's other weight gradients
  }
  return 0;
}

static int run_chain_backward(DinoPlan* p, const float* params, int B, float* grad, const TopQ& top, GemmGroup& wg,
                              cudaStream_t st, const ImplicitCtx* imp = nullptr) {
  float* ws = p->ws;
  const int s = p->s, L = p->L, M = B * s;
  // dR_l (gradient wrt the chain state entering layer l from below): l == 0 -> R0 buffer (holds G);
  // l >= 1 -> the buffer of layer l-1.  All are kept: the weight gradients run afterwards in one launch.
  auto dr_ptr = [&](int l) { return l == 0 ? ws + p->off_r0 : ws + p->layers[l - 1].dr_off; };
  auto dr_ld = [&](int l) { return l == 0 ? p->ld_in : p->layers[l - 1].ld_out; };
  for (int l = 0; l < L; ++l) {
    const LayerPlan& lp = p->layers[l];
    if (l == 0 && imp != nullptr) {
      if (run_implicit_layer0(p, params, B, grad, wg, *imp, st)) return 1;
      continue;
    }
    const float* q; long long ldq; int qmod;
    if (l == L - 1) { q = top.ptr; ldq = top.ld; qmod = top.mod; }
    else { q = ws + lp.q_off; ldq = lp.ld_out; qmod = 0; }
    if (lp.trainable && l == L - 1 && p->top_identity) {
      DINO_LAUNCH(launch_identity_seed_wgrad(dr_ptr(l), dr_ld(l), B, s, lp.n_in, grad + lp.koff, st));
    } else if (lp.trainable) {     // dW_l += dR_l^T Q_l   (reduction over all B*s chain rows, split-K)
      GemmOp op = gemm_op_zero();
      op.A = dr_ptr(l); op.a_sm = 1; op.a_sk = dr_ld(l);
      op.B = q; op.b_sk = ldq; op.b_sn = 1; op.b_kmod = qmod;
      op.M = lp.n_in; op.N = lp.n_out; op.K = M;
      op.C = grad + lp.koff; op.ldc = lp.n_out; op.atomic_c = 1;
      if (push_wgrad(p, wg, op, st)) return 1;
    }
    const bool need_rs = lp.act == DINO_ACT_SOFTPLUS;
    const bool need_out = l < L - 1;
    if (!need_rs && !need_out) continue;
    GemmOp op = gemm_op_zero();
    op.A = dr_ptr(l); op.a_sm = dr_ld(l); op.a_sk = 1;
    op.M = M; op.N = lp.n_out; op.K = lp.n_in; op.sdim = s;
    const float* skip_add = nullptr; long long ld_skip = 0;
    if (need_out && lp.skip_from >= 0) { skip_add = dr_ptr(lp.skip_from); ld_skip = dr_ld(lp.skip_from); }
    if (p->tc && qmod == 0) {
      // tensor-core path: dQ_l = dR_l W_l with the K-major copy W_l^T; the cross-row d sigma sum, the per-sample
      // scaling and the residual all happen in the epilogue of the same launch
      op.B = ws + lp.wtp_off; op.b_sk = 1; op.b_sn = lp.ld_wtp;
      if (p->split && lp.wtp_lo_off >= 0) op.B_lo = ws + lp.wtp_lo_off;
      if (need_rs) {
        op.yref = q; op.ld_y = ldq;
        op.rowsum = ws + lp.rs_off; op.ld_rs = lp.ld_out;
        op.scale = ws + lp.sg_off; op.ld_sc = lp.ld_out;
      }
      if (need_out) {
        op.C = dr_ptr(l + 1); op.ldc = dr_ld(l + 1);
        op.add_post = skip_add; op.ld_apo = ld_skip;
        op.round_tf32 = p->split ? 0 : 1;
      }
      if (gemm(p, op, st)) return 1;
      continue;
    }
    op.B = params + lp.koff; op.b_sk = lp.n_out; op.b_sn = 1;      // W_l
    if (need_rs) {
      op.yref = q; op.ld_y = ldq;       // softplus top layers are always materialised (qmod == 0)
      op.rowsum = ws + lp.rs_off; op.ld_rs = lp.ld_out;
      op.scale = ws + lp.sg_off; op.ld_sc = lp.ld_out;
    }
    if (need_out) {
      op.C = dr_ptr(l + 1); op.ldc = dr_ld(l + 1);
      op.add_post = skip_add; op.ld_apo = ld_skip;
    }
    if (gemm(p, op, st)) return 1;
  }
  return 0;
}

static int run_mlp_backward(DinoPlan* p, const float* params, int B, float* grad, GemmGroup& wg, cudaStream_t st) {
  float* ws = p->ws;
  const int L = p->L;
  if (p->mlp_fused) {
    // one launch produces da_l for every layer (and the bias gradients); the weight gradients
    // dW_l += z_l^T da_l join the chain's in the grouped launch
    DINO_LAUNCH(launch_mlp_backward_fused(p->mlp, params, ws, grad, B, p->jac_mode != DINO_JAC_NONE ? 1 : 0, st));
    for (int l = p->mlp.lowest_trainable; l < L; ++l) {
      const LayerPlan& lp = p->layers[l];
      if (!lp.trainable) continue;
      GemmOp op = gemm_op_zero();
      op.A = l == 0 ? ws + p->off_z0 : ws + p->layers[l - 1].z_off;
      op.a_sm = 1; op.a_sk = l == 0 ? p->ld_in : p->layers[l - 1].ld_out;
      op.B = ws + lp.da_off; op.b_sk = lp.ld_out; op.b_sn = 1;
      op.M = lp.n_in; op.N = lp.n_out; op.K = B;
      op.C = grad + lp.koff; op.ldc = lp.n_out; op.atomic_c = 1;
      if (push_wgrad(p, wg, op, st)) return 1;
    }
    return 0;
  }
  auto dz_ptr = [&](int l) { return ws + p->off_dz[l % 3]; };
  const bool tc_mlp = p->tc && p->prep.n > 0 && !p->chain_fused;
  int lowest_trainable = L;
  for (int l = 0; l < L; ++l)
    if (p->layers[l].trainable) { lowest_trainable = l; break; }
  for (int l = L - 1; l >= lowest_trainable; --l) {
    const LayerPlan& lp = p->layers[l];
    const float* g = dz_ptr(l + 1);
    const float* da = g;
    long long ld_da = p->ld_dz;
    bool da_kept = false;           // da_l has a buffer of its own: its weight gradient can wait for the grouped launch
    float* dbias = (lp.trainable && lp.has_bias) ? grad + lp.boff : nullptr;
    if (lp.act == DINO_ACT_SOFTPLUS) {
      float* dab = lp.da_off >= 0 ? ws + lp.da_off : ws + p->off_da;
      if (lp.da_off >= 0) { ld_da = lp.ld_out; da_kept = true; }
      DINO_LAUNCH(launch_mlp_da(g, p->ld_dz, ws + lp.sg_off, lp.ld_out, lp.rs_off >= 0 ? ws + lp.rs_off : nullptr,
                                lp.ld_out, B, lp.n_out, dab, ld_da, dbias, st));
      da = dab;
    } else if (dbias != nullptr) {
      DINO_LAUNCH(launch_mlp_da(g, p->ld_dz, nullptr, 0, nullptr, 0, B, lp.n_out, nullptr, 0, dbias, st));
    }
    if (lp.trainable) {     // dW_l += z_l^T da_l
      GemmOp op = gemm_op_zero();
      op.A = l == 0 ? ws + p->off_z0 : ws + p->layers[l - 1].z_off;
      op.a_sm = 1; op.a_sk = l == 0 ? p->ld_in : p->layers[l - 1].ld_out;
      op.B = da; op.b_sk = ld_da; op.b_sn = 1;
      op.M = lp.n_in; op.N = lp.n_out; op.K = B;
      op.C = grad + lp.koff; op.ldc = lp.n_out; op.atomic_c = 1;
      if (da_kept) {
        if (push_wgrad(p, wg, op, st)) return 1;
      } else {
        // (dz of a linear layer lives in a rotating buffer: this contraction cannot wait for the grouped launch)
        GemmGroup one;
        one.n = 1; one.ops[0] = op;
        if (flush_wgrad(p, one, st)) return 1;
      }
    }
    if (l > lowest_trainable) {   // dz_l = da_l W_l^T [+ dz above the residual block that starts here]
      GemmOp op = gemm_op_zero();
      op.A = da; op.a_sm = ld_da; op.a_sk = 1;
      op.B = params + lp.koff; op.b_sk = 1; op.b_sn = lp.n_out;
      op.M = B; op.N = lp.n_in; op.K = lp.n_out;
      op.C = dz_ptr(l); op.ldc = p->ld_dz;
      if (lp.lands_skip >= 0) { op.add_post = dz_ptr(lp.lands_skip + 1); op.ld_apo = p->ld_dz; }
      if (tc_mlp) {
        GemmOp opt = op;
        opt.B = ws + lp.wp_off; opt.b_sk = 1; opt.b_sn = lp.ld_wp;               // 16-byte-pitch copy of W_l
        if (p->split && lp.wp_lo_off >= 0) opt.B_lo = ws + lp.wp_lo_off;
        if (mlp_gemm(p, opt, op, st)) return 1;
      } else if (gemm_simt_only(p, op, st)) return 1;
    }
  }
  return 0;
}

extern "C" int dino_forward(DinoPlan* p, const float* params, const DinoBatch* b, int32_t B, float* q_out,
                            dino_stream_t stream) {
  if (!p || !params || !b || !q_out) return fail("dino_forward: null argument");
  if (!p->ws) return fail("dino_forward: workspace not bound");
  if (B < 1 || B > p->max_batch) return fail("dino_forward: batch size outside [1, max_batch]");
  if (!b->m) return fail("dino_forward: batch.m is null");
  cudaStream_t st = S(stream);
  if (run_mlp_forward(p, params, b, B, st)) return 1;
  const LayerPlan& last = p->layers[p->L - 1];
  DINO_LAUNCH(launch_copy2d(p->ws + last.z_off, last.ld_out, q_out, p->d_out, B, p->d_out, st));
  return 0;
}

extern "C" int dino_jacobian(DinoPlan* p, const float* params, const DinoBatch* b, int32_t B, float* q_out,
                             float* J_out, dino_stream_t stream) {
  if (!J_out) return fail("dino_jacobian: null output");
  if (check_call(p, params, b, B, false)) return 1;
  if (p->jac_mode == DINO_JAC_NONE) return fail("dino_jacobian: plan has no Jacobian output");
  cudaStream_t st = S(stream);
  if (!p->mlp_in_chain && run_mlp_forward(p, params, b, B, st)) return 1;
  TopQ top;
  if (p->chain_fused) {
    if (run_chain_fused(p, params, b, B, 0.f, 0, 0, nullptr, st, &top, false, p->mlp_in_chain ? 1 : 0)) return 1;
  } else if (run_chain_forward(p, params, b, B, st, &top)) return 1;
  const LayerPlan& last = p->layers[p->L - 1];
  if (q_out) DINO_LAUNCH(launch_copy2d(p->ws + last.z_off, last.ld_out, q_out, p->d_out, B, p->d_out, st));
  if (p->jac_mode == DINO_JAC_SKETCHED)
    DINO_LAUNCH(launch_sketch_close(p->ws + p->off_r0, p->ld_in, b->V, b->idx, B, p->s, p->d_in, J_out, st));
  else
    DINO_LAUNCH(launch_copy2d(p->ws + p->off_r0, p->ld_in, J_out, p->d_in, B * p->s, p->d_in, st));
  return 0;
}

// J(m) mhat by ONE forward-mode sweep: t_0 = mhat, t_{l+1} = (t_l W_l) * sg_l [+ t_skip]  (the tangent of the MLP forward).
// The reference contracts the full batch Jacobian with mhat (inferenceOracle.py:96-98); a single column costs one
// more pass of the network's own size.  mhat_rows = 1: one direction shared by all samples; = B: one per sample.
extern "C" int dino_jvp(DinoPlan* p, const float* params, const DinoBatch* b, int32_t B, const float* d_mhat,
                        int32_t mhat_rows, float* d_out, dino_stream_t stream) {
  if (!p || !params || !b || !d_mhat || !d_out) return fail("dino_jvp: null argument");
  if (!p->ws) return fail("dino_jvp: workspace not bound");
  if (B < 1 || B > p->max_batch) return fail("dino_jvp: batch size outside [1, max_batch]");
  if (!b->m) return fail("dino_jvp: batch.m is null");
  if (mhat_rows != 1 && mhat_rows != B) return fail("dino_jvp: mhat must have 1 row or one row per sample");
  cudaStream_t st = S(stream);
  if (run_mlp_forward(p, params, b, B, st)) return 1;        // leaves sg_l = act'(a_l) of every layer in the workspace
  float* ws = p->ws;
  auto t_ptr = [&](int l) { return ws + p->off_dz[l % 3]; };
  // t_0 = mhat, one row per sample (a shared direction is broadcast)
  DINO_LAUNCH(launch_scale_seed(d_mhat, mhat_rows == 1 ? 0 : p->d_in, p->d_in, nullptr, 0, B, 1, p->d_in, t_ptr(0), p->ld_dz, st));
  for (int l = 0; l < p->L; ++l) {
    const LayerPlan& lp = p->layers[l];
    GemmOp op = gemm_op_zero();
    op.A = t_ptr(l); op.a_sm = p->ld_dz; op.a_sk = 1;
    op.B = params + lp.koff; op.b_sk = lp.n_out; op.b_sn = 1;
    op.M = B; op.N = lp.n_out; op.K = lp.n_in; op.sdim = 1;
    if (lp.act == DINO_ACT_SOFTPLUS) { op.scale = ws + lp.sg_off; op.ld_sc = lp.ld_out; }
    if (lp.skip_from >= 0) {
      op.add_post = t_ptr(lp.skip_from); op.ld_apo = p->ld_dz;
    }
    if (l == p->L - 1) { op.C = d_out; op.ldc = p->d_out; }
    else { op.C = t_ptr(l + 1); op.ldc = p->ld_dz; }
    if (gemm_simt_only(p, op, st)) return 1;
  }
  return 0;
}

extern "C" int dino_h1_fwd_bwd(DinoPlan* p, const float* params, const DinoBatch* b, int32_t B, float w_l2,
                               float w_h1, float inv_gb, float* grad, float* stats, dino_stream_t stream) {
  if (!grad || !stats) return fail("dino_h1_fwd_bwd: null output");
  if (check_call(p, params, b, B, true)) return 1;
  cudaStream_t st = S(stream);
  DINO_CUDA(cudaMemsetAsync(grad, 0, (size_t)p->n_params * sizeof(float), st));
  if (p->rs_floats > 0)
    DINO_CUDA(cudaMemsetAsync(p->ws + p->off_rs_begin, 0, (size_t)p->rs_floats * sizeof(float), st));
  const bool prep_aside = p->chain_fused && p->side != nullptr;
  if (prep_aside) {
    // the padded / rounded weight tiles of the chain only depend on the parameters: build them next to the MLP forward
    DINO_CUDA(cudaEventRecord(p->ev_fork, st));
    DINO_CUDA(cudaStreamWaitEvent(p->side, p->ev_fork, 0));
    DINO_LAUNCH(launch_prep_weights(params, p->ws, p->prep, p->max_in, p->max_out, p->side));
    DINO_CUDA(cudaEventRecord(p->ev_prep, p->side));
  }
  if (!p->mlp_in_chain && run_mlp_forward(p, params, b, B, st)) return 1;
  TopQ top = {nullptr, 0, 0};
  GemmGroup wg;
  wg.n = 0;
  if (p->chain_fused) {
    if (!p->mlp_in_chain && run_losses(p, b, B, w_l2, w_h1, inv_gb, true, stats, st, true)) return 1;
    if (prep_aside) DINO_CUDA(cudaStreamWaitEvent(st, p->ev_prep, 0));
    if (run_chain_fused(p, params, b, B, 2.f * w_h1 * inv_gb, 1, 1, stats, st, &top, prep_aside, p->mlp_in_chain ? 3 : 0,
                        2.f * w_l2 * inv_gb))
      return 1;
    if (p->side != nullptr) {
      // fork: the chain's weight gradients (tensor cores, side stream) overlap the MLP backward (main stream);
      // both only ADD into the gradient, which was cleared before the fork
      GemmGroup cw;
      cw.n = 0;
      DINO_CUDA(cudaEventRecord(p->ev_fork, st));
      DINO_CUDA(cudaStreamWaitEvent(p->side, p->ev_fork, 0));
      if (push_chain_wgrads(p, B, grad, top, cw, p->side)) return 1;
      if (flush_wgrad(p, cw, p->side)) return 1;
      DINO_CUDA(cudaEventRecord(p->ev_join, p->side));
      if (run_mlp_backward(p, params, B, grad, wg, st)) return 1;
      if (flush_wgrad(p, wg, st)) return 1;
      DINO_CUDA(cudaStreamWaitEvent(st, p->ev_join, 0));
      return 0;
    }
    if (push_chain_wgrads(p, B, grad, top, wg, st)) return 1;
  } else {
    const bool imp = implicit_applicable(p, b, B, w_h1, grad);
    const ImplicitCtx cx = {b, 2.f * w_h1 * inv_gb, stats};
    if (prof_begin(p, st)) return 1;
    if (p->jac_mode != DINO_JAC_NONE && run_chain_forward(p, params, b, B, st, &top, imp)) return 1;
    if (run_losses(p, b, B, w_l2, w_h1, inv_gb, true, stats, st, imp)) return 1;
    if (p->jac_mode != DINO_JAC_NONE && run_chain_backward(p, params, B, grad, top, wg, st, imp ? &cx : nullptr)) return 1;
    if (prof_end(p, st)) return 1;
  }
  if (!p->chain_fused && p->side != nullptr && p->jac_mode != DINO_JAC_NONE && wg.n > 0) {
    // fork: network backward (and its weight gradients) on the high-priority side stream, the chain's grouped weight
    // gradients on the main stream; both only ADD into the gradient (the one plain store into it, W_1 S of the
    // Jacobian-free first layer, was enqueued before the fork)
    GemmGroup wg2;
    wg2.n = 0;
    DINO_CUDA(cudaEventRecord(p->ev_fork, st));
    DINO_CUDA(cudaStreamWaitEvent(p->side, p->ev_fork, 0));
    if (run_mlp_backward(p, params, B, grad, wg2, p->side)) return 1;
    if (flush_wgrad(p, wg2, p->side)) return 1;
    DINO_CUDA(cudaEventRecord(p->ev_join, p->side));
    if (flush_wgrad(p, wg, st)) return 1;
    DINO_CUDA(cudaStreamWaitEvent(st, p->ev_join, 0));
    return 0;
  }
  if (run_mlp_backward(p, params, B, grad, wg, st)) return 1;
  return flush_wgrad(p, wg, st);
}

extern "C" int dino_h1_eval(DinoPlan* p, const float* params, const DinoBatch* b, int32_t B, float* stats,
                            dino_stream_t stream) {
  if (!stats) return fail("dino_h1_eval: null output");
  if (check_call(p, params, b, B, true)) return 1;
  cudaStream_t st = S(stream);
  if (!p->mlp_in_chain && run_mlp_forward(p, params, b, B, st)) return 1;
  TopQ top = {nullptr, 0, 0};
  if (p->chain_fused) {
    if (!p->mlp_in_chain && run_losses(p, b, B, 1.f, 1.f, 1.f, false, stats, st, true)) return 1;
    return run_chain_fused(p, params, b, B, 2.f, 1, 0, stats, st, &top, false, p->mlp_in_chain ? 2 : 0);
  }
  if (p->jac_mode != DINO_JAC_NONE && run_chain_forward(p, params, b, B, st, &top)) return 1;
  return run_losses(p, b, B, 1.f, 1.f, 1.f, false, stats, st);
}

extern "C" int dino_adam(float* params, const float* grad, float* m, float* v, int64_t count, float lr, float b1,
                         float b2, float eps, int64_t t, dino_stream_t stream) {
  if (!params || !grad || !m || !v) return fail("dino_adam: null argument");
  if (t < 1) return fail("dino_adam: step t must be >= 1");
  // the same formula as adam_dev_kernel (1 - b^t = -expm1(t log b)), in double
  const double lr_t = (double)lr * sqrt(-expm1((double)t * log((double)b2))) / (-expm1((double)t * log((double)b1)));
  DINO_LAUNCH(launch_adam(params, grad, m, v, count, (float)lr_t, b1, b2, eps, S(stream)));
  return 0;
}

extern "C" int dino_adam_dev(float* params, const float* grad, float* m, float* v, int64_t count, float lr, float b1,
                             float b2, float eps, const int64_t* d_steps_done, dino_stream_t stream) {
  if (!params || !grad || !m || !v || !d_steps_done) return fail("dino_adam_dev: null argument");
  DINO_LAUNCH(launch_adam_dev(params, grad, m, v, count, lr, b1, b2, eps,
                              reinterpret_cast<const long long*>(d_steps_done), S(stream)));
  return 0;
}

extern "C" int dino_stage_batch(const int32_t* d_idx_all, const int32_t* d_pos, int32_t* d_idx_stage, int32_t B,
                                dino_stream_t stream) {
  if (!d_idx_all || !d_pos || !d_idx_stage) return fail("dino_stage_batch: null argument");
  DINO_LAUNCH(launch_stage_batch(d_idx_all, d_pos, d_idx_stage, B, S(stream)));
  return 0;
}

extern "C" int dino_roll_stats(float* d_stats_stage, float* d_stats_hist, int32_t* d_pos, int32_t advance,
                               int32_t* d_step, int64_t* d_steps_done, dino_stream_t stream) {
  if (!d_stats_stage || !d_stats_hist || !d_pos || !d_step || !d_steps_done) return fail("dino_roll_stats: null argument");
  DINO_LAUNCH(launch_roll_stats(d_stats_stage, d_stats_hist, d_pos, advance, d_step,
                                reinterpret_cast<long long*>(d_steps_done), S(stream)));
  return 0;
}

extern "C" int dino_sgd(float* params, const float* grad, int64_t count, float lr, dino_stream_t stream) {
  if (!params || !grad) return fail("dino_sgd: null argument");
  DINO_LAUNCH(launch_sgd(params, grad, count, lr, S(stream)));
  return 0;
}

extern "C" int dino_sample_sqnorms(const float* x, int64_t n, int64_t len, float* out, dino_stream_t stream) {
  if (!x || !out) return fail("dino_sample_sqnorms: null argument");
  DINO_LAUNCH(launch_sqnorms(x, n, len, out, S(stream)));
  return 0;
}

extern "C" int dino_gemm_nt(int engine, const float* A, int64_t lda, const float* Bt, int64_t ldb, float* C,
                            int64_t ldc, int32_t M, int32_t N, int32_t K, dino_stream_t stream) {
  if (!A || !Bt || !C) return fail("dino_gemm_nt: null argument");
  GemmOp op = gemm_op_zero();
  op.A = A; op.a_sm = lda; op.a_sk = 1;
  op.B = Bt; op.b_sk = 1; op.b_sn = ldb;
  op.M = M; op.N = N; op.K = K; op.C = C; op.ldc = ldc;
  if (engine == DINO_PREC_TF32 || engine == DINO_PREC_FP32X3) {
    if (!gemm_tc_supported(op)) return fail("dino_gemm_nt: operands do not meet the tcgen05 engine's alignment rules");
    DINO_LAUNCH(launch_gemm_tc(op, engine == DINO_PREC_FP32X3 ? 1 : 0, S(stream)));
  } else {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    DINO_LAUNCH(launch_gemm_simt(op, sms, S(stream)));
  }
  return 0;
}

extern "C" int dino_gemm_nt_acc(int engine, const float* XT, int64_t ldx, const float* YT, int64_t ldy, float* C,
                                int32_t rows, int32_t n_in, int32_t n_out, dino_stream_t stream) {
  if (!XT || !YT || !C) return fail("dino_gemm_nt_acc: null argument");
  GemmGroup g;
  g.n = 1;
  GemmOp& op = g.ops[0];
  op = gemm_op_zero();
  op.A = XT; op.a_sm = ldx; op.a_sk = 1;
  op.B = YT; op.b_sk = 1; op.b_sn = ldy;
  op.M = n_in; op.N = n_out; op.K = rows;
  op.C = C; op.ldc = n_out; op.atomic_c = 1;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (engine == DINO_PREC_TF32 || engine == DINO_PREC_FP32X3) {
    if (!wgrad_tc_supported(op)) return fail("dino_gemm_nt_acc: operands do not meet the tensor-core weight-gradient rules");
    DINO_LAUNCH(launch_wgrad_tc(g, engine == DINO_PREC_FP32X3 ? 1 : 0, sms, S(stream)));
  } else {
    DINO_LAUNCH(launch_gemm_simt(op, sms, S(stream)));
  }
  return 0;
}

extern "C" int dino_gemm_tn_acc(int engine, const float* X, int64_t ldx, const float* Y, int64_t ldy, float* C,
                                int32_t rows, int32_t n_in, int32_t n_out, dino_stream_t stream) {
  if (!X || !Y || !C) return fail("dino_gemm_tn_acc: null argument");
  GemmGroup g;
  g.n = 1;
  GemmOp& op = g.ops[0];
  op = gemm_op_zero();
  op.A = X; op.a_sm = 1; op.a_sk = ldx;
  op.B = Y; op.b_sk = ldy; op.b_sn = 1;
  op.M = n_in; op.N = n_out; op.K = rows;
  op.C = C; op.ldc = n_out; op.atomic_c = 1;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (engine == DINO_PREC_TF32 || engine == DINO_PREC_FP32X3) {
    if (!wgrad_tc_supported(op)) return fail("dino_gemm_tn_acc: operands do not meet the tensor-core weight-gradient rules");
    DINO_LAUNCH(launch_wgrad_tc(g, engine == DINO_PREC_FP32X3 ? 1 : 0, sms, S(stream)));
  } else {
    DINO_LAUNCH(launch_gemm_simt(op, sms, S(stream)));
  }
  return 0;
}

// debug aid: per-phase timestamps of the fused chain kernel (zeros unless built with -DDINO_CF_TRACE)
extern "C" int dino_debug_chain_trace(uint64_t* out_64x8) {
  if (!out_64x8) return fail("dino_debug_chain_trace: null argument");
  DINO_CUDA(cudaDeviceSynchronize());
  DINO_CUDA(chain_fused_read_trace(reinterpret_cast<unsigned long long*>(out_64x8)));
  return 0;
}

// Per-kernel timing of the dominant kernel (the fused Jacobian-chain kernel), measured with CUDA events on the
// launching stream.  enable=1 starts, enable=0 stops; out_ms_sum / out_count receive the totals so far.
// Only meaningful for eager (non-graph) steps.
extern "C" int dino_plan_profile(DinoPlan* p, int enable, double* out_ms_sum, int64_t* out_count) {
  if (!p) return fail("dino_plan_profile: null plan");
  if (enable && !p->ev0) {
    DINO_CUDA(cudaEventCreate(&p->ev0));
    DINO_CUDA(cudaEventCreate(&p->ev1));
  }
  if (p->ev_pending && cudaEventSynchronize(p->ev1) == cudaSuccess) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, p->ev0, p->ev1) == cudaSuccess) { p->prof_ms += ms; p->prof_n += 1; }
    p->ev_pending = false;
  }
  if (out_ms_sum) *out_ms_sum = p->prof_ms;
  if (out_count) *out_count = p->prof_n;
  if (enable && !p->profile) { p->prof_ms = 0.0; p->prof_n = 0; }
  p->profile = enable != 0;
  return 0;
}
