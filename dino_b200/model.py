"""DinoModel: the Keras ``Model`` surface the reference drivers use, backed by the CUDA library.

The reference builds tf.keras functional models (neuralNetworks.py:21-172) and drives them with
``compile / fit / evaluate / predict`` plus ``layers``, ``get_layer``, ``get_weights`` and
``set_weights`` (SURVEY.md 8b lists every call site).  This module offers the same duck-typed
surface.  Host code here only marshals: arrays go to HBM through torch tensors, every
arithmetic step is a call into libdino_b200.so (include/dino_b200.h).  No CPU compute path exists.
"""
from __future__ import annotations

import ctypes as C
import math
import os
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence

import numpy as np

from . import _lib
from . import parallel

ACT_NAMES = {None: _lib.ACT_LINEAR, "linear": _lib.ACT_LINEAR, "softplus": _lib.ACT_SOFTPLUS}


def _r4(n: int) -> int:
    return (n + 3) & ~3


# --------------------------------------------------------------------------------------------
# optimisers (the objects trainingUtilities.py:89-94 constructs)
# --------------------------------------------------------------------------------------------
class Adam:
    """Keras Adam: beta_1=0.9, beta_2=0.999, epsilon=1e-7 placed outside the bias correction."""

    def __init__(self, learning_rate=1e-3, beta_1=0.9, beta_2=0.999, epsilon=1e-7):
        self.learning_rate, self.beta_1, self.beta_2, self.epsilon = learning_rate, beta_1, beta_2, epsilon
        self.iterations = 0
        self._m = self._v = None

    kind = "adam"


class SGD:
    def __init__(self, learning_rate=1e-2):
        self.learning_rate = learning_rate
        self.iterations = 0

    kind = "sgd"


def _as_optimizer(opt):
    if isinstance(opt, (Adam, SGD)):
        return opt
    if isinstance(opt, str):
        if opt.lower() == "adam":
            return Adam()
        if opt.lower() == "sgd":
            return SGD()
    raise TypeError(f"unsupported optimizer {opt!r}: use dino_b200.Adam / dino_b200.SGD / 'adam' / 'sgd'")


# --------------------------------------------------------------------------------------------
# Keras-like layer / tensor placeholders
# --------------------------------------------------------------------------------------------
@dataclass
class TensorSpec:
    shape: tuple          # (None, d) / (None, s, d) : static shapes read at inferenceOracle.py:63-66
    name: str = ""


class DinoLayer:
    """One entry of ``model.layers``.  Dense layers own a slice of the parameter arena; the input
    and Add pseudo-layers carry no weights (Keras lists them too, and the reference pickles
    ``{layer.name: layer.get_weights()}`` over all of them, jacobianTrainingDriver.py:326-328)."""

    def __init__(self, core, name, index=None):
        self._core, self.name, self._index = core, name, index

    @property
    def trainable(self):
        return True if self._index is None else bool(self._core.trainable[self._index])

    @trainable.setter
    def trainable(self, value):
        if self._index is not None:
            self._core.trainable[self._index] = bool(value)
            self._core.invalidate_plans()

    def get_weights(self) -> List[np.ndarray]:
        if self._index is None:
            return []
        return self._core.layer_weights(self._index)

    def set_weights(self, weights: Sequence[np.ndarray]):
        if self._index is None:
            assert len(weights) == 0
            return
        self._core.set_layer_weights(self._index, weights)

    def count_params(self):
        return sum(int(w.size) for w in self.get_weights())


class NetCore:
    """Shared state of a network: layer table, parameter arena (host master copy + device mirror)."""

    def __init__(self, d_in: int, name_prefix: str = ""):
        self.d_in = int(d_in)
        self.name_prefix = name_prefix
        self.names: List[str] = []
        self.n_in: List[int] = []
        self.n_out: List[int] = []
        self.act: List[int] = []
        self.use_bias: List[bool] = []
        self.trainable: List[bool] = []
        self.skip_from: List[int] = []
        self.keras_layers: List[DinoLayer] = [DinoLayer(self, name_prefix + "input_data")]
        self.koff: List[int] = []
        self.boff: List[int] = []
        self.n_params = 0
        self.host = np.zeros(0, dtype=np.float32)
        self._dev = None                 # torch tensor mirror
        self._dev_newer = False          # device copy has updates the host copy lacks
        self._plans: Dict[tuple, "_Plan"] = {}

    # ---- construction -------------------------------------------------------------------------
    def add_dense(self, n_out, name, activation=None, use_bias=True, trainable=True, skip_from=-1,
                  kernel=None, bias=None):
        n_in = self.n_out[-1] if self.n_out else self.d_in
        self.names.append(name)
        self.n_in.append(int(n_in))
        self.n_out.append(int(n_out))
        self.act.append(ACT_NAMES[activation])
        self.use_bias.append(bool(use_bias))
        self.trainable.append(bool(trainable))
        self.skip_from.append(int(skip_from))
        self.koff.append(self.n_params)
        self.n_params += _r4(n_in * n_out)
        if use_bias:
            self.boff.append(self.n_params)
            self.n_params += _r4(n_out)
        else:
            self.boff.append(-1)
        self.host = np.concatenate([self.host, np.zeros(self.n_params - self.host.size, dtype=np.float32)])
        idx = len(self.names) - 1
        self.keras_layers.append(DinoLayer(self, name, idx))
        if kernel is not None:
            self.set_layer_weights(idx, [kernel] if bias is None else [kernel, bias])
        return idx

    def add_pseudo(self, name):
        self.keras_layers.append(DinoLayer(self, name))

    @property
    def d_out(self):
        return self.n_out[-1]

    @property
    def n_layers(self):
        return len(self.names)

    # ---- weights ------------------------------------------------------------------------------
    def _sync_host(self):
        if self._dev_newer and self._dev is not None:
            self.host = self._dev.detach().cpu().numpy().copy()
            self._dev_newer = False

    def layer_weights(self, l):
        self._sync_host()
        k, n = self.n_in[l], self.n_out[l]
        out = [self.host[self.koff[l]:self.koff[l] + k * n].reshape(k, n).copy()]
        if self.use_bias[l]:
            out.append(self.host[self.boff[l]:self.boff[l] + n].copy())
        return out

    def set_layer_weights(self, l, weights):
        self._sync_host()
        k, n = self.n_in[l], self.n_out[l]
        expect = 2 if self.use_bias[l] else 1
        if len(weights) != expect:
            raise ValueError(f"layer {self.names[l]} expects {expect} weight arrays, got {len(weights)}")
        W = np.asarray(weights[0], dtype=np.float32)
        if W.shape != (k, n):
            raise ValueError(f"layer {self.names[l]}: kernel shape {W.shape} != {(k, n)}")
        self.host[self.koff[l]:self.koff[l] + k * n] = W.reshape(-1)
        if self.use_bias[l]:
            b = np.asarray(weights[1], dtype=np.float32)
            if b.shape != (n,):
                raise ValueError(f"layer {self.names[l]}: bias shape {b.shape} != {(n,)}")
            self.host[self.boff[l]:self.boff[l] + n] = b
        if self._dev is not None:
            import torch
            self._dev.copy_(torch.from_numpy(self.host))

    def device_params(self):
        import torch
        if self._dev is None:
            self._dev = torch.from_numpy(self.host).to(_device())
        return self._dev

    def mark_device_updated(self):
        self._dev_newer = True

    def invalidate_plans(self):
        for p in self._plans.values():
            p.close()
        self._plans.clear()


def _device():
    import torch
    if not torch.cuda.is_available():
        raise _lib.DinoLibraryError("dino_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def _stream():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


class _Plan:
    """A DinoPlan handle plus its workspace tensor."""

    def __init__(self, core: NetCore, jac_mode: int, seed_dim: int, max_batch: int, precision: int):
        import torch
        lib = _lib.load()
        L = core.n_layers
        arr = (_lib.DinoLayerSpec * L)()
        for l in range(L):
            arr[l] = _lib.DinoLayerSpec(core.n_out[l], core.act[l], int(core.use_bias[l]), int(core.trainable[l]),
                                        core.skip_from[l])
        spec = _lib.DinoNetSpec(core.d_in, L, arr, jac_mode, seed_dim, max_batch, precision)
        handle = C.c_void_p()
        dev = _device()
        _lib.check(lib.dino_plan_create(C.byref(spec), dev.index, C.byref(handle)), "dino_plan_create")
        self.handle = handle
        self.lib = lib
        if lib.dino_plan_param_count(handle) != core.n_params:
            raise _lib.DinoLibraryError("parameter arena layout mismatch between host and library")
        ko, bo = (C.c_int64 * L)(), (C.c_int64 * L)()
        _lib.check(lib.dino_plan_layer_offsets(handle, ko, bo), "dino_plan_layer_offsets")
        if list(ko) != core.koff or list(bo) != core.boff:
            raise _lib.DinoLibraryError("parameter arena offsets mismatch between host and library")
        nbytes = lib.dino_plan_workspace_bytes(handle)
        self.workspace = torch.empty(max(nbytes, 256) // 4 + 64, dtype=torch.float32, device=dev)
        _lib.check(lib.dino_plan_bind_workspace(handle, C.c_void_p(self.workspace.data_ptr()),
                                                C.c_size_t(self.workspace.numel() * 4), _stream()),
                   "dino_plan_bind_workspace")
        self.max_batch = max_batch

    def close(self):
        if self.handle is not None:
            self.lib.dino_plan_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class DeviceDataset:
    """(m, q, J[, U, V]) of one fit/evaluate call, resident in HBM as fp32 (numpy fp64 sources are
    cast on upload, exactly as Keras casts to floatx on feed)."""

    ld_J = 0        # row pitch of J in floats (0: dense) and per-row squared norms: set by _prepare_wide_J
    jrn2 = None

    def __init__(self, jac_mode: int, x: Sequence[np.ndarray], y: Optional[Sequence[np.ndarray]], seed=None):
        import torch
        lib = _lib.load()
        dev = _device()

        def up(a):
            if isinstance(a, torch.Tensor):          # already on a device (construction/deviceData.py): no host round trip
                return a.to(device=dev, dtype=torch.float32).contiguous()
            a = np.asarray(a)
            if a.dtype != np.float32:
                a = a.astype(np.float32)
            return torch.from_numpy(np.ascontiguousarray(a)).to(dev)

        self.m = up(x[0])
        self.n = int(self.m.shape[0])
        self.U = self.V = self.q = self.J = self.qn2 = self.jn2 = None
        self.ld_J, self.jrn2 = 0, None
        self.seed = None if seed is None else up(seed)
        if jac_mode == _lib.JAC_SKETCHED:
            if len(x) != 3:
                raise ValueError("sketched-Jacobian models take inputs [m, U, V]")
            self.U, self.V = up(x[1]), up(x[2])
        elif len(x) != 1:
            raise ValueError("this model takes one input array [m]")
        if y is not None:
            self.q = up(y[0])
            self.qn2 = torch.empty(self.n, dtype=torch.float32, device=dev)
            _lib.check(lib.dino_sample_sqnorms(C.c_void_p(self.q.data_ptr()), self.n, self.q[0].numel(),
                                               C.c_void_p(self.qn2.data_ptr()), _stream()), "dino_sample_sqnorms")
            if jac_mode != _lib.JAC_NONE:
                self.J = up(y[1])
                self.jn2 = torch.empty(self.n, dtype=torch.float32, device=dev)
                _lib.check(lib.dino_sample_sqnorms(C.c_void_p(self.J.data_ptr()), self.n, self.J[0].numel(),
                                                   C.c_void_p(self.jn2.data_ptr()), _stream()), "dino_sample_sqnorms")
                self._prepare_wide_J(jac_mode)

    @classmethod
    def from_device(cls, jac_mode: int, m, q=None, J=None, seed=None, U=None, V=None):
        """A dataset over tensors that already live on the device (fp32, contiguous): nothing is copied."""
        import torch
        lib = _lib.load()
        ds = cls.__new__(cls)
        ds.m, ds.q, ds.J, ds.U, ds.V, ds.seed = m, q, J, U, V, seed
        ds.n = int(m.shape[0])
        ds.qn2 = ds.jn2 = None
        ds.ld_J, ds.jrn2 = 0, None
        for t in (m, q, J, U, V, seed):
            if t is not None and (t.dtype != torch.float32 or not t.is_contiguous() or not t.is_cuda):
                raise ValueError("DeviceDataset.from_device needs contiguous fp32 CUDA tensors")
        if q is not None:
            ds.qn2 = torch.empty(ds.n, dtype=torch.float32, device=m.device)
            _lib.check(lib.dino_sample_sqnorms(C.c_void_p(q.data_ptr()), ds.n, q[0].numel(), C.c_void_p(ds.qn2.data_ptr()),
                                               _stream()), "dino_sample_sqnorms")
        if J is not None and jac_mode != _lib.JAC_NONE:
            ds.jn2 = torch.empty(ds.n, dtype=torch.float32, device=m.device)
            _lib.check(lib.dino_sample_sqnorms(C.c_void_p(J.data_ptr()), ds.n, J[0].numel(), C.c_void_p(ds.jn2.data_ptr()),
                                               _stream()), "dino_sample_sqnorms")
            ds._prepare_wide_J(jac_mode)
        return ds

    def _prepare_wide_J(self, jac_mode: int):
        """Targets of a FULL Jacobian over an unreduced input (d_in in the thousands): rows on a 16-byte pitch (zero
        padded) and the squared norm of every row, which is what lets dino_h1_fwd_bwd take the first layer without
        forming J_hat (include/dino_b200.h: DinoBatch.ld_J, J_row_sqnorm).  The one case where a dataset
        built by from_device is copied: a row length that is not a multiple of 4 floats (d_M = 4225)."""
        import torch
        if jac_mode != _lib.JAC_FULL or self.J is None or self.J.dim() != 3:
            return
        d_in = int(self.J.shape[2])
        if d_in < int(os.environ.get("DINO_B200_IMPLICIT_MIN", "1024")) or os.environ.get("DINO_B200_NO_IMPLICIT0"):
            return
        ld = (d_in + 3) & ~3
        if ld != d_in:
            Jp = torch.zeros(self.J.shape[0], self.J.shape[1], ld, dtype=torch.float32, device=self.J.device)
            Jp[:, :, :d_in].copy_(self.J)
            self.J = Jp
        self.ld_J = ld
        rows = self.n * int(self.J.shape[1])
        self.jrn2 = torch.empty(rows, dtype=torch.float32, device=self.J.device)
        _lib.check(_lib.load().dino_sample_sqnorms(C.c_void_p(self.J.data_ptr()), rows, ld, C.c_void_p(self.jrn2.data_ptr()),
                                                   _stream()), "dino_sample_sqnorms")

    def batch(self, idx_ptr: Optional[int], row_offset: int = 0) -> _lib.DinoBatch:
        """DinoBatch for rows addressed by a device int32 index array (idx_ptr) or, when idx_ptr is
        None, the contiguous rows starting at row_offset."""
        def p(t, per_row=None):
            if t is None:
                return None
            if idx_ptr is None and row_offset:
                return t.data_ptr() + row_offset * t[0].numel() * 4 if t.dim() > 1 else t.data_ptr() + row_offset * 4
            return t.data_ptr()
        jrn2 = None
        if self.jrn2 is not None:
            jrn2 = self.jrn2.data_ptr() + (row_offset * int(self.J.shape[1]) * 4 if idx_ptr is None and row_offset else 0)
        return _lib.DinoBatch(p(self.m), p(self.q), p(self.J), p(self.qn2), p(self.jn2), idx_ptr,
                              p(self.U), p(self.V), None if self.seed is None else self.seed.data_ptr(),
                              self.ld_J, jrn2, self.n)


class TrainEngine:
    """Static device state of a training run and the CUDA graphs of its step.

    Everything a step consumes lives at fixed device addresses: the epoch's permutation, a cursor into
    it, the staged sample indices of the current minibatch, the statistics row, the flat gradient and
    the optimiser step counter.  The step (stage indices -> dino_h1_fwd_bwd -> [all-reduce] ->
    optimiser -> roll cursors) therefore has no host-side arguments that change from step to step and
    is captured once per minibatch size in a CUDA graph; afterwards a step costs the host one graph
    launch.  The first step of every size runs eagerly (it also triggers one-time kernel attribute
    set-up inside the library, which may not happen during capture)."""

    def __init__(self, model: "DinoModel", ds: "DeviceDataset", batch_size: int, use_graph: bool = True):
        import torch
        self.model, self.ds = model, ds
        self.dp = parallel.context()
        self.lib = _lib.load()
        core = model.core
        self.params = core.device_params()
        dev = self.params.device
        self.batch_size = int(batch_size)
        self.steps_cap = (ds.n + self.batch_size - 1) // self.batch_size
        local_max = -(-min(self.batch_size, ds.n) // self.dp.world)
        self.plan = model._plan(max(local_max, 1))
        model._ensure_opt_state()
        opt = model.optimizer
        if getattr(opt, "_t_dev", None) is None or opt._t_dev.device != dev:
            opt._t_dev = torch.tensor([opt.iterations], dtype=torch.int64, device=dev)
        self.idx_all = torch.zeros(ds.n, dtype=torch.int32, device=dev)
        self.cursors = torch.zeros(2, dtype=torch.int32, device=dev)          # [pos, step]
        self.idx_stage = torch.zeros(max(local_max, 1), dtype=torch.int32, device=dev)
        self.stats_stage = torch.zeros(_lib.STAT_SLOTS, dtype=torch.float32, device=dev)
        self.stats_hist = torch.zeros(self.steps_cap + 1, _lib.STAT_SLOTS, dtype=torch.float32, device=dev)
        self.grad = torch.zeros(core.n_params, dtype=torch.float32, device=dev)
        self.batch_struct = ds.batch(self.idx_stage.data_ptr())
        self.use_graph = use_graph and os.environ.get("DINO_B200_NO_GRAPH", "0") != "1"
        self.graphs = {}           # Bg -> (CUDAGraph, key, kernels per replay)
        self.replayed_kernels = 0  # kernels of ours executed through graph replays
        self.captured_kernels = 0  # kernels recorded during capture (the library counted them, they did not run)
        self.seen = set()
        self.host_pos = 0
        self.host_step = 0

    def set_order(self, order: np.ndarray):
        """Install the sample order of the next epoch and rewind the cursors."""
        import torch
        t = torch.from_numpy(np.ascontiguousarray(order, dtype=np.int32))
        self.idx_all[:t.numel()].copy_(t, non_blocking=False)
        self.cursors.zero_()
        self.stats_stage.zero_()
        self.n_order = int(t.numel())
        self.host_pos = 0
        self.host_step = 0

    def rewind(self):
        """Walk the installed order again from the start (asynchronous; no host->device traffic)."""
        self.cursors.zero_()
        self.stats_stage.zero_()
        self.host_pos = 0
        self.host_step = 0

    def remaining(self) -> int:
        return self.n_order - self.host_pos

    def _key(self):
        opt = self.model.optimizer
        return (opt.kind, float(opt.learning_rate), tuple(self.model.loss_weights))

    def _enqueue(self, Bg: int):
        lib, m = self.lib, self.model
        opt, P = m.optimizer, m.core.n_params
        st = _stream()
        off, Bl = parallel.shard_bounds(Bg, self.dp.rank, self.dp.world)
        pos_ptr, step_ptr = self.cursors.data_ptr(), self.cursors.data_ptr() + 4
        if Bl > 0:
            _lib.check(lib.dino_stage_batch(C.c_void_p(self.idx_all.data_ptr() + 4 * off), C.c_void_p(pos_ptr),
                                            C.c_void_p(self.idx_stage.data_ptr()), Bl, st), "dino_stage_batch")
            _lib.check(lib.dino_h1_fwd_bwd(self.plan.handle, C.c_void_p(self.params.data_ptr()),
                                           C.byref(self.batch_struct), Bl, m.loss_weights[0], m.loss_weights[1],
                                           1.0 / Bg, C.c_void_p(self.grad.data_ptr()),
                                           C.c_void_p(self.stats_stage.data_ptr()), st), "dino_h1_fwd_bwd")
        else:
            self.grad.zero_()
        if self.dp.world > 1:
            self.dp.all_reduce_sum(self.grad)
        if opt.kind == "adam":
            _lib.check(lib.dino_adam_dev(C.c_void_p(self.params.data_ptr()), C.c_void_p(self.grad.data_ptr()),
                                         C.c_void_p(opt._m.data_ptr()), C.c_void_p(opt._v.data_ptr()), P,
                                         opt.learning_rate, opt.beta_1, opt.beta_2, opt.epsilon,
                                         C.c_void_p(opt._t_dev.data_ptr()), st), "dino_adam_dev")
        else:
            _lib.check(lib.dino_sgd(C.c_void_p(self.params.data_ptr()), C.c_void_p(self.grad.data_ptr()), P,
                                    opt.learning_rate, st), "dino_sgd")
        _lib.check(lib.dino_roll_stats(C.c_void_p(self.stats_stage.data_ptr()), C.c_void_p(self.stats_hist.data_ptr()),
                                       C.c_void_p(pos_ptr), Bg, C.c_void_p(step_ptr),
                                       C.c_void_p(opt._t_dev.data_ptr()), st), "dino_roll_stats")

    def step(self, Bg: int = None):
        """One optimiser step on the next Bg samples of the installed order.  Asynchronous."""
        import torch
        Bg = self.batch_size if Bg is None else int(Bg)
        if Bg > self.remaining() or self.host_step >= self.steps_cap + 1:
            raise RuntimeError("TrainEngine.step: not enough samples left in the installed order")
        if self.plan.handle is None:
            # the network changed under the engine (layer.trainable toggled -> plans invalidated): take a fresh plan
            # and drop the graphs that replay the old workspace
            self.plan = self.model._plan(self.plan.max_batch)
            self.graphs.clear()
            self.seen.clear()
        key = self._key()
        entry = self.graphs.get(Bg)
        if entry is not None and entry[1] != key:
            entry = None
            self.graphs.pop(Bg)
        if entry is not None:
            entry[0].replay()
            self.replayed_kernels += entry[2]
        elif self.use_graph and (Bg, key) in self.seen:
            try:
                g = torch.cuda.CUDAGraph()
                cur = torch.cuda.current_stream()
                side = torch.cuda.Stream()
                side.wait_stream(cur)
                before = _lib.launch_count()
                with torch.cuda.graph(g, stream=side):
                    self._enqueue(Bg)
                cur.wait_stream(side)
                kernels = _lib.launch_count() - before          # our kernel nodes in the captured step
                self.graphs[Bg] = (g, key, kernels)
                self.captured_kernels += kernels                # counted by the library but not executed
                g.replay()
                self.replayed_kernels += kernels
            except _lib.DinoLibraryError:
                raise                           # a real error of the step itself: never swallow it
            except RuntimeError as exc:         # stream capture refused (e.g. a collective that cannot be captured)
                import warnings
                warnings.warn(f"dino_b200: CUDA-graph capture of the training step failed ({exc}); staying eager")
                self.use_graph = False
                torch.cuda.synchronize()
                self._enqueue(Bg)
        else:
            self._enqueue(Bg)
            self.seen.add((Bg, key))
        self.host_pos += Bg
        self.host_step += 1
        self.model.optimizer.iterations += 1
        self.model.core.mark_device_updated()

    def epoch_stats(self) -> np.ndarray:
        """Per-step statistics rows of the steps run since set_order (synchronises)."""
        if self.dp.world > 1:
            h = self.stats_hist[:self.host_step].clone()
            self.dp.all_reduce_sum(h)
            return h.cpu().numpy()
        return self.stats_hist[:self.host_step].cpu().numpy()


class History:
    def __init__(self):
        self.history: Dict[str, List[float]] = {}
        self.epoch: List[int] = []

    def _append(self, epoch, logs):
        self.epoch.append(epoch)
        for k, v in logs.items():
            self.history.setdefault(k, []).append(v)


def _combine_stats(stats: np.ndarray, weights, jac: bool):
    """Keras aggregation (TF-knowledge, Keras 2.7): the losses are batch-size-weighted means of the per-batch values
    (LossesContainer passes sample_weight = batch size); the function metrics l2_accuracy / f_accuracy_matrix return
    one scalar per batch and go through MeanMetricWrapper, so every batch counts ONCE, also a short last one.
    stats: [n_batches, STAT_SLOTS] per-batch sums from the loss kernels."""
    s = stats.astype(np.float64)
    cnt = s[:, _lib.STAT_COUNT]
    keep = cnt > 0
    s, cnt = s[keep], cnt[keep]
    wts = cnt / cnt.sum()
    uni = np.full(len(cnt), 1.0 / max(len(cnt), 1))
    l2 = s[:, _lib.STAT_L2_SUM] / cnt
    out_q = [l2, 1.0 - np.sqrt(l2), 1.0 - np.sqrt(s[:, _lib.STAT_EQ_SUM] / s[:, _lib.STAT_NQ_SUM])]
    if not jac:
        loss = weights[0] * l2
        return [float((loss * wts).sum()), float((out_q[1] * uni).sum())]
    h1 = s[:, _lib.STAT_H1_SUM] / cnt
    loss = weights[0] * l2 + weights[1] * h1
    accs = [out_q[1], out_q[2], 1.0 - np.sqrt(s[:, _lib.STAT_JROW_SUM] / s[:, _lib.STAT_JROWS]), 1.0 - np.sqrt(h1)]
    return [float((v * wts).sum()) for v in (loss, l2, h1)] + [float((v * uni).sum()) for v in accs]


_LOSS_NAMES = {"normalized_mse", "normalized_mse_matrix"}


class DinoModel:
    """Duck-typed ``tf.keras.Model`` with one input (or [m, U, V]) and one or two outputs."""

    def __init__(self, core: NetCore, jac_mode: str = "none", seed_dim: int = 0, output_basis=None,
                 precision: Optional[str] = None):
        self.core = core
        self.jac_mode_name = jac_mode
        self.jac_mode = _lib.JAC_MODES[jac_mode]
        self.seed_dim = {"none": 0, "full": core.d_out}.get(jac_mode, seed_dim)
        # Phi [d_Q, r_out] of equip_model_with_output_reduced_jacobian; the chain is seeded with Phi^T
        self.output_basis = None if output_basis is None else np.asarray(output_basis, dtype=np.float32)
        self.precision = precision or os.environ.get("DINO_B200_PRECISION", "fp32x3")
        self.optimizer = None
        self.loss_weights = [1.0, 1.0]
        self._compiled = False
        p = core.name_prefix
        self.inputs = [TensorSpec((None, core.d_in), p + "input_data")]
        self.outputs = [TensorSpec((None, core.d_out), core.names[-1])]
        if jac_mode == "sketched":
            self.inputs += [TensorSpec((None, core.d_out, seed_dim), p + "_input_U"),
                            TensorSpec((None, core.d_in, seed_dim), p + "_input_V")]
            self.outputs.append(TensorSpec((None, seed_dim, seed_dim), p + "reduced_jacobian"))
        elif jac_mode != "none":
            self.outputs.append(TensorSpec((None, self.seed_dim, core.d_in), p + "jacobian"))
        self.stop_training = False

    # ---- Keras structural surface ----------------------------------------------------------------
    @property
    def layers(self):
        return self.core.keras_layers

    def get_layer(self, name=None, index=None):
        if index is not None:
            return self.core.keras_layers[index]
        for layer in self.core.keras_layers:
            if layer.name == name:
                return layer
        raise ValueError(f"No such layer: {name}.")

    @property
    def input(self):
        return self.inputs[0] if len(self.inputs) == 1 else self.inputs

    @property
    def output(self):
        return self.outputs[0] if len(self.outputs) == 1 else self.outputs

    @property
    def input_shape(self):
        return self.inputs[0].shape

    @property
    def output_shape(self):
        return self.outputs[0].shape if len(self.outputs) == 1 else [o.shape for o in self.outputs]

    def get_weights(self):
        out = []
        for layer in self.layers:
            out += layer.get_weights()
        return out

    def set_weights(self, weights):
        i = 0
        for layer in self.layers:
            n = len(layer.get_weights())
            layer.set_weights(weights[i:i + n])
            i += n

    def count_params(self):
        return sum(layer.count_params() for layer in self.layers)

    def summary(self, print_fn=print):
        lines = [f"Model: dino_b200 ({self.jac_mode_name} Jacobian output)" if self.jac_mode else "Model: dino_b200",
                 "_" * 65, f"{'Layer':<40}{'Output Shape':<15}{'Param #':>10}", "=" * 65]
        c = self.core
        lines.append(f"{c.keras_layers[0].name:<40}{str((None, c.d_in)):<15}{0:>10}")
        for layer in c.keras_layers[1:]:
            shape = (None, c.n_out[layer._index]) if layer._index is not None else ""
            lines.append(f"{layer.name:<40}{str(shape):<15}{layer.count_params():>10}")
        total = self.count_params()
        train = sum(l.count_params() for l in c.keras_layers if l._index is not None and l.trainable)
        lines += ["=" * 65, f"Total params: {total}", f"Trainable params: {train}",
                  f"Non-trainable params: {total - train}"]
        for ln in lines:
            print_fn(ln)

    # ---- compile / metrics -------------------------------------------------------------------------
    def compile(self, optimizer="adam", loss=None, loss_weights=None, metrics=None, **_ignored):
        losses = loss if isinstance(loss, (list, tuple)) else [loss]
        for f in losses:
            name = f if isinstance(f, str) else getattr(f, "__name__", None)
            if name not in _LOSS_NAMES:
                raise NotImplementedError(f"loss {name!r}: only normalized_mse / normalized_mse_matrix "
                                          "(trainingUtilities.py:25-37) exist as kernels")
        if len(losses) != len(self.outputs):
            raise ValueError(f"{len(losses)} losses for {len(self.outputs)} model outputs")
        self.optimizer = _as_optimizer(optimizer)
        if loss_weights is None:
            loss_weights = [1.0] * len(self.outputs)
        self.loss_weights = [float(w) for w in loss_weights] + [0.0] * (2 - len(loss_weights))
        self._compiled = True

    @property
    def metrics_names(self):
        o = [t.name for t in self.outputs]
        if len(o) == 1:
            return ["loss", "l2_accuracy"]
        return ["loss", o[0] + "_loss", o[1] + "_loss", o[0] + "_l2_accuracy", o[0] + "_f_accuracy_matrix",
                o[1] + "_l2_accuracy", o[1] + "_f_accuracy_matrix"]

    # ---- device plumbing ---------------------------------------------------------------------------
    def _plan(self, max_batch: int) -> _Plan:
        key = (self.jac_mode, self.seed_dim, _lib.PRECISIONS[self.precision])
        plan = self.core._plans.get(key)
        if plan is None or plan.handle is None or plan.max_batch < max_batch:
            # A larger plan REPLACES the cached one but never destroys it: a live TrainEngine (or a captured CUDA
            # graph of fit_generator) may still hold the smaller plan and its workspace; that object is released
            # when its last user lets go of it (_Plan.__del__).
            plan = _Plan(self.core, self.jac_mode, self.seed_dim, max_batch, _lib.PRECISIONS[self.precision])
            self.core._plans[key] = plan
        return plan

    def _dataset(self, x, y):
        x = list(x) if isinstance(x, (list, tuple)) else [x]
        if y is not None:
            y = list(y) if isinstance(y, (list, tuple)) else [y]
            if len(y) != len(self.outputs):
                raise ValueError(f"{len(y)} target arrays for {len(self.outputs)} model outputs")
        seed = None if self.output_basis is None else np.ascontiguousarray(self.output_basis.T)
        return DeviceDataset(self.jac_mode, x, y, seed)

    # ---- evaluate / predict ------------------------------------------------------------------------
    def evaluate(self, x, y, verbose=0, batch_size=32, **_ignored):
        import torch
        ds = x if isinstance(x, DeviceDataset) else self._dataset(x, y)
        stats = self._eval_stats(ds, batch_size)
        vals = _combine_stats(stats, self.loss_weights, self.jac_mode != _lib.JAC_NONE)
        if verbose:
            print(" - ".join(f"{k}: {v:.4e}" for k, v in zip(self.metrics_names, vals)))
        return vals

    def _eval_stats(self, ds: DeviceDataset, batch_size: int) -> np.ndarray:
        import torch
        lib = _lib.load()
        plan = self._plan(min(batch_size, ds.n))
        params = self.core.device_params()
        nb = (ds.n + batch_size - 1) // batch_size
        stats = torch.zeros(nb, _lib.STAT_SLOTS, dtype=torch.float32, device=params.device)
        st = _stream()
        for i in range(nb):
            lo = i * batch_size
            B = min(batch_size, ds.n - lo)
            batch = ds.batch(None, lo)
            _lib.check(lib.dino_h1_eval(plan.handle, C.c_void_p(params.data_ptr()), C.byref(batch), B,
                                        C.c_void_p(stats[i].data_ptr()), st), "dino_h1_eval")
        return stats.cpu().numpy()

    def predict(self, x, batch_size=32, **_ignored):
        out = self.predict_device(x, batch_size)
        if isinstance(out, list):
            return [t.cpu().numpy() for t in out]
        return out.cpu().numpy()

    def predict_device(self, x, batch_size=32):
        """``predict`` whose results stay on the device (torch tensors); ``x`` may be a DeviceDataset."""
        import torch
        lib = _lib.load()
        ds = x if isinstance(x, DeviceDataset) else self._dataset(x, None)
        plan = self._plan(min(batch_size, ds.n))
        params = self.core.device_params()
        q = torch.empty(ds.n, self.core.d_out, dtype=torch.float32, device=params.device)
        J = None
        if self.jac_mode != _lib.JAC_NONE:
            J = torch.empty((ds.n,) + tuple(self.outputs[1].shape[1:]), dtype=torch.float32, device=params.device)
        st = _stream()
        for lo in range(0, ds.n, batch_size):
            B = min(batch_size, ds.n - lo)
            batch = ds.batch(None, lo)
            if J is None:
                _lib.check(lib.dino_forward(plan.handle, C.c_void_p(params.data_ptr()), C.byref(batch), B,
                                            C.c_void_p(q[lo].data_ptr()), st), "dino_forward")
            else:
                _lib.check(lib.dino_jacobian(plan.handle, C.c_void_p(params.data_ptr()), C.byref(batch), B,
                                             C.c_void_p(q[lo].data_ptr()), C.c_void_p(J[lo].data_ptr()), st),
                           "dino_jacobian")
        if J is None:
            return q
        return [q, J]

    def jvp(self, x, mhat, batch_size=32):
        """J_hat(m) mhat for every row of x: one forward-mode sweep (dino_jvp).  mhat: [d_in] or [1, d_in] (shared
        direction) or [n, d_in] (one per sample).  Returns [n, d_out]."""
        import torch
        lib = _lib.load()
        ds = x if isinstance(x, DeviceDataset) else self._dataset(x, None)
        mh = np.ascontiguousarray(np.atleast_2d(np.asarray(mhat, dtype=np.float32)))
        if mh.shape[1] != self.core.d_in or mh.shape[0] not in (1, ds.n):
            raise ValueError(f"mhat has shape {mh.shape}; expected [1 or {ds.n}, {self.core.d_in}]")
        plan = self._plan(min(batch_size, ds.n))
        params = self.core.device_params()
        mh_d = torch.from_numpy(mh).to(params.device)
        out = torch.empty(ds.n, self.core.d_out, dtype=torch.float32, device=params.device)
        st = _stream()
        for lo in range(0, ds.n, batch_size):
            B = min(batch_size, ds.n - lo)
            batch = ds.batch(None, lo)
            shared = mh.shape[0] == 1 and ds.n != 1
            rows, src = (1, mh_d) if shared else (B, mh_d[lo])
            _lib.check(lib.dino_jvp(plan.handle, C.c_void_p(params.data_ptr()), C.byref(batch), B, C.c_void_p(src.data_ptr()),
                                    rows, C.c_void_p(out[lo].data_ptr()), st), "dino_jvp")
        return out.cpu().numpy()

    # ---- one optimiser step ------------------------------------------------------------------------
    def _ensure_opt_state(self):
        import torch
        opt, P = self.optimizer, self.core.n_params
        dev = self.core.device_params().device
        if opt.kind == "adam" and (opt._m is None or opt._m.numel() != P):
            opt._m = torch.zeros(P, dtype=torch.float32, device=dev)
            opt._v = torch.zeros(P, dtype=torch.float32, device=dev)

    def _train_step(self, plan, batch, Bl, Bg, grad, stats_row, dp, st):
        """fwd + Jacobian + losses + double backward on this rank's Bl samples of a global minibatch
        of Bg, gradient all-reduce (world > 1), optimiser update.  Asynchronous: no host sync."""
        lib = plan.lib
        params = self.core.device_params()
        opt, P = self.optimizer, self.core.n_params
        if Bl > 0:
            _lib.check(lib.dino_h1_fwd_bwd(plan.handle, C.c_void_p(params.data_ptr()), C.byref(batch), Bl,
                                           self.loss_weights[0], self.loss_weights[1], 1.0 / Bg,
                                           C.c_void_p(grad.data_ptr()), C.c_void_p(stats_row.data_ptr()), st),
                       "dino_h1_fwd_bwd")
        else:
            grad.zero_()
        if dp.world > 1:
            dp.all_reduce_sum(grad)
        opt.iterations += 1
        if getattr(opt, "_t_dev", None) is not None:
            opt._t_dev += 1
        if opt.kind == "adam":
            _lib.check(lib.dino_adam(C.c_void_p(params.data_ptr()), C.c_void_p(grad.data_ptr()),
                                     C.c_void_p(opt._m.data_ptr()), C.c_void_p(opt._v.data_ptr()), P,
                                     opt.learning_rate, opt.beta_1, opt.beta_2, opt.epsilon, opt.iterations, st),
                       "dino_adam")
        else:
            _lib.check(lib.dino_sgd(C.c_void_p(params.data_ptr()), C.c_void_p(grad.data_ptr()), P,
                                    opt.learning_rate, st), "dino_sgd")
        self.core.mark_device_updated()

    def make_dataset(self, x, y):
        """Upload (x, y) to HBM once; the result can be passed to fit / evaluate / train_on_indices."""
        return self._dataset(x, y)

    def make_engine(self, ds, batch_size, use_graph=True) -> TrainEngine:
        """A TrainEngine over a resident dataset: ``set_order(perm)`` then ``step()`` per minibatch."""
        if not self._compiled:
            raise RuntimeError("You must compile your model before training/testing.")
        return TrainEngine(self, ds, batch_size, use_graph)

    def train_on_batch(self, x, y, **_ignored):
        """Keras ``train_on_batch``: one optimiser step on HOST arrays.  Under data parallelism each
        rank passes ITS shard of the global minibatch (equal shard sizes).  The shard is copied
        host->device (pinned torch tensors are copied asynchronously), the step runs, and the
        step's [loss, l2, h1, metrics...] are read back (device->host), like Keras does."""
        import torch
        if not self._compiled:
            raise RuntimeError("You must compile your model before training/testing.")
        lib = _lib.load()
        dp = parallel.context()
        x = list(x) if isinstance(x, (list, tuple)) else [x]
        y = list(y) if isinstance(y, (list, tuple)) else [y]
        dev = self.core.device_params().device
        stage = getattr(self, "_stage", None)
        if stage is None:
            stage = self._stage = {}

        def up(key, a):
            t = a if isinstance(a, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32))
            buf = stage.get(key)
            if buf is None or buf.shape != t.shape:
                buf = stage[key] = torch.empty(t.shape, dtype=torch.float32, device=dev)
            buf.copy_(t, non_blocking=True)
            return buf

        ds = DeviceDataset.__new__(DeviceDataset)
        ds.m = up("m", x[0]); ds.n = Bl = int(ds.m.shape[0])
        ds.U = ds.V = ds.J = ds.jn2 = ds.seed = None
        if self.jac_mode == _lib.JAC_SKETCHED:
            ds.U, ds.V = up("U", x[1]), up("V", x[2])
        if self.output_basis is not None:
            ds.seed = up("seed", np.ascontiguousarray(self.output_basis.T))
        ds.q = up("q", y[0])
        st = _stream()
        norms = stage.get("norms")
        if norms is None or norms.shape[1] != Bl:
            norms = stage["norms"] = torch.empty(2, Bl, dtype=torch.float32, device=dev)
            stage["grad"] = torch.zeros(self.core.n_params + _lib.STAT_SLOTS, dtype=torch.float32, device=dev)
            stage["stats"] = stage["grad"][self.core.n_params:]          # statistics at the tail: one all-reduce carries both
        ds.qn2 = norms[0]
        _lib.check(lib.dino_sample_sqnorms(C.c_void_p(ds.q.data_ptr()), Bl, ds.q[0].numel(),
                                           C.c_void_p(ds.qn2.data_ptr()), st), "dino_sample_sqnorms")
        if self.jac_mode != _lib.JAC_NONE:
            ds.J = up("J", y[1])
            ds.jn2 = norms[1]
            _lib.check(lib.dino_sample_sqnorms(C.c_void_p(ds.J.data_ptr()), Bl, ds.J[0].numel(),
                                               C.c_void_p(ds.jn2.data_ptr()), st), "dino_sample_sqnorms")
        stats = stage["stats"]
        stats.zero_()
        plan = self._plan(Bl)
        self._ensure_opt_state()
        self._train_step(plan, ds.batch(None), Bl, Bl * dp.world, stage["grad"], stats, dp, st)   # (its all-reduce covers the tail)
        host = stats.cpu().numpy()[None, :]          # device->host read of the step's result
        self.last_h2d_bytes = sum(int(t.numel()) * 4 for k, t in stage.items() if k in ("m", "q", "J", "U", "V"))
        self.last_d2h_bytes = int(stats.numel()) * 4
        return _combine_stats(host, self.loss_weights, self.jac_mode != _lib.JAC_NONE)

    def fit_generator(self, generator, steps_per_epoch=None, **_ignored):
        """Keras ``fit_generator`` for ONE pass over ``generator``: every item is ``(x, y)`` of HOST arrays
        (pinned torch tensors are copied asynchronously), one optimiser step per item, like ``train_on_batch``.
        The input pipeline is overlapped the way ``tf.data`` prefetching does it: the host->device copy of
        minibatch k+1 runs on a copy stream while minibatch k is computed, and the [loss, ...] row of step k
        is read back (device->host) after step k+1 has been queued.  Returns the list of per-step rows."""
        import torch
        if not self._compiled:
            raise RuntimeError("You must compile your model before training/testing.")
        lib = _lib.load()
        dp = parallel.context()
        dev = self.core.device_params().device
        jac = self.jac_mode != _lib.JAC_NONE
        compute = torch.cuda.current_stream()
        pipe = getattr(self, "_pipe", None)
        if pipe is None:
            pipe = self._pipe = {"copy": torch.cuda.Stream(), "sets": [{}, {}], "grad": None}
        copy_stream = pipe["copy"]
        if pipe["grad"] is None:
            # the step's statistics ride at the tail of the gradient buffer: ONE all-reduce per step carries both
            pipe["grad"] = torch.zeros(self.core.n_params + _lib.STAT_SLOTS, dtype=torch.float32, device=dev)
        stats_tail = pipe["grad"][self.core.n_params:]
        self._ensure_opt_state()
        st = _stream()
        rows, pending, h2d = [], None, 0

        def stage(S, key, a):
            t = a if isinstance(a, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32))
            buf = S.get(key)
            if buf is None or buf.shape != t.shape:
                buf = S[key] = torch.empty(t.shape, dtype=torch.float32, device=dev)
            buf.copy_(t, non_blocking=True)
            return buf

        for k, (x, y) in enumerate(generator):
            if steps_per_epoch is not None and k >= steps_per_epoch:
                break
            x = list(x) if isinstance(x, (list, tuple)) else [x]
            y = list(y) if isinstance(y, (list, tuple)) else [y]
            S = pipe["sets"][k & 1]
            # ---- copy stream: this minibatch's inputs, once the step that last used this buffer set is done ----
            if "consumed" in S:
                copy_stream.wait_event(S["consumed"])
            with torch.cuda.stream(copy_stream):
                ds = DeviceDataset.__new__(DeviceDataset)
                ds.m = stage(S, "m", x[0]); ds.n = Bl = int(ds.m.shape[0])
                ds.U = ds.V = ds.J = ds.jn2 = ds.seed = None
                if self.jac_mode == _lib.JAC_SKETCHED:
                    ds.U, ds.V = stage(S, "U", x[1]), stage(S, "V", x[2])
                ds.q = stage(S, "q", y[0])
                if jac:
                    ds.J = stage(S, "J", y[1])
                S.setdefault("copied", torch.cuda.Event()).record(copy_stream)
            h2d = sum(int(t.numel()) * 4 for kk, t in S.items() if kk in ("m", "q", "J", "U", "V"))
            if self.output_basis is not None:
                if "seed" not in pipe:
                    pipe["seed"] = torch.from_numpy(np.ascontiguousarray(self.output_basis.T, dtype=np.float32)).to(dev)
                ds.seed = pipe["seed"]
            # ---- compute stream: norms, the step, the result row into pinned host memory ----
            compute.wait_event(S["copied"])
            if "norms" not in S or S["norms"].shape[1] != Bl:
                S["norms"] = torch.empty(2, Bl, dtype=torch.float32, device=dev)
                S["host"] = torch.zeros(_lib.STAT_SLOTS, dtype=torch.float32).pin_memory()
                S.pop("graph", None)
            opt = self.optimizer
            if getattr(opt, "_t_dev", None) is None or opt._t_dev.device != dev:
                opt._t_dev = torch.tensor([opt.iterations], dtype=torch.int64, device=dev)
            plan = self._plan(Bl)

            def enqueue():
                # everything of one step that does not depend on the host: capturable, replayable
                ds.qn2 = S["norms"][0]
                _lib.check(lib.dino_sample_sqnorms(C.c_void_p(ds.q.data_ptr()), Bl, ds.q[0].numel(),
                                                   C.c_void_p(ds.qn2.data_ptr()), _stream()), "dino_sample_sqnorms")
                if jac:
                    ds.jn2 = S["norms"][1]
                    _lib.check(lib.dino_sample_sqnorms(C.c_void_p(ds.J.data_ptr()), Bl, ds.J[0].numel(),
                                                       C.c_void_p(ds.jn2.data_ptr()), _stream()), "dino_sample_sqnorms")
                stats_tail.zero_()
                params = self.core.device_params()
                _lib.check(lib.dino_h1_fwd_bwd(plan.handle, C.c_void_p(params.data_ptr()), C.byref(ds.batch(None)), Bl,
                                               self.loss_weights[0], self.loss_weights[1], 1.0 / (Bl * dp.world),
                                               C.c_void_p(pipe["grad"].data_ptr()), C.c_void_p(stats_tail.data_ptr()),
                                               _stream()), "dino_h1_fwd_bwd")
                if dp.world > 1:
                    dp.all_reduce_sum(pipe["grad"])
                if opt.kind == "adam":
                    _lib.check(lib.dino_adam_dev(C.c_void_p(params.data_ptr()), C.c_void_p(pipe["grad"].data_ptr()),
                                                 C.c_void_p(opt._m.data_ptr()), C.c_void_p(opt._v.data_ptr()),
                                                 self.core.n_params, opt.learning_rate, opt.beta_1, opt.beta_2, opt.epsilon,
                                                 C.c_void_p(opt._t_dev.data_ptr()), _stream()), "dino_adam_dev")
                else:
                    _lib.check(lib.dino_sgd(C.c_void_p(params.data_ptr()), C.c_void_p(pipe["grad"].data_ptr()),
                                            self.core.n_params, opt.learning_rate, _stream()), "dino_sgd")
                opt._t_dev += 1
                S["host"].copy_(stats_tail, non_blocking=True)      # device->host read of the step's result

            gkey = (Bl, opt.kind, float(opt.learning_rate), tuple(self.loss_weights), self.core.device_params().data_ptr(),
                    id(plan))
            entry = S.get("graph")
            if entry is not None and entry[1] != gkey:
                entry = None
            use_graph = os.environ.get("DINO_B200_NO_GRAPH", "0") != "1" and not pipe.get("no_graph", False)
            if entry is not None:
                entry[0].replay()
            elif use_graph and S.get("seen") == gkey:
                try:                                    # second use of this buffer set with this shape: capture the step
                    g = torch.cuda.CUDAGraph()
                    side = torch.cuda.Stream()
                    side.wait_stream(compute)
                    with torch.cuda.graph(g, stream=side):
                        enqueue()
                    compute.wait_stream(side)
                    S["graph"] = (g, gkey, plan)          # the graph replays this plan's workspace: keep it alive
                    g.replay()
                except _lib.DinoLibraryError:
                    raise
                except RuntimeError as exc:
                    import warnings
                    warnings.warn(f"dino_b200: CUDA-graph capture of the pipelined step failed ({exc}); staying eager")
                    pipe["no_graph"] = True
                    torch.cuda.synchronize()
                    enqueue()
            else:
                enqueue()
                S["seen"] = gkey
            opt.iterations += 1
            self.core.mark_device_updated()
            S.setdefault("consumed", torch.cuda.Event()).record(compute)
            S.setdefault("done", torch.cuda.Event()).record(compute)
            # ---- the previous step's row is on the host by now (or soon): collect it ----
            if pending is not None:
                pending["done"].synchronize()
                rows.append(_combine_stats(pending["host"].numpy().copy()[None, :], self.loss_weights, jac))
            pending = S
        if pending is not None:
            pending["done"].synchronize()
            rows.append(_combine_stats(pending["host"].numpy().copy()[None, :], self.loss_weights, jac))
        self.last_h2d_bytes = h2d
        self.last_d2h_bytes = _lib.STAT_SLOTS * 4
        return rows

    # ---- fit -----------------------------------------------------------------------------------------
    def fit(self, x, y, validation_data=None, epochs=1, batch_size=32, verbose=0, callbacks=None,
            shuffle=True, initial_epoch=0, perms=None, **_ignored):
        """Keras ``fit`` semantics (trainingUtilities.py:135-138): per epoch a permutation drawn
        from numpy's global RNG (``perms[e]`` overrides it), minibatches of ``batch_size`` in that
        order, a short last batch averaged over its own size, validation after every epoch.
        With torch.distributed initialised every global minibatch is split contiguously over the
        ranks and the flat gradient is sum-all-reduced before the (replicated) optimiser step."""
        if not self._compiled:
            raise RuntimeError("You must compile your model before training/testing.")
        ds = x if isinstance(x, DeviceDataset) else self._dataset(x, y)
        val = None
        if validation_data is not None:
            val = self._dataset(validation_data[0], validation_data[1])
        dp = parallel.context()
        n = ds.n
        eng = TrainEngine(self, ds, batch_size)
        steps = (n + batch_size - 1) // batch_size
        hist = History()
        names = self.metrics_names
        jac = self.jac_mode != _lib.JAC_NONE
        callbacks = callbacks or []
        for cb in callbacks:
            if hasattr(cb, "set_model"):
                cb.set_model(self)
            if hasattr(cb, "on_train_begin"):
                cb.on_train_begin({})
        for epoch in range(initial_epoch, epochs):
            if perms is not None:
                order = np.asarray(perms[epoch - initial_epoch])
            elif shuffle:
                order = np.random.permutation(n)
            else:
                order = np.arange(n)
            if dp.world > 1:
                order = dp.broadcast_order(order)
            eng.set_order(order)
            for cb in callbacks:
                if hasattr(cb, "on_epoch_begin"):
                    cb.on_epoch_begin(epoch, {})
            for step in range(steps):
                eng.step(min(batch_size, n - step * batch_size))
            vals = _combine_stats(eng.epoch_stats(), self.loss_weights, jac)
            logs = dict(zip(names, vals))
            if val is not None:
                vv = _combine_stats(self._eval_stats(val, batch_size), self.loss_weights, jac)
                logs.update({"val_" + k: v for k, v in zip(names, vv)})
            hist._append(epoch, logs)
            if verbose and dp.rank == 0:
                print(f"Epoch {epoch + 1}/{epochs} - " + " - ".join(f"{k}: {v:.4e}" for k, v in logs.items()))
            for cb in callbacks:
                if hasattr(cb, "on_epoch_end"):
                    cb.on_epoch_end(epoch, logs)
            if self.stop_training:
                break
        for cb in callbacks:
            if hasattr(cb, "on_train_end"):
                cb.on_train_end({})
        self.history = hist
        return hist
