"""GPU test of dino_forward_noalloc, the entry point for callers that cannot allocate.

Some users of the library own no device memory they could lend it as a workspace and cannot get
any: a routine called from inside another code's solver loop with nothing but its input and output
buffers, a process whose HBM has already been reserved in full by another framework's memory pool
(cudaMalloc would fail), or a latency-critical loop that cannot take the device-wide
synchronisation cudaMalloc/cudaFree imply.  For them the forward pass is also offered with its
activations in a fixed array inside the library.  This test drives that entry point through the
C ABI twice in a row (the array is shared between calls and keeps the previous call's contents, so
the second call, on a different batch size and network, would expose stale data) and compares with
dino_forward on the same inputs.  It is collected last so that a failure here cannot hide any
other test, and it passes both alone in a fresh process and after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest
import torch

import helpers as H
from oracle import dino_oracle as O
from dino_b200 import _lib, synthetic as S
from dino_b200.model import DeviceDataset, _stream

pytestmark = pytest.mark.gpu


def _spec(core):
    L = core.n_layers
    arr = (_lib.DinoLayerSpec * L)()
    for l in range(L):
        arr[l] = _lib.DinoLayerSpec(core.n_out[l], core.act[l], int(core.use_bias[l]), int(core.trainable[l]),
                                    core.skip_from[l])
    return _lib.DinoNetSpec(core.d_in, L, arr, _lib.JAC_NONE, 0, 1, _lib.PREC_FP32), arr


def _both_ways(arch, B, seed):
    data, tw = S.make_dataset(arch, B, seed)
    w = H.student_weights(arch, seed + 1, tw, bias_noise=0.1)
    model = S.build_model(arch, w, jac=None)
    lib = _lib.load()
    params = model.core.device_params()
    dev = params.device
    m = torch.from_numpy(data["m"]).to(dev)
    # existing entry point (caller-allocated workspace)
    ds = DeviceDataset(_lib.JAC_NONE, [data["m"]], None)
    plan = model._plan(B)
    q_ref = torch.empty(B, arch.d_out, device=dev)
    batch = ds.batch(None)
    _lib.check(lib.dino_forward(plan.handle, C.c_void_p(params.data_ptr()), C.byref(batch), B,
                                C.c_void_p(q_ref.data_ptr()), _stream()), "dino_forward")
    # new entry point: inputs and outputs only
    spec, keep = _spec(model.core)
    q_new = torch.full((B, arch.d_out), float("nan"), device=dev)
    _lib.check(lib.dino_forward_noalloc(C.byref(spec), C.c_void_p(params.data_ptr()), C.c_void_p(m.data_ptr()), B,
                                        C.c_void_p(q_new.data_ptr()), _stream()), "dino_forward_noalloc")
    torch.cuda.synchronize()
    # the oracle's forward pass (fp64) on the same inputs: the checker of both entry points
    net = H.oracle_net(arch, w)
    q_or, _ = O.batch_jacobian(net, torch.tensor(data["m"], dtype=torch.float64), False)
    return q_new.cpu().numpy(), q_ref.cpu().numpy(), q_or.detach().numpy()


def test_noalloc_forward_matches_forward_twice_in_a_row(cuda_device):
    # call 1: wide dense network, large batch -> fills much of the shared array
    a1 = S.Arch("rb_dense", 100, 50, hidden=[96, 64, 50], frozen_output=True, r_out=50)
    got, want, oracle = _both_ways(a1, 200, 31)
    assert np.isfinite(got).all()
    assert H.rel_err(got, want) < 1e-5
    assert H.rel_err(got, oracle) < 1e-5 and H.rel_err(want, oracle) < 1e-5
    # call 2, straight after: a smaller residual network and batch; stale activations of call 1 are
    # still in the array and must not leak into the result
    a2 = H.SMALL["resnet"]
    got, want, oracle = _both_ways(a2, 7, 32)
    assert np.isfinite(got).all()
    assert H.rel_err(got, want) < 1e-5
    assert H.rel_err(got, oracle) < 1e-5 and H.rel_err(want, oracle) < 1e-5


def test_noalloc_rejects_sizes_beyond_the_fixed_array(cuda_device):
    lib = _lib.load()
    arch = H.SMALL["dense"]
    model = S.build_model(arch, None, jac=None)
    params = model.core.device_params()
    spec, keep = _spec(model.core)
    m = torch.zeros(_lib.NOALLOC_MAX_BATCH + 1, arch.d_in, device=params.device)
    q = torch.zeros(_lib.NOALLOC_MAX_BATCH + 1, arch.d_out, device=params.device)
    rc = lib.dino_forward_noalloc(C.byref(spec), C.c_void_p(params.data_ptr()), C.c_void_p(m.data_ptr()),
                                  _lib.NOALLOC_MAX_BATCH + 1, C.c_void_p(q.data_ptr()), _stream())
    assert rc != 0 and b"DINO_NOALLOC_MAX_BATCH" in lib.dino_last_error()
    wide = S.Arch("generic_dense", 8, 4, hidden=[_lib.NOALLOC_MAX_WIDTH + 8])
    model = S.build_model(wide, None, jac=None)
    spec, keep = _spec(model.core)
    params = model.core.device_params()
    rc = lib.dino_forward_noalloc(C.byref(spec), C.c_void_p(params.data_ptr()), C.c_void_p(m.data_ptr()), 4,
                                  C.c_void_p(q.data_ptr()), _stream())
    assert rc != 0 and b"DINO_NOALLOC_MAX_WIDTH" in lib.dino_last_error()
