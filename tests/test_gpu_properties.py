"""-m gpu: size-independent properties of the H1 step at BASELINE.json's full cfg2 size (per-GPU minibatch 256),
where the fp64 oracle is too slow to be the checker: additivity of the gradient over a split minibatch,
agreement of the three entry points on the same statistics, and agreement of the two engines."""
import ctypes as C

import numpy as np
import pytest
import torch

import helpers as H
from dino_b200 import _lib, synthetic as S
from dino_b200.model import DeviceDataset, _stream

pytestmark = pytest.mark.gpu

B_FULL = 256


def _setup(precision):
    arch = S.CONFIGS["cfg2"]["arch"]
    data, tw = S.make_dataset(arch, B_FULL, 31)
    w = H.student_weights(arch, 32, tw, bias_noise=0.1)
    model = S.build_model(arch, w, precision=precision)
    ds = DeviceDataset(model.jac_mode, [data["m"]], [data["q"], data["J"]])
    return arch, data, model, ds


def _step(model, ds, idx, inv_gb, B_plan=B_FULL):
    """dino_h1_fwd_bwd on the samples `idx`; returns (gradient, statistics) as numpy."""
    lib = _lib.load()
    plan = model._plan(B_plan)
    params = model.core.device_params()
    grad = torch.zeros(model.core.n_params, device=params.device)
    stats = torch.zeros(_lib.STAT_SLOTS, device=params.device)
    didx = torch.tensor(np.asarray(idx), dtype=torch.int32, device=params.device)
    batch = ds.batch(didx.data_ptr())
    _lib.check(lib.dino_h1_fwd_bwd(plan.handle, C.c_void_p(params.data_ptr()), C.byref(batch), len(idx), 1.0, 1.0,
                                   inv_gb, C.c_void_p(grad.data_ptr()), C.c_void_p(stats.data_ptr()), _stream()), "step")
    torch.cuda.synchronize()
    return grad.cpu().numpy().astype(np.float64), stats.cpu().numpy().astype(np.float64)


@pytest.mark.parametrize("precision,tol", [("fp32x3", 2e-5), ("fp32", 2e-5), ("tf32", 2e-5)])
def test_gradient_is_additive_over_a_split_minibatch(cuda_device, precision, tol):
    """grad(A u B) = grad(A) + grad(B) with the same 1/B_global scaling: what the data-parallel all-reduce relies on.
    Each sample goes through the same arithmetic whichever tile it lands in, so only the summation order differs."""
    arch, data, model, ds = _setup(precision)
    idx = np.random.default_rng(3).permutation(B_FULL)
    g_full, s_full = _step(model, ds, idx, 1.0 / B_FULL)
    g_a, s_a = _step(model, ds, idx[:100], 1.0 / B_FULL)
    g_b, s_b = _step(model, ds, idx[100:], 1.0 / B_FULL)
    assert np.isfinite(g_full).all() and np.abs(g_full).max() > 0
    assert H.rel_err(g_a + g_b, g_full) < tol
    np.testing.assert_allclose(s_a + s_b, s_full, rtol=1e-4, atol=1e-6)


@pytest.mark.parametrize("precision,tol", [("fp32x3", 1e-5), ("fp32", 1e-5), ("tf32", 1e-5)])
def test_three_entry_points_agree_on_the_statistics(cuda_device, precision, tol):
    """The H1 / L2 sums of dino_h1_eval, of dino_h1_fwd_bwd and the ones recomputed on the host from
    dino_jacobian's output are the same numbers (trainingUtilities.py:25-37 on the same predictions)."""
    arch, data, model, ds = _setup(precision)
    lib = _lib.load()
    idx = np.arange(B_FULL)
    _, s_train = _step(model, ds, idx, 1.0 / B_FULL)
    plan = model._plan(B_FULL)
    params = model.core.device_params()
    stats = torch.zeros(_lib.STAT_SLOTS, device=params.device)
    didx = torch.tensor(idx, dtype=torch.int32, device=params.device)
    _lib.check(lib.dino_h1_eval(plan.handle, C.c_void_p(params.data_ptr()), C.byref(ds.batch(didx.data_ptr())), B_FULL,
                                C.c_void_p(stats.data_ptr()), _stream()), "eval")
    s_eval = stats.cpu().numpy().astype(np.float64)
    for slot in (_lib.STAT_L2_SUM, _lib.STAT_H1_SUM, _lib.STAT_EQ_SUM, _lib.STAT_NQ_SUM, _lib.STAT_COUNT):
        assert abs(s_eval[slot] - s_train[slot]) <= tol * abs(s_train[slot]) + 1e-12, slot
    q, J = model.predict(data["m"], batch_size=B_FULL)
    q64, J64 = q.astype(np.float64), J.astype(np.float64)
    tq, tJ = data["q"].astype(np.float64), data["J"].astype(np.float64)
    l2 = (((q64 - tq) ** 2).sum(1) / (tq ** 2).sum(1)).sum()
    h1 = (((J64 - tJ) ** 2).sum((1, 2)) / (tJ ** 2).sum((1, 2))).sum()
    # the predicted Jacobian is not rounded on its way out, the statistics are taken before the rounding too
    assert abs(l2 - s_eval[_lib.STAT_L2_SUM]) < 1e-4 * l2
    assert abs(h1 - s_eval[_lib.STAT_H1_SUM]) < 1e-4 * h1


def test_tensor_core_engine_agrees_with_the_fp32_engine_at_full_size(cuda_device):
    """tf32 chain vs the exact-fp32 engine on the full cfg2 minibatch: Jacobian 5e-3, gradient 1e-2 (the bars of
    tests/test_gpu_tc.py against the fp64 oracle, here with the fp32 engine - itself at 1e-4 - as the checker)."""
    arch, data, m32, ds32 = _setup("fp32")
    _, _, mtf, dstf = _setup("tf32")
    idx = np.arange(B_FULL)
    g32, s32 = _step(m32, ds32, idx, 1.0 / B_FULL)
    gtf, stf = _step(mtf, dstf, idx, 1.0 / B_FULL)
    assert H.rel_err(gtf, g32) < 1e-2
    assert abs(stf[_lib.STAT_H1_SUM] - s32[_lib.STAT_H1_SUM]) < 5e-3 * s32[_lib.STAT_H1_SUM]
    assert abs(stf[_lib.STAT_L2_SUM] - s32[_lib.STAT_L2_SUM]) < 1e-5 * s32[_lib.STAT_L2_SUM]      # the network itself stays fp32
    q32, J32 = m32.predict(data["m"], batch_size=B_FULL)
    qtf, Jtf = mtf.predict(data["m"], batch_size=B_FULL)
    assert H.rel_err(qtf, q32) < 1e-5
    assert H.rel_err(Jtf, J32) < 5e-3


def test_fp32x3_agrees_with_the_fp32_engine_at_the_bench_minibatch(cuda_device):
    """The bench workload (cfg5: width 1024, per-GPU minibatch 256 -> 25600 chain rows, the CTA-pair kernels, the
    2048-row accumulator cap of the weight gradients): fp32x3 against the SIMT fp32 engine - itself within 1e-4 of the
    oracle at this shape (tests/test_gpu_shapes.py) - at the full minibatch the oracle cannot do in seconds: 1e-4."""
    arch = S.CONFIGS["cfg5"]["arch"]
    B = 256
    ds_dev, tw, m_host = S.make_device_dataset(arch, B, 77)
    w = S.student_weights(arch, 78, tw, bias_noise=0.1)
    idx = np.arange(B)
    out = {}
    for precision in ("fp32", "fp32x3"):
        model = S.build_model(arch, w, precision=precision)
        out[precision] = _step(model, ds_dev, idx, 1.0 / B, B_plan=B)
        model.core.invalidate_plans()
    (g32, s32), (gx3, sx3) = out["fp32"], out["fp32x3"]
    assert np.isfinite(gx3).all() and np.abs(g32).max() > 0
    assert H.rel_err(gx3, g32) < 1e-4
    for slot in (_lib.STAT_L2_SUM, _lib.STAT_H1_SUM):
        assert abs(sx3[slot] - s32[slot]) < 1e-4 * abs(s32[slot])
