"""-m gpu parity tests at the shapes of BASELINE.json's configurations, CUDA path vs the fp64 oracle (never vs another
engine of this library), in every precision mode.

Per configuration and mode: predicted q and Jacobian, the loss, and every trainable gradient of one minibatch (one SGD
step with lr = 1 moves each parameter by minus its gradient).  Bars (relative to the largest entry of the reference):

  fp32x3 (tcgen05, three MMAs on hi/lo splits; the default)   1e-4   <- north_star's fp32 bar
  fp32   (SIMT FFMA)                                          1e-4
  tf32   (tcgen05, one MMA; reported variant)                 Jacobian 5e-3, gradients 1e-2

cfg2 runs at BASELINE's full minibatch (256); the wide configurations run at the minibatch the oracle's double backward
finishes in seconds (the kernels' tiling does not depend on the minibatch beyond the number of 128-row tiles)."""
import numpy as np
import pytest
import torch

import helpers as H
from oracle import dino_oracle as O
from dino_b200 import synthetic as S
from dino_b200 import Adam, SGD
from dino_b200.construction import (equip_model_with_output_reduced_jacobian, normalized_mse, normalized_mse_matrix,
                                    l2_accuracy, f_accuracy_matrix)

pytestmark = pytest.mark.gpu

BARS = {"fp32x3": (1e-4, 1e-4), "fp32": (1e-4, 1e-4), "tf32": (5e-3, 1e-2)}     # (Jacobian / loss, gradients)
CASES = [("cfg1", 32), ("cfg2", 256), ("cfg3", 8), ("cfg3b", 8), ("cfg5", 6), ("cfg4", 3)]


def _compile(model, opt=None):
    model.compile(optimizer=opt or Adam(1e-3), loss=[normalized_mse, normalized_mse_matrix], loss_weights=[1.0, 1.0],
                  metrics=[l2_accuracy, f_accuracy_matrix])


def _check_step(arch, B, precision, seed):
    tol_j, tol_g = BARS[precision]
    data, tw = S.make_dataset(arch, B, seed)
    w = S.student_weights(arch, seed + 1, tw, bias_noise=0.1)
    net = H.oracle_net(arch, w)
    qo, Jo = O.batch_jacobian(net, torch.tensor(data["m"], dtype=torch.float64), False)
    (lt, l2, h1), grads = O.h1_step_grads(net, H.oracle_batch(data), (1.0, 1.0), vectorised=True)
    model = S.build_model(arch, w, precision=precision)
    q, J = model.predict(data["m"], batch_size=B)
    assert np.isfinite(J).all()
    assert H.rel_err(q, qo.detach().numpy()) < (1e-5 if precision != "tf32" else 1e-3)
    assert H.rel_err(J, Jo.numpy()) < tol_j
    _compile(model, SGD(1.0))
    before = {l.name: l.get_weights() for l in model.layers}
    hist = model.fit([data["m"]], [data["q"], data["J"]], epochs=1, batch_size=B, perms=[np.arange(B)])
    assert abs(hist.history["loss"][0] - lt) < tol_j * abs(lt)
    gi, worst = 0, 0.0
    for l, lname in enumerate(model.core.names):
        if not model.core.trainable[l]:
            continue
        for a, b in zip(before[lname], model.get_layer(lname).get_weights()):
            err = H.rel_err(a - b, grads[gi].numpy())
            worst = max(worst, err)
            assert err < tol_g, (lname, err)
            gi += 1
    return worst


@pytest.mark.parametrize("precision", ["fp32x3", "tf32", "fp32"])
@pytest.mark.parametrize("cfg,B", CASES)
def test_baseline_shape_step_vs_oracle(cuda_device, cfg, B, precision):
    arch = S.CONFIGS[cfg]["arch"]
    _check_step(arch, B, precision, seed=31 + len(cfg) + B)


@pytest.mark.parametrize("precision", ["fp32x3", "tf32"])
@pytest.mark.parametrize("shape", ["narrow", "wide"])
def test_output_reduced_seed_on_the_tensor_cores(cuda_device, precision, shape):
    """Phi^T-seeded chain (equip_model_with_output_reduced_jacobian) through the fused kernel (narrow) and the
    layer-by-layer tensor-core path (wide) vs the oracle: Jacobian and parameters after one epoch."""
    tol_j, tol_g = BARS[precision]
    arch = (S.Arch("rb_dense", 24, 20, hidden=[16, 12], frozen_output=True, r_out=12) if shape == "narrow"
            else S.Arch("rb_dense", 40, 24, hidden=[150, 136, 12], frozen_output=True, r_out=12))
    n = 21
    data, tw = S.make_dataset(arch, n, 2)
    w = S.student_weights(arch, 3, tw, bias_noise=0.1)
    rng = np.random.default_rng(5)
    Phi = np.linalg.qr(rng.normal(size=(arch.d_out, 6)))[0]
    PhiTJ = np.einsum("ij,kil->kjl", Phi, data["J"]).astype(np.float32)
    base = S.build_model(arch, w, jac=None, precision=precision)
    model = equip_model_with_output_reduced_jacobian(base, Phi)
    assert model.precision == precision
    q, J = model.predict(data["m"])
    net = H.oracle_net(arch, w)
    qo, Jo = O.output_reduced_jacobian(net, torch.tensor(data["m"], dtype=torch.float64), torch.tensor(Phi), False)
    assert H.rel_err(J, Jo.numpy()) < tol_j
    _compile(model, SGD(1e-2))
    perms = [np.arange(n)]
    model.fit([data["m"]], [data["q"], PhiTJ], epochs=1, batch_size=8, perms=perms)
    d2 = dict(m=data["m"], q=data["q"], J=PhiTJ)
    O.fit(net, H.oracle_batch(d2), epochs=1, batch_size=8, lr=1e-2, opt="sgd", jac_mode="out_reduced", Phi=torch.tensor(Phi),
          perms=perms)
    ow = net.get_weights()
    for layer in model.layers:
        w0 = w.get(layer.name)
        for i, (a, b) in enumerate(zip(layer.get_weights(), ow.get(layer.name, []))):
            if np.abs(b - w0[i]).max() > 0:
                assert H.rel_err(a - w0[i], b - w0[i]) < tol_g, layer.name      # the UPDATE, relative to its own size


# (measured, profiles/r02a_precision.md: fp32 5e-6, fp32x3 1.4e-4, tf32 9e-3 -- Adam divides by sqrt(v), which turns the
#  per-step gradient differences of 4e-7 / 6e-6 / 8e-4 into larger relative differences of the smallest updates)
@pytest.mark.parametrize("precision,tol", [("fp32x3", 5e-4), ("fp32", 1e-4), ("tf32", 2e-2)])
def test_parameters_after_100_adam_steps(cuda_device, precision, tol):
    """cfg1's network (r_in = r_out = 50, depth 6): 100 Adam steps (4 epochs of 25 minibatches of 16, fixed
    permutations, validation data supplied as train_h1_network does) end at the oracle's parameters."""
    arch = S.CONFIGS["cfg1"]["arch"]
    n = 400
    data, tw = S.make_dataset(arch, n, 17)
    w = S.student_weights(arch, 18, tw)
    perms = [np.random.default_rng(e).permutation(n) for e in range(4)]
    model = S.build_model(arch, w, precision=precision)
    _compile(model, Adam(1e-3))
    val = ([data["m"][:64]], [data["q"][:64], data["J"][:64]])
    hist = model.fit([data["m"]], [data["q"], data["J"]], validation_data=val, epochs=4, batch_size=16, perms=perms)
    net = H.oracle_net(arch, w)
    ohist, _ = O.fit(net, H.oracle_batch(data), epochs=4, batch_size=16, lr=1e-3, opt="adam", perms=perms, vectorised=True)
    ow = net.get_weights()
    for layer in model.layers:
        w0 = w.get(layer.name)
        for i, (a, b) in enumerate(zip(layer.get_weights(), ow.get(layer.name, []))):
            assert H.rel_err(a - w0[i], b - w0[i]) < tol, layer.name            # what 100 steps did to the parameters
    last = np.mean([h[0] for h in ohist[-25:]])
    assert abs(hist.history["loss"][-1] - last) < max(tol, 1e-4) * abs(last)
    assert "val_loss" in hist.history


def test_engine_survives_a_larger_plan_requested_by_evaluate(cuda_device):
    """train_h1_network always passes validation_data, and under data parallelism the validation batch is larger than
    the engine's shard: an evaluate that needs a larger plan must not pull the plan, the workspace or the captured
    graphs from under a live TrainEngine."""
    arch = H.SMALL["dense_frozen"]
    data, tw = S.make_dataset(arch, 96, 5)
    w = S.student_weights(arch, 6, tw, bias_noise=0.1)
    runs = []
    for grow in (False, True):
        model = S.build_model(arch, w)
        _compile(model, Adam(1e-3))
        ds = model.make_dataset([data["m"][:32]], [data["q"][:32], data["J"][:32]])
        eng = model.make_engine(ds, 8)
        eng.set_order(np.arange(32))
        eng.step(); eng.step()                      # the second step of a size captures its CUDA graph
        if grow:
            ev = model.evaluate([data["m"]], [data["q"], data["J"]], batch_size=64)      # plan for 64 > 8
            assert np.isfinite(ev).all()
        eng.step(); eng.step()
        torch.cuda.synchronize()
        runs.append([lw for l in model.layers for lw in l.get_weights()])
    for a_, b_ in zip(*runs):            # (not bit-equal: the split-K reductions add in a run-dependent order)
        assert np.allclose(a_, b_, rtol=1e-4, atol=1e-6)


def test_fit_with_validation_data(cuda_device):
    arch = H.SMALL["dense_frozen"]
    data, tw = S.make_dataset(arch, 96, 5)
    w = S.student_weights(arch, 6, tw, bias_noise=0.1)
    model = S.build_model(arch, w)
    _compile(model, Adam(1e-3))
    val = ([data["m"][32:]], [data["q"][32:], data["J"][32:]])
    hist = model.fit([data["m"][:32]], [data["q"][:32], data["J"][:32]], validation_data=val, epochs=3, batch_size=8)
    assert len(hist.history["val_loss"]) == 3 and np.isfinite(hist.history["val_loss"]).all()


def test_two_plans_on_two_devices_in_one_process(cuda_device):
    """Kernel attributes (dynamic shared memory) are per DEVICE and launch descriptors are not process-global: one
    process drives a plan on cuda:0 and another on cuda:1, alternating, in both tensor-core modes."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs in one box")
    results = {}
    for precision in ("fp32x3", "tf32"):
        for cfg, B in (("cfg2", 64), ("cfg3", 4)):
            arch = S.CONFIGS[cfg]["arch"]
            data, tw = S.make_dataset(arch, B, 3)
            w = S.student_weights(arch, 4, tw, bias_noise=0.1)
            for dev in (1, 0, 1):
                torch.cuda.set_device(dev)
                model = S.build_model(arch, w, precision=precision)
                _compile(model, SGD(1.0))
                hist = model.fit([data["m"]], [data["q"], data["J"]], epochs=1, batch_size=B, perms=[np.arange(B)])
                results.setdefault((precision, cfg), []).append(hist.history["loss"][0])
    torch.cuda.set_device(0)
    for key, losses in results.items():
        assert np.isfinite(losses).all() and max(losses) - min(losses) < 1e-5 * abs(losses[0]), (key, losses)


@pytest.mark.parametrize("precision", ["fp32x3", "tf32"])
@pytest.mark.parametrize("d_out,hidden", [(72, [40, 36]), (96, [40, 36]), (48, [264, 40])])
def test_first_layer_without_its_jacobian(cuda_device, precision, d_out, hidden, monkeypatch):
    """Full Jacobian over an unreduced input (d_in >= 1024): the step that never forms J_hat (csrc/plan.cu,
    run_implicit_layer0: the target gathered per sample by TMA, J W and J^T Q instead of G W and G^T Q) against the
    oracle and against the explicit path of the same library, over shuffled minibatches with a ragged last one: loss,
    every metric of the history (the per-row Jacobian accuracy comes from row dot products there) and the weights."""
    tol_j, tol_g = BARS[precision]
    # 72 rows per sample: row tiles stop at sample ends (one part-empty tile each); 96: 128-row tiles of consecutive rows,
    # loaded as 32-row-granular boxes from up to three samples; 1030 % 4 != 0: the target is re-pitched on upload
    # hidden[0] = 264: a first layer wider than two 128-column tiles (the Gram matrices take the wide weight-gradient tiles)
    arch = S.Arch("generic_dense", 1030, d_out, hidden=hidden)
    n, B = 13, 5
    data, tw = S.make_dataset(arch, n, 77)
    w = S.student_weights(arch, 78, tw, bias_noise=0.1)
    rng = np.random.default_rng(3)
    perms = [rng.permutation(n) for _ in range(2)]

    def run(explicit):
        if explicit:
            monkeypatch.setenv("DINO_B200_NO_IMPLICIT0", "1")
        else:
            monkeypatch.delenv("DINO_B200_NO_IMPLICIT0", raising=False)
        model = S.build_model(arch, w, precision=precision)
        _compile(model, SGD(0.5))
        hist = model.fit([data["m"]], [data["q"], data["J"]], epochs=2, batch_size=B, perms=perms,
                         validation_data=([data["m"]], [data["q"], data["J"]]))
        return hist.history, {l.name: l.get_weights() for l in model.layers}

    h_imp, w_imp = run(False)
    h_exp, w_exp = run(True)
    net = H.oracle_net(arch, w)
    O.fit(net, H.oracle_batch(data), epochs=2, batch_size=B, lr=0.5, opt="sgd", perms=perms, vectorised=True)
    ow = net.get_weights()
    for k in h_exp:
        if k.startswith("val_"):
            continue       # evaluation always forms J_hat
        assert np.allclose(h_imp[k], h_exp[k], rtol=20 * tol_j, atol=2 * tol_j), (k, h_imp[k], h_exp[k])
    for name, ws0 in w.items():
        for i, c in enumerate(ws0):
            assert H.rel_err(w_imp[name][i] - c, w_exp[name][i] - c) < 2 * tol_g, name
            assert H.rel_err(w_imp[name][i] - c, ow[name][i] - c) < 2 * tol_g, name


def test_first_layer_without_its_jacobian_over_100_adam_steps(cuda_device, monkeypatch):
    """100 Adam steps (25 epochs of 4 shuffled minibatches) of a full-Jacobian network over a 1030-wide input: the
    Jacobian-free first layer and the explicit path end at the same loss and the same parameters (fp32x3; Adam's
    division by sqrt(v) amplifies per-step differences of 1e-6, hence 2e-3 on the parameter change)."""
    arch = S.Arch("generic_dense", 1030, 96, hidden=[40, 36])
    n, B = 64, 16
    data, tw = S.make_dataset(arch, n, 5)
    w = S.student_weights(arch, 6, tw, bias_noise=0.1)
    rng = np.random.default_rng(9)
    perms = [rng.permutation(n) for _ in range(25)]
    res = []
    for explicit in (False, True):
        if explicit:
            monkeypatch.setenv("DINO_B200_NO_IMPLICIT0", "1")
        else:
            monkeypatch.delenv("DINO_B200_NO_IMPLICIT0", raising=False)
        model = S.build_model(arch, w, precision="fp32x3")
        _compile(model, Adam(1e-3))
        hist = model.fit([data["m"]], [data["q"], data["J"]], epochs=25, batch_size=B, perms=perms)
        res.append((hist.history["loss"], {l.name: l.get_weights() for l in model.layers}))
    (la, wa), (lb, wb) = res
    assert lb[-1] < 0.9 * lb[0]                                   # it trains
    assert np.allclose(la, lb, rtol=1e-3, atol=0)
    for name, ws0 in w.items():
        for i, c in enumerate(ws0):
            assert H.rel_err(wa[name][i] - c, wb[name][i] - c) < 2e-3, name
