"""-m gpu parity tests: the CUDA path (through the C ABI, via DinoModel) against the CPU oracle on
the same seeded inputs, in both fp32-grade modes (tcgen05 fp32x3 and the SIMT FFMA engine).  The oracle
computes in fp64; north_star asks for 1e-4 relative in fp32, the asserts below are tighter where fp32
round-off allows."""
import numpy as np
import pytest
import torch

import helpers as H
from oracle import dino_oracle as O
from dino_b200 import synthetic as S
from dino_b200 import _lib, Adam, SGD
from dino_b200.construction import (equip_model_with_full_jacobian, equip_model_with_output_reduced_jacobian,
                                    equip_model_with_sketched_jacobian, normalized_mse, normalized_mse_matrix,
                                    l2_accuracy, f_accuracy_matrix)

pytestmark = pytest.mark.gpu
TOL = 1e-4


@pytest.fixture(params=["fp32x3", "fp32"], autouse=True)
def precision(request, monkeypatch):
    """Both 1e-4 modes: the tensor-core fp32x3 arithmetic (the default of DinoModel) and the SIMT fp32 engine."""
    monkeypatch.setenv("DINO_B200_PRECISION", request.param)
    return request.param



def _problem(arch, n, seed):
    data, tw = S.make_dataset(arch, n, seed)
    w = H.student_weights(arch, seed + 1, tw, bias_noise=0.1)
    return data, w


def _compile(model, weights=(1.0, 1.0), opt=None):
    model.compile(optimizer=opt or Adam(1e-3), loss=[normalized_mse, normalized_mse_matrix],
                  loss_weights=list(weights), metrics=[l2_accuracy, f_accuracy_matrix])


@pytest.mark.parametrize("name", list(H.SMALL))
def test_forward_and_jacobian(cuda_device, name):
    arch = H.SMALL[name]
    data, w = _problem(arch, 19, 3)
    model = S.build_model(arch, w)
    qh, Jh = model.predict(data["m"], batch_size=8)          # ragged last batch (19 = 8+8+3)
    net = H.oracle_net(arch, w)
    qo, Jo = O.batch_jacobian(net, torch.tensor(data["m"], dtype=torch.float64), False)
    assert H.rel_err(qh, qo.detach().numpy()) < 1e-5
    assert H.rel_err(Jh, Jo.numpy()) < 1e-5


@pytest.mark.parametrize("name", list(H.SMALL))
@pytest.mark.parametrize("B", [1, 5, 32])
def test_step_gradients(cuda_device, name, B):
    """loss and d loss / d theta of one minibatch vs autograd double-backward."""
    import ctypes as C
    arch = H.SMALL[name]
    data, w = _problem(arch, 40, 7)
    model = S.build_model(arch, w)
    net = H.oracle_net(arch, w)
    idx = np.random.default_rng(0).permutation(40)[:B]
    (lt, l2, h1), grads = O.h1_step_grads(net, H.oracle_batch(data, idx), (0.7, 1.3))
    from dino_b200.model import DeviceDataset, _stream
    ds = DeviceDataset(model.jac_mode, [data["m"]], [data["q"], data["J"]])
    plan = model._plan(B)
    params = model.core.device_params()
    grad = torch.zeros(model.core.n_params, device=params.device)
    stats = torch.zeros(_lib.STAT_SLOTS, device=params.device)
    didx = torch.tensor(idx, dtype=torch.int32, device=params.device)
    batch = ds.batch(didx.data_ptr())
    lib = _lib.load()
    _lib.check(lib.dino_h1_fwd_bwd(plan.handle, C.c_void_p(params.data_ptr()), C.byref(batch), B, 0.7, 1.3, 1.0 / B,
                                   C.c_void_p(grad.data_ptr()), C.c_void_p(stats.data_ptr()), _stream()), "step")
    st = stats.cpu().numpy()
    assert abs(st[_lib.STAT_L2_SUM] / B - l2) < TOL * abs(l2)
    assert abs(st[_lib.STAT_H1_SUM] / B - h1) < TOL * abs(h1)
    g = grad.cpu().numpy()
    core, gi = model.core, 0
    for l, name_l in enumerate(core.names):
        k, n = core.n_in[l], core.n_out[l]
        gW = g[core.koff[l]:core.koff[l] + k * n].reshape(k, n)
        if not core.trainable[l]:
            assert np.all(gW == 0)
            continue
        assert H.rel_err(gW, grads[gi].numpy()) < TOL, name_l
        gi += 1
        if core.use_bias[l]:
            assert H.rel_err(g[core.boff[l]:core.boff[l] + n], grads[gi].numpy()) < TOL, name_l + " bias"
            gi += 1


@pytest.mark.parametrize("name", ["dense_frozen", "resnet", "generic"])
@pytest.mark.parametrize("opt", ["adam", "sgd"])
def test_fit_parameters_after_n_steps(cuda_device, name, opt):
    """Parameters after 2 epochs (ragged last batch, fixed permutations) match the oracle's fit."""
    arch = H.SMALL[name]
    data, w = _problem(arch, 37, 11)
    perms = [np.random.default_rng(e).permutation(37) for e in range(2)]
    lr = 1e-3 if opt == "adam" else 1e-2
    model = S.build_model(arch, w)
    _compile(model, (1.0, 1.0), Adam(lr) if opt == "adam" else SGD(lr))
    hist = model.fit([data["m"]], [data["q"], data["J"]], epochs=2, batch_size=8, perms=perms)
    net = H.oracle_net(arch, w)
    ohist, _ = O.fit(net, H.oracle_batch(data), epochs=2, batch_size=8, lr=lr, opt=opt, perms=perms)
    ow = net.get_weights()
    for layer in model.layers:
        for a, b in zip(layer.get_weights(), ow.get(layer.name, [])):
            assert H.rel_err(a, b) < TOL, layer.name
    # epoch loss = sample-weighted mean of the per-batch losses
    sizes = np.array([8, 8, 8, 8, 5] * 2, dtype=float)
    per_epoch = [np.average([h[0] for h in ohist[e * 5:(e + 1) * 5]], weights=sizes[:5]) for e in range(2)]
    assert np.allclose(hist.history["loss"], per_epoch, rtol=TOL)


@pytest.mark.parametrize("name", ["dense", "resnet"])
def test_evaluate_matches_keras_aggregation(cuda_device, name):
    arch = H.SMALL[name]
    data, w = _problem(arch, 45, 5)
    model = S.build_model(arch, w)
    _compile(model, (0.5, 2.0))
    vals = model.evaluate([data["m"]], [data["q"], data["J"]], batch_size=32)
    ref = O.evaluate(H.oracle_net(arch, w), H.oracle_batch(data), batch_size=32, loss_weights=(0.5, 2.0))
    assert np.allclose(vals, ref, rtol=TOL, atol=1e-6)


def test_output_reduced_jacobian(cuda_device):
    arch = H.SMALL["dense_frozen"]
    data, w = _problem(arch, 21, 2)
    rng = np.random.default_rng(5)
    Phi = np.linalg.qr(rng.normal(size=(arch.d_out, 4)))[0]
    PhiTJ = np.einsum("ij,kil->kjl", Phi, data["J"]).astype(np.float32)
    base = S.build_model(arch, w, jac=None)
    model = equip_model_with_output_reduced_jacobian(base, Phi)
    q, J = model.predict(data["m"])
    net = H.oracle_net(arch, w)
    qo, Jo = O.output_reduced_jacobian(net, torch.tensor(data["m"], dtype=torch.float64), torch.tensor(Phi), False)
    assert H.rel_err(J, Jo.numpy()) < 1e-5
    _compile(model)
    perms = [np.arange(21)]
    model.fit([data["m"]], [data["q"], PhiTJ], epochs=1, batch_size=8, perms=perms)
    d2 = dict(m=data["m"], q=data["q"], J=PhiTJ)
    O.fit(net, H.oracle_batch(d2), epochs=1, batch_size=8, jac_mode="out_reduced", Phi=torch.tensor(Phi), perms=perms)
    ow = net.get_weights()
    for layer in model.layers:
        for a, b in zip(layer.get_weights(), ow.get(layer.name, [])):
            assert H.rel_err(a, b) < TOL, layer.name


@pytest.mark.parametrize("name", ["generic", "dense_frozen"])
def test_sketched_jacobian(cuda_device, name):
    arch = H.SMALL[name]
    data, w = _problem(arch, 18, 4)
    rng = np.random.default_rng(9)
    k = 3
    U = rng.normal(size=(18, arch.d_out, k)).astype(np.float32)
    V = rng.normal(size=(18, arch.d_in, k)).astype(np.float32)
    sig = np.zeros((18, k, k), np.float32)
    sig[:, np.arange(k), np.arange(k)] = rng.uniform(0.5, 2.0, size=(18, k))
    base = S.build_model(arch, w, jac=None)
    model = equip_model_with_sketched_jacobian(base, k)
    q, Sg = model.predict([data["m"], U, V])
    net = H.oracle_net(arch, w)
    t = lambda a: torch.tensor(a, dtype=torch.float64)
    qo, So = O.sketched_jacobian(net, t(data["m"]), t(U), t(V), False)
    assert H.rel_err(Sg, So.numpy()) < 1e-5
    _compile(model)
    perms = [np.arange(18)]
    model.fit([data["m"], U, V], [data["q"], sig], epochs=1, batch_size=8, perms=perms)
    ob = O.Batch(t(data["m"]), t(data["q"]), t(sig), t(U), t(V))
    O.fit(net, ob, epochs=1, batch_size=8, jac_mode="sketched", perms=perms)
    ow = net.get_weights()
    for layer in model.layers:
        for a, b in zip(layer.get_weights(), ow.get(layer.name, [])):
            assert H.rel_err(a, b) < TOL, layer.name


def test_l2_only_training(cuda_device):
    arch = H.SMALL["generic"]
    data, w = _problem(arch, 33, 6)
    model = S.build_model(arch, w, jac=None)
    model.compile(optimizer=Adam(1e-3), loss=normalized_mse, metrics=[l2_accuracy])
    perms = [np.arange(33)]
    model.fit(data["m"], data["q"], epochs=1, batch_size=16, perms=perms)
    net = H.oracle_net(arch, w)
    opt = O.KerasAdam(net.trainable_tensors(), 1e-3)
    m, q = torch.tensor(data["m"], dtype=torch.float64), torch.tensor(data["q"], dtype=torch.float64)
    for lo in range(0, 33, 16):
        params = net.trainable_tensors()
        for p in params:
            p.requires_grad_(True)
        loss = O.normalized_mse(q[lo:lo + 16], O.forward(net, m[lo:lo + 16]))
        g = torch.autograd.grad(loss, params)
        for p in params:
            p.requires_grad_(False)
        opt.step(g)
    ow = net.get_weights()
    for layer in model.layers:
        for a, b in zip(layer.get_weights(), ow.get(layer.name, [])):
            assert H.rel_err(a, b) < TOL, layer.name
    ev = model.evaluate(data["m"], data["q"])
    assert len(ev) == 2


def test_zero_loss_when_targets_are_predictions(cuda_device):
    arch = H.SMALL["dense"]
    data, w = _problem(arch, 16, 8)
    model = S.build_model(arch, w)
    q, J = model.predict(data["m"])
    _compile(model)
    vals = model.evaluate([data["m"]], [q, J])
    assert abs(vals[0]) < 1e-10 and abs(vals[3] - 1) < 1e-5 and abs(vals[6] - 1) < 1e-5


def test_bench_config_shapes_gradients(cuda_device):
    """cfg2 (the bench workload) at B=64: gradient parity at the real widths (100/50)."""
    import ctypes as C
    arch = S.CONFIGS["cfg2"]["arch"]
    data, w = _problem(arch, 64, 21)
    model = S.build_model(arch, w)
    net = H.oracle_net(arch, w)
    (lt, l2, h1), grads = O.h1_step_grads(net, H.oracle_batch(data), (1.0, 1.0), vectorised=True)
    _compile(model)
    model.optimizer = SGD(1.0)           # one SGD step with lr 1: delta theta = -grad
    before = {l.name: l.get_weights() for l in model.layers}
    model.fit([data["m"]], [data["q"], data["J"]], epochs=1, batch_size=64, perms=[np.arange(64)])
    gi = 0
    for l, name in enumerate(model.core.names):
        if not model.core.trainable[l]:
            continue
        after = model.get_layer(name).get_weights()
        for a, b in zip(before[name], after):
            assert H.rel_err(a - b, grads[gi].numpy()) < 2e-4, name
            gi += 1


def test_adam_step_size_host_and_device_agree_with_double(cuda_device):
    """lr_t = lr sqrt(1 - b2^t) / (1 - b1^t) over 1 .. 10^5 steps: dino_adam (step number from the host) and
    dino_adam_dev (step number read on the device) against the oracle's double-precision value.  One parameter with
    g = 1, m = v = 0 moves by -lr_t (1 - b1) / (sqrt(1 - b2) + eps), which gives lr_t back."""
    import ctypes as C
    import math
    from dino_b200.model import _stream
    lib = _lib.load()
    lr, b1, b2, eps = 1e-3, 0.9, 0.999, 1e-7
    f32 = lambda x: float(np.float32(x))
    unit = (1.0 - f32(b1)) / (math.sqrt(1.0 - f32(b2)) + f32(eps))
    for t in (1, 2, 3, 10, 100, 1000, 6931, 10000, 31623, 100000):
        want = f32(lr) * math.sqrt(1.0 - f32(b2) ** t) / (1.0 - f32(b1) ** t)
        got = []
        for dev_counter in (False, True):
            p = torch.zeros(4, device="cuda"); g = torch.ones(4, device="cuda")
            m = torch.zeros(4, device="cuda"); v = torch.zeros(4, device="cuda")
            if dev_counter:
                done = torch.tensor([t - 1], dtype=torch.int64, device="cuda")
                _lib.check(lib.dino_adam_dev(C.c_void_p(p.data_ptr()), C.c_void_p(g.data_ptr()), C.c_void_p(m.data_ptr()),
                                             C.c_void_p(v.data_ptr()), 4, lr, b1, b2, eps, C.c_void_p(done.data_ptr()),
                                             _stream()), "dino_adam_dev")
            else:
                _lib.check(lib.dino_adam(C.c_void_p(p.data_ptr()), C.c_void_p(g.data_ptr()), C.c_void_p(m.data_ptr()),
                                         C.c_void_p(v.data_ptr()), 4, lr, b1, b2, eps, t, _stream()), "dino_adam")
            got.append(-float(p[0].item()) / unit)
        assert abs(got[0] - want) < 2e-6 * want, (t, got, want)
        assert abs(got[1] - want) < 2e-6 * want, (t, got, want)
        assert abs(got[0] - got[1]) < 1e-6 * want, (t, got)
