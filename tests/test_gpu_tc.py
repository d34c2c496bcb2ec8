"""-m gpu tests of the tcgen05 TF32 engine (csrc/gemm_tc.cu) and of the tf32 precision mode of the step.

TF32 keeps 10 mantissa bits of each operand (products are exact, accumulation is fp32), so one
contraction carries a relative error of ~2^-11 ~ 5e-4 per operand in the worst case; the asserts use
2e-3 of the result scale for a single GEMM and 1e-2 for gradients that went through a chain of ~20
TF32 contractions.  The fp32 engine is the 1e-4 parity mode (tests/test_gpu_parity.py)."""
import ctypes as C

import numpy as np
import pytest
import torch

import helpers as H
from oracle import dino_oracle as O
from dino_b200 import synthetic as S
from dino_b200 import _lib, Adam
from dino_b200.model import _stream
from dino_b200.construction import normalized_mse, normalized_mse_matrix, l2_accuracy, f_accuracy_matrix

pytestmark = pytest.mark.gpu


def _gemm(engine, A, Bt, M, N, K, lda, ldb, ldc):
    lib = _lib.load()
    Cm = torch.full((M, ldc), float("nan"), device=A.device)
    _lib.check(lib.dino_gemm_nt(engine, C.c_void_p(A.data_ptr()), lda, C.c_void_p(Bt.data_ptr()), ldb,
                                C.c_void_p(Cm.data_ptr()), ldc, M, N, K, _stream()), "dino_gemm_nt")
    torch.cuda.synchronize()
    return Cm[:, :N].cpu().numpy()


@pytest.mark.parametrize("M,N,K", [(128, 64, 32), (100, 50, 50), (1600, 50, 100), (300, 100, 50), (257, 130, 70),
                                   (512, 400, 200), (1000, 1024, 1024), (128, 8, 8), (5, 52, 300), (8192, 520, 100), (19000, 1000, 260)])
def test_tf32_gemm_matches_fp64(cuda_device, M, N, K):
    rng = np.random.default_rng(M * 7 + N * 3 + K)
    r4 = lambda n: (n + 3) & ~3
    lda, ldb, ldc = r4(K), r4(K), r4(N)
    A = np.zeros((M, lda), np.float32); A[:, :K] = rng.normal(size=(M, K))
    Bt = np.zeros((N, ldb), np.float32); Bt[:, :K] = rng.normal(size=(N, K))
    # poison the padding: the engine must never read columns >= K as data
    A[:, K:] = 1e30; Bt[:, K:] = 1e30
    dA, dB = torch.from_numpy(A).cuda(), torch.from_numpy(Bt).cuda()
    want = A[:, :K].astype(np.float64) @ Bt[:, :K].astype(np.float64).T
    got_simt = _gemm(_lib.PREC_FP32, dA, dB, M, N, K, lda, ldb, ldc)
    assert np.abs(got_simt - want).max() < 1e-5 * np.sqrt(K) * 4
    got = _gemm(_lib.PREC_TF32, dA, dB, M, N, K, lda, ldb, ldc)
    assert np.isfinite(got).all()
    assert np.abs(got - want).max() < 2e-3 * np.sqrt(K) * 4, np.abs(got - want).max()
    # fp32x3: three MMAs on hi/lo splits of both operands -> fp32-grade products
    got3 = _gemm(_lib.PREC_FP32X3, dA, dB, M, N, K, lda, ldb, ldc)
    assert np.isfinite(got3).all()
    assert np.abs(got3 - want).max() < 1e-5 * np.sqrt(K) * 4, np.abs(got3 - want).max()


@pytest.mark.parametrize("rows,n_in,n_out", [(64, 32, 32), (240, 13, 8), (1000, 50, 50), (12800, 100, 50), (300, 128, 128),
                                             (77, 5, 100), (520, 300, 200), (1000, 130, 257)])
def test_tf32_weight_gradient_contraction(cuda_device, rows, n_in, n_out):
    """C += XT YT^T (split over the reduction, red.global.add epilogue) vs fp64, onto a non-zero C."""
    lib = _lib.load()
    rng = np.random.default_rng(rows + n_in + n_out)
    ld = (rows + 3) & ~3
    XT = np.full((n_in, ld), 1e30, np.float32); XT[:, :rows] = rng.normal(size=(n_in, rows))
    YT = np.full((n_out, ld), 1e30, np.float32); YT[:, :rows] = rng.normal(size=(n_out, rows))
    C0 = rng.normal(size=(n_in, n_out)).astype(np.float32)
    want = C0 + XT[:, :rows].astype(np.float64) @ YT[:, :rows].astype(np.float64).T
    for engine, tol in ((_lib.PREC_FP32, 1e-5), (_lib.PREC_TF32, 2e-3), (_lib.PREC_FP32X3, 1e-5)):
        dX, dY, dC = torch.from_numpy(XT).cuda(), torch.from_numpy(YT).cuda(), torch.from_numpy(C0.copy()).cuda()
        _lib.check(lib.dino_gemm_nt_acc(engine, C.c_void_p(dX.data_ptr()), ld, C.c_void_p(dY.data_ptr()), ld,
                                        C.c_void_p(dC.data_ptr()), rows, n_in, n_out, _stream()), "dino_gemm_nt_acc")
        torch.cuda.synchronize()
        got = dC.cpu().numpy()
        assert np.isfinite(got).all()
        assert np.abs(got - want).max() < tol * np.sqrt(rows) * 4, (engine, np.abs(got - want).max())


@pytest.mark.parametrize("rows,n_in,n_out", [(64, 32, 32), (240, 13, 8), (1000, 50, 50), (12800, 100, 50), (300, 128, 128),
                                             (77, 5, 100), (520, 300, 200), (1000, 130, 257), (4096, 300, 260),
                                             (6000, 1024, 520), (2500, 129, 400)])
def test_tf32_weight_gradient_contraction_row_major(cuda_device, rows, n_in, n_out):
    """C += X^T Y with the factors row-major (MN-major tensor-core operands) vs fp64, onto a non-zero C."""
    lib = _lib.load()
    rng = np.random.default_rng(rows + n_in + n_out + 1)
    ldx, ldy = (n_in + 3) & ~3, (n_out + 7) & ~3
    X = np.full((rows, ldx), 1e30, np.float32); X[:, :n_in] = rng.normal(size=(rows, n_in))
    Y = np.full((rows, ldy), 1e30, np.float32); Y[:, :n_out] = rng.normal(size=(rows, n_out))
    C0 = rng.normal(size=(n_in, n_out)).astype(np.float32)
    want = C0 + X[:, :n_in].astype(np.float64).T @ Y[:, :n_out].astype(np.float64)
    for engine, tol in ((_lib.PREC_FP32, 1e-5), (_lib.PREC_TF32, 2e-3), (_lib.PREC_FP32X3, 1e-5)):
        dX, dY, dC = torch.from_numpy(X).cuda(), torch.from_numpy(Y).cuda(), torch.from_numpy(C0.copy()).cuda()
        _lib.check(lib.dino_gemm_tn_acc(engine, C.c_void_p(dX.data_ptr()), ldx, C.c_void_p(dY.data_ptr()), ldy,
                                        C.c_void_p(dC.data_ptr()), rows, n_in, n_out, _stream()), "dino_gemm_tn_acc")
        torch.cuda.synchronize()
        got = dC.cpu().numpy()
        assert np.isfinite(got).all()
        assert np.abs(got - want).max() < tol * np.sqrt(rows) * 4, (engine, np.abs(got - want).max())


def _compile(model):
    model.compile(optimizer=Adam(1e-3), loss=[normalized_mse, normalized_mse_matrix], loss_weights=[1.0, 1.0],
                  metrics=[l2_accuracy, f_accuracy_matrix])


@pytest.mark.parametrize("name", ["dense_frozen", "resnet", "generic"])
def test_tf32_mode_jacobian_and_gradients(cuda_device, name):
    arch = H.SMALL[name]
    data, tw = S.make_dataset(arch, 24, 3)
    w = H.student_weights(arch, 4, tw, bias_noise=0.1)
    model = S.build_model(arch, w, precision="tf32")
    q, J = model.predict(data["m"], batch_size=24)
    net = H.oracle_net(arch, w)
    qo, Jo = O.batch_jacobian(net, torch.tensor(data["m"], dtype=torch.float64), False)
    assert H.rel_err(q, qo.detach().numpy()) < 1e-5            # the MLP itself stays fp32
    assert H.rel_err(J, Jo.numpy()) < 5e-3
    # one SGD step with lr 1: delta theta = -gradient
    from dino_b200 import SGD
    (lt, l2, h1), grads = O.h1_step_grads(net, H.oracle_batch(data), (1.0, 1.0))
    _compile(model)
    model.optimizer = SGD(1.0)
    before = {l.name: l.get_weights() for l in model.layers}
    hist = model.fit([data["m"]], [data["q"], data["J"]], epochs=1, batch_size=24, perms=[np.arange(24)])
    assert abs(hist.history["loss"][0] - lt) < 5e-3 * abs(lt)
    gi = 0
    for l, lname in enumerate(model.core.names):
        if not model.core.trainable[l]:
            continue
        for a, b in zip(before[lname], model.get_layer(lname).get_weights()):
            assert H.rel_err(a - b, grads[gi].numpy()) < 1e-2, lname
            gi += 1


def test_tf32_mode_cfg2_widths(cuda_device):
    arch = S.CONFIGS["cfg2"]["arch"]
    data, tw = S.make_dataset(arch, 64, 21)
    w = H.student_weights(arch, 22, tw, bias_noise=0.1)
    m32 = S.build_model(arch, w, precision="fp32")
    mtf = S.build_model(arch, w, precision="tf32")
    q0, J0 = m32.predict(data["m"], batch_size=64)
    q1, J1 = mtf.predict(data["m"], batch_size=64)
    assert H.rel_err(J1, J0) < 5e-3
    _compile(m32); _compile(mtf)
    e0 = m32.evaluate([data["m"]], [data["q"], data["J"]], batch_size=64)
    e1 = mtf.evaluate([data["m"]], [data["q"], data["J"]], batch_size=64)
    assert np.allclose(e0, e1, rtol=5e-3, atol=1e-5)


def test_tf32_mode_wide_network_layerwise_path(cuda_device):
    """Widths > 128 take the layer-by-layer tensor-core path (gemm_tc + transposes + tiled wgrad_tc)."""
    arch = S.Arch("rb_dense", 40, 20, hidden=[150, 200, 20], frozen_output=True, r_out=20)
    data, tw = S.make_dataset(arch, 16, 9)
    w = H.student_weights(arch, 10, tw, bias_noise=0.1)
    from dino_b200 import SGD
    net = H.oracle_net(arch, w)
    (lt, l2, h1), grads = O.h1_step_grads(net, H.oracle_batch(data), (1.0, 1.0), vectorised=True)
    model = S.build_model(arch, w, precision="tf32")
    _compile(model)
    model.optimizer = SGD(1.0)
    before = {l.name: l.get_weights() for l in model.layers}
    hist = model.fit([data["m"]], [data["q"], data["J"]], epochs=1, batch_size=16, perms=[np.arange(16)])
    assert abs(hist.history["loss"][0] - lt) < 5e-3 * abs(lt)
    gi = 0
    for l, lname in enumerate(model.core.names):
        if not model.core.trainable[l]:
            continue
        for a, b in zip(before[lname], model.get_layer(lname).get_weights()):
            assert H.rel_err(a - b, grads[gi].numpy()) < 1e-2, lname
            gi += 1
