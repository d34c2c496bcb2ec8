"""-m gpu tests of the array methods of InferenceOracle (inferenceOracle.py:105-139) against the oracle: ``Jvp`` as one
forward-mode sweep (dino_jvp), ``Jtvp`` as a reverse chain with a single seed row, ``J`` as the full chain."""
import numpy as np
import pytest
import torch

import helpers as H
from oracle import dino_oracle as O
from dino_b200 import synthetic as S
from dino_b200.inference import InferenceOracle

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("precision", ["fp32x3", "fp32"])
@pytest.mark.parametrize("name", ["dense_frozen", "resnet", "resnet_square", "generic", "wide"])
def test_jvp_and_jtvp_without_the_jacobian(cuda_device, name, precision):
    arch = H.SMALL[name] if name != "wide" else S.Arch("rb_dense", 40, 24, hidden=[150, 136, 12], frozen_output=True, r_out=12)
    data, tw = S.make_dataset(arch, 9, 3)
    w = S.student_weights(arch, 4, tw, bias_noise=0.1)
    net = H.oracle_net(arch, w)
    _, Jo = O.batch_jacobian(net, torch.tensor(data["m"], dtype=torch.float64), False)
    Jo = Jo.numpy()
    model = S.build_model(arch, w, precision=precision)
    io = InferenceOracle(None, model)
    rng = np.random.default_rng(1)
    for i in (0, 5):
        mh, qh = rng.normal(size=arch.d_in), rng.normal(size=arch.d_out)
        assert H.rel_err(io.Jvp(data["m"][i], mh), Jo[i] @ mh) < 1e-5
        assert H.rel_err(io.Jtvp(data["m"][i], qh), Jo[i].T @ qh) < 1e-5
        assert H.rel_err(io.J(data["m"][i]), Jo[i]) < 1e-5
    # batched directions through the model method: one per sample, and one shared by all samples
    Mh = rng.normal(size=(9, arch.d_in)).astype(np.float32)
    assert H.rel_err(model.jvp(data["m"], Mh, batch_size=4), np.einsum("bij,bj->bi", Jo, Mh)) < 1e-5
    assert H.rel_err(model.jvp(data["m"], Mh[2], batch_size=4), np.einsum("bij,j->bi", Jo, Mh[2])) < 1e-5


@pytest.mark.parametrize("q,m,noise", [(10, 40, 0.3), (50, 100, 0.05), (7, 4225, 0.5), (20, 37, 1e-3)])
def test_jacobian_error_metrics_on_the_device(cuda_device, q, m, noise):
    """The seven metrics of compute_jacobian_errors (jacobianTests.py:104-126) from three d_Q x d_Q device Gram
    contractions per sample vs the oracle's restatement of the reference's lines (float64 numpy, SVD of the truth)."""
    from dino_b200.evaluation.jacobianTests import jacobian_error_metrics_device
    rng = np.random.default_rng(q + m)
    n = 5
    J = rng.normal(size=(n, q, m))
    P = J + noise * rng.normal(size=(n, q, m))
    got = jacobian_error_metrics_device(torch.tensor(P, dtype=torch.float32).cuda(), torch.tensor(J, dtype=torch.float32).cuda())
    for b in range(n):
        want = O.jacobian_errors(P[b].astype(np.float32).astype(np.float64), J[b].astype(np.float32).astype(np.float64))
        for k, v in want.items():
            assert abs(got[k][b] - v) < 2e-4 * max(v, 1e-3), (k, got[k][b], v)
