"""The hand-derived reverse-chain / double-backward stages (tests/pipeline_model.py, the numpy model
of what csrc/plan.cu enqueues) against the autograd oracle, fp64, every Jacobian mode and layer kind."""
import numpy as np
import pytest
import torch

import pipeline_model as P
from oracle import dino_oracle as O

rng = np.random.default_rng(0)


def to_layers(net):
    return [P.Layer(L.kernel.numpy(), None if L.bias is None else L.bias.numpy(), L.act, L.trainable, L.skip_from)
            for L in net.layers]


def nets():
    Phi0 = np.linalg.qr(rng.normal(size=(9, 6)))[0]
    last = [Phi0.T, rng.normal(size=9)]
    yield "dense", O.build_projected_dense(rng, reduced_input_dim=7, last_layer_weights=last,
                                           hidden_layer_dimensions=[6, 6, 6])
    yield "resnet", O.build_projected_resnet(rng, reduced_input_dim=7, last_layer_weights=last, ranks=[3, 3, 4],
                                             zeros=False)
    yield "generic", O.build_generic_dense(rng, 11, 5, [8, 8])


@pytest.mark.parametrize("mode", ["full", "out_reduced", "sketched"])
@pytest.mark.parametrize("kind", ["dense", "resnet", "generic"])
def test_pipeline_matches_autograd(kind, mode):
    net = dict(nets())[kind]
    for L in net.layers:
        if L.bias is not None and L.trainable:
            L.bias += torch.tensor(rng.normal(0, 0.1, L.bias.shape))
    B, din, dout = 5, net.d_in, net.d_out
    m, q = rng.normal(size=(B, din)), rng.normal(size=(B, dout))
    Phi = U = V = seed = None
    if mode == "full":
        J = rng.normal(size=(B, dout, din))
    elif mode == "out_reduced":
        Phi = np.linalg.qr(rng.normal(size=(dout, 3)))[0]
        J, seed = rng.normal(size=(B, 3, din)), Phi.T
    else:
        k = 4
        U, V = rng.normal(size=(B, dout, k)), rng.normal(size=(B, din, k))
        J, seed = rng.normal(size=(B, k, k)), U.transpose(0, 2, 1)
    t = lambda a: None if a is None else torch.tensor(a)
    (lt, l2, h1), grads = O.h1_step_grads(net, O.Batch(t(m), t(q), t(J), t(U), t(V)), (0.7, 1.3), mode, t(Phi))
    out = P.step(to_layers(net), m, q, J, (0.7, 1.3), seed=seed, V=V)
    assert abs(out["loss3"][0] - lt) < 1e-12 * abs(lt)
    gi = 0
    for l, L in enumerate(net.layers):
        if not L.trainable:
            continue
        ref = grads[gi].numpy(); gi += 1
        assert np.abs(ref - out["dW"][l]).max() < 1e-12 * (np.abs(ref).max() + 1e-30)
        if L.bias is not None:
            ref = grads[gi].numpy(); gi += 1
            assert np.abs(ref - out["db"][l]).max() < 1e-12 * (np.abs(ref).max() + 1e-30)


def test_first_layer_without_its_jacobian_is_the_same_step():
    """The identities behind csrc/plan.cu: run_implicit_layer0 (G W = c (Q M - T), G^T Q = W S - J^T (c Q), the per-row
    error from row dot products), fp64: the same loss and gradients as the explicit stages, and as autograd."""
    net = dict(nets())["generic"]
    B = 4
    m, q = rng.normal(size=(B, net.d_in)), rng.normal(size=(B, net.d_out))
    J = rng.normal(size=(B, net.d_out, net.d_in))
    layers = to_layers(net)
    a = P.step(layers, m, q, J, w=(0.7, 1.3))
    b = P.step(layers, m, q, J, w=(0.7, 1.3), implicit_first_layer=True)
    assert np.allclose(a["loss3"], b["loss3"], rtol=1e-12, atol=0)
    for x, y in zip(a["dW"] + [v for v in a["db"] if v is not None], b["dW"] + [v for v in b["db"] if v is not None]):
        assert np.abs(x - y).max() <= 1e-12 * max(np.abs(x).max(), 1e-30)
    batch = O.Batch(torch.tensor(m), torch.tensor(q), torch.tensor(J))
    (lt, _, _), grads = O.h1_step_grads(net, batch, (0.7, 1.3))
    assert abs(b["loss3"][0] - lt) < 1e-12 * abs(lt)
    mine = [g for l, L in enumerate(layers) for g in ([b["dW"][l]] + ([b["db"][l]] if L.b is not None else [])) if L.trainable]
    for x, y in zip(mine, grads):
        assert np.abs(x - y.numpy()).max() <= 1e-10 * max(np.abs(x).max(), 1e-30)
