"""Known-answer tests that pin the oracle where the reference offers no vectors."""
import math

import numpy as np
import torch

from oracle import dino_oracle as O

rng = np.random.default_rng(1)
T = lambda a: torch.tensor(np.asarray(a), dtype=torch.float64)


def _dense(rng, **kw):
    net = O.build_projected_dense(rng, **kw)
    for L in net.layers:
        if L.bias is not None and L.trainable:
            L.bias += T(rng.normal(0, 0.2, L.bias.shape))
    return net


def test_linear_network_jacobian_is_the_weight_product():
    net = O.build_generic_dense(rng, 6, 4, [5, 5])
    for L in net.layers:
        L.act = O.ACT_LINEAR
    _, J = O.batch_jacobian(net, T(rng.normal(size=(3, 6))), False)
    W = net.layers[0].kernel @ net.layers[1].kernel @ net.layers[2].kernel
    assert torch.allclose(J, W.T.expand(3, 4, 6), atol=1e-14)


def test_jacobian_matches_central_finite_differences():
    net = _dense(rng, reduced_input_dim=5, reduced_output_dim=4, hidden_layer_dimensions=[6, 4])
    m = T(rng.normal(size=(2, 5)))
    _, J = O.batch_jacobian(net, m, False)
    h = 1e-6
    for j in range(5):
        e = torch.zeros(5, dtype=torch.float64); e[j] = h
        fd = (O.forward(net, m + e) - O.forward(net, m - e)) / (2 * h)
        assert torch.allclose(J[:, :, j], fd, atol=1e-7)


def test_vectorised_and_looped_reverse_passes_agree():
    net = _dense(rng, reduced_input_dim=5, reduced_output_dim=4, hidden_layer_dimensions=[6, 4])
    m = T(rng.normal(size=(3, 5)))
    _, J1 = O.batch_jacobian(net, m, False, vectorised=False)
    _, J2 = O.batch_jacobian(net, m, False, vectorised=True)
    assert torch.allclose(J1, J2, atol=1e-15)


def test_zero_loss_and_gradient_when_targets_equal_predictions():
    net = _dense(rng, reduced_input_dim=5, reduced_output_dim=4, hidden_layer_dimensions=[6, 4])
    m = T(rng.normal(size=(4, 5)))
    q, J = O.batch_jacobian(net, m, False)
    (lt, l2, h1), grads = O.h1_step_grads(net, O.Batch(m, q.detach(), J.detach()))
    assert lt == 0 and all(float(g.abs().max()) == 0 for g in grads)


def test_normalized_mse_matrix_of_zero_prediction_is_one():
    J = T(rng.normal(size=(3, 4, 5)))
    assert abs(float(O.normalized_mse_matrix(J, torch.zeros_like(J))) - 1.0) < 1e-15
    assert abs(float(O.f_accuracy_matrix(J, torch.zeros_like(J)))) < 1e-15


def test_reduced_versus_stitched_network_identities():
    """J_full V == J_r and ||J - Phi J_r V^T||^2 = ||J||^2 - ||Phi^T J V||^2 + ||Phi^T J V - J_r||^2."""
    dM, dQ, ri, ro = 9, 7, 4, 3
    V = np.linalg.qr(rng.normal(size=(dM, ri)))[0]
    Phi = np.linalg.qr(rng.normal(size=(dQ, ro)))[0]
    small = _dense(rng, reduced_input_dim=ri, reduced_output_dim=ro, hidden_layer_dimensions=[5, ro])
    big = O.build_projected_dense(rng, input_projector=V, last_layer_weights=[Phi.T, np.zeros(dQ)],
                                  hidden_layer_dimensions=[5, ro])
    by_name = {L.name: L for L in small.layers}
    for L in big.layers:
        if L.name in by_name:
            L.kernel, L.bias = by_name[L.name].kernel.clone(), None if by_name[L.name].bias is None else by_name[L.name].bias.clone()
    m = rng.normal(size=(2, dM))
    _, Jf = O.batch_jacobian(big, T(m), False)
    _, Jr = O.batch_jacobian(small, T(m @ V), False)
    assert torch.allclose(Jf @ T(V), T(Phi) @ Jr, atol=1e-13)
    Jt = rng.normal(size=(2, dQ, dM))
    lhs = ((Jt - (Phi @ Jr.numpy() @ V.T)) ** 2).sum((1, 2))
    red = np.einsum("qr,bqm,ms->brs", Phi, Jt, V)
    rhs = (Jt ** 2).sum((1, 2)) - (red ** 2).sum((1, 2)) + ((red - Jr.numpy()) ** 2).sum((1, 2))
    assert np.allclose(lhs, rhs, rtol=1e-12)


def test_sketched_output_equals_sketch_of_the_full_jacobian():
    net = _dense(rng, reduced_input_dim=6, reduced_output_dim=4, hidden_layer_dimensions=[5, 4])
    m, U, V = T(rng.normal(size=(3, 6))), T(rng.normal(size=(3, 4, 2))), T(rng.normal(size=(3, 6, 2)))
    _, S = O.sketched_jacobian(net, m, U, V, False)
    _, J = O.batch_jacobian(net, m, False)
    assert torch.allclose(S, U.transpose(1, 2) @ J @ V, atol=1e-13)


def test_keras_adam_three_hand_computed_steps():
    p = torch.tensor([1.0, -2.0], dtype=torch.float64)
    opt = O.KerasAdam([p], lr=0.1)
    m = np.zeros(2); v = np.zeros(2); x = np.array([1.0, -2.0])
    for t, g in enumerate([np.array([0.5, -1.0]), np.array([0.25, 2.0]), np.array([-1.0, 0.1])], start=1):
        opt.step([T(g)])
        m = 0.9 * m + 0.1 * g
        v = 0.999 * v + 0.001 * g * g
        x = x - 0.1 * math.sqrt(1 - 0.999 ** t) / (1 - 0.9 ** t) * m / (np.sqrt(v) + 1e-7)   # eps OUTSIDE
        assert np.allclose(p.numpy(), x, rtol=1e-14)
    # the torch.optim.Adam placement of eps gives a different (wrong for Keras) third step
    assert abs(p.numpy()[0] - 0.7) > 1e-3


def test_resnet_zero_initialised_blocks_are_identity_at_init():
    Phi = np.linalg.qr(rng.normal(size=(6, 4)))[0]
    net = O.build_projected_resnet(rng, reduced_input_dim=4, last_layer_weights=[Phi.T, np.zeros(6)], ranks=[3, 3])
    m = T(rng.normal(size=(2, 4)))
    _, J = O.batch_jacobian(net, m, False)
    compat = [L for L in net.layers if L.name.endswith("output_compat_layer")][0].kernel
    assert torch.allclose(J, (compat @ T(Phi.T)).T.expand(2, 6, 4), atol=1e-14)


def test_fit_short_last_batch_and_evaluate_weighting():
    net = _dense(rng, reduced_input_dim=5, reduced_output_dim=4, hidden_layer_dimensions=[6, 4])
    n = 11
    data = O.Batch(T(rng.normal(size=(n, 5))), T(rng.normal(size=(n, 4))), T(rng.normal(size=(n, 4, 5))))
    hist, opt = O.fit(net, data, epochs=1, batch_size=4)
    assert len(hist) == 3 and opt.t == 3
    ev = O.evaluate(net, data, batch_size=4)
    per = [O.evaluate_batch(net, O.slice_batch(data, np.arange(s, min(n, s + 4)))) for s in (0, 4, 8)]
    # losses: batch-size weights; function metrics (MeanMetricWrapper on per-batch scalars): every batch counts once
    per = np.array(per)
    want = list(((4 * per[0] + 4 * per[1] + 3 * per[2]) / 11)[:3]) + list(per.mean(0)[3:])
    assert np.allclose(ev, want, rtol=1e-13)
