"""Oracle and host mirror vs the golden vectors produced by the REFERENCE's own functions
(tests/golden/make_golden.py).  CPU only; reads nothing under /root/reference."""
import os

import numpy as np
import torch

from oracle import dino_oracle as O
import dino_b200.construction as DC

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_golden.npz"))
T = lambda a: torch.tensor(np.asarray(a), dtype=torch.float64)


def test_loss_and_metric_formulas_oracle():
    q, qh, J, Jh = (T(G[k]) for k in ("loss_q", "loss_qh", "loss_J", "loss_Jh"))
    assert np.isclose(float(O.normalized_mse(q, qh)), G["normalized_mse_q"], rtol=1e-13)
    assert np.isclose(float(O.l2_accuracy(q, qh)), G["l2_accuracy_q"], rtol=1e-13)
    assert np.isclose(float(O.f_accuracy_matrix(q, qh)), G["f_accuracy_matrix_q"], rtol=1e-13)
    assert np.isclose(float(O.normalized_mse_matrix(J, Jh)), G["normalized_mse_matrix_J"], rtol=1e-13)
    assert np.isclose(float(O.l2_accuracy(J, Jh)), G["l2_accuracy_J"], rtol=1e-13)
    assert np.isclose(float(O.f_accuracy_matrix(J, Jh)), G["f_accuracy_matrix_J"], rtol=1e-13)


def test_loss_and_metric_formulas_host_mirror():
    q, qh, J, Jh = (G[k] for k in ("loss_q", "loss_qh", "loss_J", "loss_Jh"))
    assert np.isclose(DC.normalized_mse(q, qh), G["normalized_mse_q"], rtol=1e-13)
    assert np.isclose(DC.l2_accuracy(q, qh), G["l2_accuracy_q"], rtol=1e-13)
    assert np.isclose(DC.f_accuracy_matrix(q, qh), G["f_accuracy_matrix_q"], rtol=1e-13)
    assert np.isclose(DC.normalized_mse_matrix(J, Jh), G["normalized_mse_matrix_J"], rtol=1e-13)
    assert np.isclose(DC.l2_accuracy(J, Jh), G["l2_accuracy_J"], rtol=1e-13)
    assert np.isclose(DC.f_accuracy_matrix(J, Jh), G["f_accuracy_matrix_J"], rtol=1e-13)


def test_keras_metric_aggregation_from_kernel_statistics():
    """The per-batch sums the loss kernels emit reproduce all six reference numbers."""
    from dino_b200.model import _combine_stats
    from dino_b200 import _lib
    q, qh, J, Jh = (G[k] for k in ("loss_q", "loss_qh", "loss_J", "loss_Jh"))
    B = q.shape[0]
    st = np.zeros((1, _lib.STAT_SLOTS))
    eq, nq = ((q - qh) ** 2).sum(1), (q ** 2).sum(1)
    eJ, nJ = ((J - Jh) ** 2).sum((1, 2)), (J ** 2).sum((1, 2))
    st[0, _lib.STAT_L2_SUM], st[0, _lib.STAT_H1_SUM] = (eq / nq).sum(), (eJ / nJ).sum()
    st[0, _lib.STAT_EQ_SUM], st[0, _lib.STAT_NQ_SUM] = eq.sum(), nq.sum()
    st[0, _lib.STAT_JROW_SUM] = (((J - Jh) ** 2).sum(2) / (J ** 2).sum(2)).sum()
    st[0, _lib.STAT_COUNT], st[0, _lib.STAT_JROWS] = B, B * J.shape[1]
    v = _combine_stats(st, [0.5, 2.0], True)
    want = [0.5 * G["normalized_mse_q"] + 2.0 * G["normalized_mse_matrix_J"], G["normalized_mse_q"],
            G["normalized_mse_matrix_J"], G["l2_accuracy_q"], G["f_accuracy_matrix_q"], G["l2_accuracy_J"],
            G["f_accuracy_matrix_J"]]
    assert np.allclose(v, want, rtol=1e-12)


def test_keras_metric_aggregation_with_a_short_last_batch():
    """n % batch_size != 0: the three losses are batch-size weighted, the four accuracies count each batch once."""
    from dino_b200.model import _combine_stats
    from dino_b200 import _lib
    st = np.zeros((2, _lib.STAT_SLOTS))
    for i, (cnt, l2, h1) in enumerate([(32, 0.04, 0.09), (5, 0.16, 0.25)]):
        st[i, _lib.STAT_COUNT], st[i, _lib.STAT_JROWS] = cnt, cnt * 3
        st[i, _lib.STAT_L2_SUM], st[i, _lib.STAT_H1_SUM] = l2 * cnt, h1 * cnt
        st[i, _lib.STAT_EQ_SUM], st[i, _lib.STAT_NQ_SUM] = l2 * cnt, cnt
        st[i, _lib.STAT_JROW_SUM] = h1 * cnt * 3
    v = _combine_stats(st, [1.0, 1.0], True)
    wl2, wh1 = (32 * 0.04 + 5 * 0.16) / 37, (32 * 0.09 + 5 * 0.25) / 37
    assert np.allclose(v[:3], [wl2 + wh1, wl2, wh1], rtol=1e-12)
    assert np.allclose(v[3:], [1 - (0.2 + 0.4) / 2, 1 - (0.2 + 0.4) / 2, 1 - (0.3 + 0.5) / 2, 1 - (0.3 + 0.5) / 2], rtol=1e-12)


def _svd_data():
    return {k: G["svd_" + k] for k in ("m_data", "q_data", "U_data", "sigma_data", "V_data")}


def test_remap_jacobian_data():
    d = _svd_data()
    assert np.allclose(O.remap_jacobian(d["U_data"], d["sigma_data"], d["V_data"]), G["remap_J"], rtol=1e-13, atol=1e-14)
    assert np.allclose(DC.remap_jacobian_data(d)["J_data"], G["remap_J"], rtol=1e-13, atol=1e-14)


def test_remap_with_orth_V():
    d = _svd_data()
    dQ = d["q_data"].shape[1]
    d5 = {k: (v[..., :dQ] if k in ("U_data", "sigma_data", "V_data") else v) for k, v in d.items()}
    out = DC.remap_jacobian_data_with_orth_V(d5)
    assert np.allclose(out["J_data"], G["orth_J"], atol=1e-13)
    assert np.allclose(out["JVVT_data"], G["orth_JVVT"], atol=1e-12)
    assert np.allclose(np.abs(out["V_data"]), np.abs(G["orth_V"]), atol=1e-12)


def test_flatten_data_variants():
    d = _svd_data()
    cases = {"a": dict(target_rank=8, batch_rank=4, order_random=True, seed=0, burn_in=0),
             "b": dict(target_rank=6, batch_rank=2, order_random=True, seed=3, burn_in=2),
             "c": dict(target_rank=8, batch_rank=8, order_random=False)}
    for tag, kw in cases.items():
        mine = DC.flatten_data(d, **kw)
        orc = O.flatten(d["m_data"], d["q_data"], d["U_data"], d["sigma_data"], d["V_data"], kw["target_rank"],
                        kw["batch_rank"], kw.get("order_random", True), kw.get("seed", 0), kw.get("burn_in", 0))
        for i, k in enumerate(("m_data", "q_data", "U_data", "sigma_data", "V_data")):
            assert np.array_equal(mine[k], G[f"flat_{tag}_{k}"]), (tag, k)
            assert np.array_equal(orc[i], G[f"flat_{tag}_{k}"]), (tag, k)


def test_train_test_split():
    tr, te = DC.train_test_split(_svd_data(), n_train=6)
    for k in tr:
        assert np.array_equal(tr[k], G["split_train_" + k]) and np.array_equal(te[k], G["split_test_" + k])


def test_prune_the_data():
    V, Phi, bias, J = G["prune_V"], G["prune_Phi"], G["prune_bias"], G["remap_J"]
    d = _svd_data()
    for tag, (ip, op) in (("in", (True, False)), ("inout", (True, True))):
        trd = {"m_data": d["m_data"][:5].copy(), "q_data": d["q_data"][:5].copy(), "J_data": J[:5].copy()}
        ted = {"m_data": d["m_data"][5:].copy(), "q_data": d["q_data"][5:].copy(), "J_data": J[5:].copy()}
        trd, ted = DC.prune_the_data({"input": V, "output": Phi, "last_layer_bias": bias}, trd, ted,
                                     input_pruning=ip, output_pruning=op)
        for k in ("m_data", "q_data", "J_data", "J_full", "m_full"):
            assert np.allclose(trd[k], G[f"prune_{tag}_train_{k}"], rtol=1e-12, atol=1e-13), (tag, k)
            assert np.allclose(ted[k], G[f"prune_{tag}_test_{k}"], rtol=1e-12, atol=1e-13), (tag, k)
        m, q, Jp = O.prune(V, Phi, bias, d["m_data"][:5], d["q_data"][:5], J[:5], ip, op)
        assert np.allclose(m, G[f"prune_{tag}_train_m_data"], atol=1e-13)
        assert np.allclose(q, G[f"prune_{tag}_train_q_data"], atol=1e-13)
        assert np.allclose(Jp, G[f"prune_{tag}_train_J_data"], atol=1e-13)


def test_settings_dictionaries_match_reference_keys_and_defaults():
    js = DC.jacobian_network_settings({})
    opt = js.pop("opt_parameters")
    js.pop("problem_settings")
    assert sorted(js.keys()) == list(G["settings_keys"])
    assert [repr(js[k]) for k in sorted(js.keys())] == list(G["settings_repr"])
    assert sorted(opt.keys()) == list(G["opt_keys"])
    assert [repr(opt[k]) for k in sorted(opt.keys())] == list(G["opt_repr"])


def test_observable_settings_dictionary_matches_reference_keys_and_defaults():
    """observable_network_settings (observableConstruction.py:38-78): every key and default, pinned to the reference."""
    osd = DC.observable_network_settings({})
    opt = osd.pop("opt_parameters")
    osd.pop("problem_settings")
    assert sorted(osd.keys()) == list(G["obs_settings_keys"])
    assert [repr(osd[k]) for k in sorted(osd.keys())] == list(G["obs_settings_repr"])
    assert sorted(opt.keys()) == list(G["opt_keys"])
