"""Generate tests/golden/reference_golden.npz by running the REFERENCE's own functions.

Run in the build container only (needs /root/reference, which does not exist on the GPU box):

    python tests/golden/make_golden.py

TensorFlow is not installable here, so the reference modules are imported with a stand-in
``tensorflow`` module whose four elementwise/reduction ops (square, reduce_mean, reduce_sum, sqrt)
are numpy's.  That is enough to execute, unmodified, the reference's
  * loss and metric functions               (dino/construction/trainingUtilities.py:25-51)
  * prune_the_data                          (dino/construction/jacobianTrainingDriver.py:93-185)
  * remap_jacobian_data, flatten_data, train_test_split, remap_jacobian_data_with_orth_V
                                            (dino/construction/dataUtilities.py:348-548)
on seeded inputs.  Inputs and outputs are stored; tests/test_golden.py replays them against the
oracle and against dino_b200's host mirror.  No reference source is copied: it is imported in place.
"""
import importlib.util
import os
import sys
import types

import numpy as np

REF = "/root/reference/dino/construction"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_golden.npz")


def _fake_tf():
    tf = types.ModuleType("tensorflow")
    tf.square, tf.sqrt = np.square, np.sqrt
    tf.reduce_mean = lambda x, axis=None: np.mean(x, axis=axis)
    tf.reduce_sum = lambda x, axis=None: np.sum(x, axis=axis)
    tf.random = types.SimpleNamespace(set_seed=lambda s: None)
    tf.keras = types.SimpleNamespace()
    tf.__version__ = "2.7.0"
    return tf


def load_reference():
    sys.modules["tensorflow"] = _fake_tf()
    for pkg in ("refdino", "refdino.construction"):
        m = types.ModuleType(pkg)
        m.__path__ = []
        sys.modules[pkg] = m
    mods = {}
    for name in ("dataUtilities", "neuralNetworks", "neuralNetworkUtilities", "trainingUtilities",
                 "jacobianConstruction", "jacobianTrainingDriver", "observableConstruction"):
        spec = importlib.util.spec_from_file_location("refdino.construction." + name, os.path.join(REF, name + ".py"))
        mod = importlib.util.module_from_spec(spec)
        sys.modules[spec.name] = mod
        spec.loader.exec_module(mod)
        mods[name] = mod
    return mods


def main():
    ref = load_reference()
    tu, du, drv = ref["trainingUtilities"], ref["dataUtilities"], ref["jacobianTrainingDriver"]
    rng = np.random.default_rng(20240611)
    g = {}
    # ---- losses / metrics -----------------------------------------------------------------------
    B, dQ, dM = 6, 5, 7
    q, qh = rng.normal(size=(B, dQ)), rng.normal(size=(B, dQ))
    J, Jh = rng.normal(size=(B, dQ, dM)), rng.normal(size=(B, dQ, dM))
    g.update(loss_q=q, loss_qh=qh, loss_J=J, loss_Jh=Jh)
    g["normalized_mse_q"] = tu.normalized_mse(q, qh)
    g["l2_accuracy_q"] = tu.l2_accuracy(q, qh)
    g["f_accuracy_matrix_q"] = tu.f_accuracy_matrix(q, qh)
    g["normalized_mse_matrix_J"] = tu.normalized_mse_matrix(J, Jh)
    g["l2_accuracy_J"] = tu.l2_accuracy(J, Jh)
    g["f_accuracy_matrix_J"] = tu.f_accuracy_matrix(J, Jh)
    # ---- SVD-factor data: remap / flatten / split ------------------------------------------------
    n, R = 9, 8
    data = {"m_data": rng.normal(size=(n, dM)), "q_data": rng.normal(size=(n, dQ)),
            "U_data": rng.normal(size=(n, dQ, R)), "sigma_data": rng.uniform(0.1, 2.0, size=(n, R)),
            "V_data": rng.normal(size=(n, dM, R))}
    for k, v in data.items():
        g["svd_" + k] = v
    g["remap_J"] = du.remap_jacobian_data(data)["J_data"]
    # the reference's orth-V variant only works when the stored rank equals d_Q (dataUtilities.py:541-542)
    data5 = {k: (v[..., :dQ] if k in ("U_data", "sigma_data", "V_data") else v) for k, v in data.items()}
    orth = du.remap_jacobian_data_with_orth_V(data5)
    g["orth_V"], g["orth_JVVT"], g["orth_J"] = orth["V_data"], orth["JVVT_data"], orth["J_data"]
    for tag, kw in (("a", dict(target_rank=8, batch_rank=4, order_random=True, seed=0, burn_in=0)),
                    ("b", dict(target_rank=6, batch_rank=2, order_random=True, seed=3, burn_in=2)),
                    ("c", dict(target_rank=8, batch_rank=8, order_random=False))):
        fl = du.flatten_data(data, **kw)
        for k, v in fl.items():
            g[f"flat_{tag}_{k}"] = v
    tr, te = du.train_test_split(data, n_train=6)
    for k in data:
        g["split_train_" + k], g["split_test_" + k] = tr[k], te[k]
    # ---- prune_the_data --------------------------------------------------------------------------
    V = np.linalg.qr(rng.normal(size=(dM, 4)))[0]
    Phi = np.linalg.qr(rng.normal(size=(dQ, 3)))[0]
    bias = rng.normal(size=dQ)
    proj = {"input": V, "output": Phi, "last_layer_bias": bias}
    g.update(prune_V=V, prune_Phi=Phi, prune_bias=bias)
    for tag, (ip, op) in (("in", (True, False)), ("inout", (True, True))):
        trd = {"m_data": data["m_data"][:5].copy(), "q_data": data["q_data"][:5].copy(), "J_data": g["remap_J"][:5].copy()}
        ted = {"m_data": data["m_data"][5:].copy(), "q_data": data["q_data"][5:].copy(), "J_data": g["remap_J"][5:].copy()}
        trd, ted = drv.prune_the_data(proj, trd, ted, input_pruning=ip, output_pruning=op)
        for k in ("m_data", "q_data", "J_data", "J_full", "m_full"):
            g[f"prune_{tag}_train_{k}"], g[f"prune_{tag}_test_{k}"] = trd[k], ted[k]
    # ---- settings dictionaries (keys and defaults) -------------------------------------------------
    js = ref["jacobianConstruction"].jacobian_network_settings({})
    opt = js.pop("opt_parameters")
    js.pop("problem_settings")
    g["settings_keys"] = np.array(sorted(js.keys()))
    g["settings_repr"] = np.array([repr(js[k]) for k in sorted(js.keys())])
    g["opt_keys"] = np.array(sorted(opt.keys()))
    g["opt_repr"] = np.array([repr(opt[k]) for k in sorted(opt.keys())])
    osd = ref["observableConstruction"].observable_network_settings({})
    osd.pop("opt_parameters")
    osd.pop("problem_settings")
    g["obs_settings_keys"] = np.array(sorted(osd.keys()))
    g["obs_settings_repr"] = np.array([repr(osd[k]) for k in sorted(osd.keys())])
    np.savez_compressed(OUT, **g)
    print("wrote", OUT, len(g), "arrays", os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
