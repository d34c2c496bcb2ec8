"""CPU tests of the host mirror: C-ABI symbols, builders, weights, on-disk formats, driver stages
that run before the first kernel.  No compute call is made without a GPU."""
import os
import pickle
import re

import numpy as np
import pytest

import dino_b200 as D
from dino_b200 import _lib, parallel, synthetic as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_loads_and_exports_every_declared_symbol():
    header = open(os.path.join(ROOT, "include", "dino_b200.h")).read()
    declared = set(re.findall(r"\b(dino_[a-z0-9_]+)\s*\(", header))
    assert declared, "no declarations parsed"
    lib = _lib.load()
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/dino_b200.h but not exported"
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    assert lib.dino_version().startswith(b"dino_b200")


def test_missing_library_is_a_loud_error(monkeypatch):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libdino_b200.so")
    with pytest.raises(_lib.DinoLibraryError):
        _lib.load()


def test_no_product_module_imports_the_oracle():
    """Nothing under dino_b200/ or dino/ imports, loads or executes anything under oracle/ (the word itself also names
    the reference's InferenceOracle objects, so the check is on import statements and paths)."""
    import re
    pat = re.compile(r"^\s*(from|import)\s+oracle\b|dino_oracle|oracle/|[\"']oracle[\"']", re.M)
    for top in ("dino_b200", "dino"):
        for base, _, files in os.walk(os.path.join(ROOT, top)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h")):
                    assert not pat.search(open(os.path.join(base, f)).read()), (base, f)


def test_projected_dense_structure_and_names():
    V = np.linalg.qr(np.random.default_rng(0).normal(size=(30, 8)))[0]
    Phi = np.linalg.qr(np.random.default_rng(1).normal(size=(12, 5)))[0]
    net = D.projected_dense(input_projector=V, last_layer_weights=[Phi.T, np.ones(12)],
                            hidden_layer_dimensions=3 * [5], name_prefix="jacobian_")
    names = [l.name for l in net.layers]
    assert names == ["jacobian_input_data", "jacobian_input_proj_layer", "jacobian_dense_reduction_layer",
                     "jacobian_inner_layer_0", "jacobian_inner_layer_1", "jacobian_inner_layer_2",
                     "jacobian_output_compat_layer", "jacobian_output_layer"]
    assert not net.get_layer("jacobian_input_proj_layer").trainable
    assert not net.get_layer("jacobian_output_layer").trainable
    assert np.allclose(net.get_layer("jacobian_input_proj_layer").get_weights()[0], V, atol=1e-7)
    assert len(net.get_layer("jacobian_output_compat_layer").get_weights()) == 1      # use_bias=False
    assert len(net.inputs) == 1 and len(net.outputs) == 1
    assert net.inputs[0].shape == (None, 30) and net.outputs[0].shape == (None, 12)
    j = D.equip_model_with_full_jacobian(net)
    assert len(j.outputs) == 2 and j.outputs[1].shape == (None, 12, 30)
    sk = D.equip_model_with_sketched_jacobian(net, 4)
    assert len(sk.inputs) == 3 and sk.outputs[1].shape == (None, 4, 4)
    orj = D.equip_model_with_output_reduced_jacobian(net, Phi)
    assert orj.outputs[1].shape == (None, 5, 30)
    # the equipped models share the layers (and weights) of the base model, like Keras
    net.get_layer("jacobian_inner_layer_0").set_weights([np.zeros((8, 5)), np.ones(5)])
    assert np.all(j.get_layer("jacobian_inner_layer_0").get_weights()[1] == 1)
    with pytest.raises(AssertionError):
        D.projected_dense(reduced_input_dim=4, reduced_output_dim=3, hidden_layer_dimensions=[5])


def test_resnet_and_generic_builders():
    Phi = np.linalg.qr(np.random.default_rng(1).normal(size=(12, 5)))[0]
    net = D.projected_resnet(None, [Phi.T, np.zeros(12)], ranks=[8, 8], name_prefix="p_", reduced_input_dim=9)
    names = [l.name for l in net.layers]
    assert names == ["p_input_data", "p_input_resnet_interface", "p_0low_rank_residual_in", "p_0low_rank_residual_out",
                     "p_add0", "p_1low_rank_residual_in", "p_1low_rank_residual_out", "p_add1",
                     "p_output_compat_layer", "p_output_layer"]
    assert np.all(net.get_layer("p_0low_rank_residual_out").get_weights()[0] == 0)       # zero-initialised
    assert net.get_layer("p_add0").get_weights() == []
    assert net.core.skip_from == [-1, -1, 1, -1, 3, -1, -1]
    g = D.generic_dense(20, 3, [7, 7])
    assert [l.name for l in g.layers] == ["input_data", "inner_layer_0", "inner_layer_1", "final_dense"]
    with pytest.raises(Exception):
        D.choose_network({"depth": 2, "truncation_dimension": 3, "architecture": "as_dense", "name_prefix": ""})


def test_weight_pickle_round_trip_through_loader(tmp_path):
    rng = np.random.default_rng(3)
    V = np.linalg.qr(rng.normal(size=(30, 8)))[0]
    Phi = np.linalg.qr(rng.normal(size=(12, 5)))[0]
    settings = D.observable_network_settings({})
    settings.update(architecture="rb_dense", depth=3, name_prefix="jacobian_", reduced_input_dim=None,
                    reduced_output_dim=None)
    proj = {"input": V, "output": Phi, "last_layer_bias": rng.normal(size=12)}
    net = D.choose_network(settings, proj)
    for layer in net.layers:
        layer.set_weights([rng.normal(size=w.shape).astype(np.float32) for w in layer.get_weights()])
    weights = {layer.name: layer.get_weights() for layer in net.layers}
    path = str(tmp_path / "w.pkl")
    with open(path, "wb") as f:
        pickle.dump(weights, f, pickle.HIGHEST_PROTOCOL)
    loaded = D.observable_network_loader(settings, file_name=path)
    assert settings["compat_layer"] is True
    for layer in loaded.layers:
        for a, b in zip(layer.get_weights(), weights[layer.name]):
            assert np.array_equal(a, b)


def _write_data_dir(tmp_path, n=12, dM=10, dQ=6, R=4, legacy=False):
    rng = np.random.default_rng(5)
    d = str(tmp_path) + "/"
    parts = [(0, 7), (7, n)]
    m, q = rng.normal(size=(n, dM)), rng.normal(size=(n, dQ))
    U, s, V = rng.normal(size=(n, dQ, R)), rng.uniform(0.5, 1.5, size=(n, R)), rng.normal(size=(n, dM, R))
    for r, (a, b) in enumerate(parts):
        if legacy:
            np.save(d + f"ms_on_proc_{r}.npy", m[a:b]); np.save(d + f"qs_on_proc_{r}.npy", q[a:b])
            np.save(d + f"Us_on_proc_{r}.npy", U[a:b]); np.save(d + f"sigmas_on_proc_{r}.npy", s[a:b])
            np.save(d + f"Vs_on_proc_{r}.npy", V[a:b])
        else:
            np.savez(d + f"mq_on_proc{r}.npz", m_data=m[a:b], q_data=q[a:b])
            np.savez(d + f"J_on_proc{r}.npz", U_data=U[a:b], sigma_data=s[a:b], V_data=V[a:b])
    np.save(d + "AS_input_projector.npy", rng.normal(size=(dM, 6)))
    np.save(d + "AS_output_projector.npy", rng.normal(size=(dQ, 5)))
    np.save(d + "AS_d_GN.npy", np.array([1, .5, .1, 1e-3, 1e-5, 1e-6]))
    np.save(d + "AS_d_NG.npy", np.array([1, .5, .1, 1e-3, 1e-6]))
    np.save(d + "KLE_projector.npy", rng.normal(size=(dM, 6)))
    np.save(d + "KLE_d.npy", np.array([1, .5, .1, 1e-3, 1e-5, 1e-6]))
    return d, dict(m=m, q=q, U=U, s=s, V=V)


@pytest.mark.parametrize("legacy", [False, True])
def test_load_data_formats(tmp_path, legacy):
    d, ref = _write_data_dir(tmp_path, legacy=legacy)
    out = D.load_data(d, derivatives=True, n_data=11)
    assert np.array_equal(out["m_data"], ref["m"][:11]) and np.array_equal(out["q_data"], ref["q"][:11])
    assert np.array_equal(out["U_data"], ref["U"][:11]) and np.array_equal(out["sigma_data"], ref["s"][:11])
    with pytest.raises(AssertionError):
        D.load_data(d, n_data=100)


def test_get_projectors_and_driver_stages_before_training(tmp_path):
    d, ref = _write_data_dir(tmp_path)
    p = D.get_projectors(d, fixed_input_rank=4, fixed_output_rank=3)
    assert p["AS_input"].shape == (10, 4) and p["AS_output"].shape == (6, 3) and p["KLE"].shape == (10, 4)
    p = D.get_projectors(d)
    assert p["AS_input"].shape == (10, 4) and p["AS_output"].shape == (6, 4)       # spectrum > 1e-4
    settings = D.jacobian_network_settings({})
    settings.update(data_dir=d, architecture="rb_dense", train_full_jacobian=True, train_data_size=8,
                    test_data_size=4, fixed_input_rank=4, fixed_output_rank=3, depth=2,
                    reduced_input_training=True, reduced_output_training=False)
    dd = D.get_the_data(settings)
    tr, te = dd["train_dict"], dd["test_dict"]
    assert tr["J_data"].shape == (8, 6, 10) and te["m_data"].shape == (4, 10)
    assert np.array_equal(tr["m_data"], ref["m"][-8:])                               # LAST n_train rows train
    proj = D.setup_reduced_bases(settings, tr)
    assert np.allclose(proj["input"].T @ proj["input"], np.eye(4), atol=1e-12)
    assert np.allclose(proj["last_layer_bias"], ref["q"][-8:].mean(0))
    settings["reduced_input_dim"] = 4
    tr, te = D.prune_the_data(proj, tr, te, input_pruning=True, output_pruning=False)
    assert tr["m_data"].shape == (8, 4) and tr["J_data"].shape == (8, 6, 4) and tr["m_full"].shape == (8, 10)
    net = D.setup_the_dino(settings, tr, proj, reduced_input_training=True)
    assert len(net.outputs) == 2 and net.outputs[1].shape == (None, 6, 4)
    assert settings["opt_parameters"]["train_full_jacobian"] is True
    # kle basis selection works despite the reference's 'KLE' / 'KLE_input' key mismatch
    settings["input_basis"] = "kle"
    assert D.setup_reduced_bases(settings, dd["train_dict"])["input"].shape == (10, 4)


def test_invalid_optimizer_raises_typeerror_like_the_reference():
    from dino_b200.construction.trainingUtilities import _make_optimizer
    with pytest.raises(TypeError):
        _make_optimizer({"keras_opt": "lbfgs", "keras_alpha": 1e-3})


def test_shard_bounds_cover_the_batch_exactly():
    for B in (1, 5, 32, 255, 256):
        for world in (1, 2, 3, 8):
            spans = [parallel.shard_bounds(B, r, world) for r in range(world)]
            assert spans[0][0] == 0 and sum(c for _, c in spans) == B
            for (o1, c1), (o2, _) in zip(spans, spans[1:]):
                assert o1 + c1 == o2
            assert max(c for _, c in spans) - min(c for _, c in spans) <= 1


def test_synthetic_flop_and_byte_model_matches_survey():
    assert abs(S.flops_per_sample(S.CONFIGS["cfg1"]["arch"]) / 5.87e6 - 1) < 0.01
    assert abs(S.flops_per_sample(S.CONFIGS["cfg2"]["arch"]) / 8.67e6 - 1) < 0.01
    assert S.bytes_per_sample(S.CONFIGS["cfg2"]["arch"]) == 20600
    assert S.bytes_per_sample(S.CONFIGS["cfg3"]["arch"]) == 81200


def test_ctypes_structs_match_the_c_header(tmp_path):
    """The ctypes mirrors of the C-ABI structs (dino_b200/_lib.py) have the size and field offsets gcc gives the structs of
    include/dino_b200.h (a field added on one side only would shift every pointer after it)."""
    import ctypes as C
    import shutil
    import subprocess
    from dino_b200 import _lib
    if shutil.which("gcc") is None:
        pytest.skip("no gcc")
    structs = {"DinoBatch": _lib.DinoBatch, "DinoNetSpec": _lib.DinoNetSpec, "DinoLayerSpec": _lib.DinoLayerSpec}
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "dino_b200.h"', 'int main(void) {']
    for name, cls in structs.items():
        lines.append(f'  printf("{name} size %zu\\n", sizeof({name}));')
        for fld, _ in cls._fields_:
            lines.append(f'  printf("{name} {fld} %zu\\n", offsetof({name}, {fld}));')
    lines += ['  return 0;', '}']
    src = tmp_path / "abi.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "abi"
    inc = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include")
    subprocess.run(["gcc", "-I", inc, str(src), "-o", str(exe)], check=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split("\n")
    seen = 0
    for ln in out:
        if not ln.strip():
            continue
        name, fld, val = ln.split()
        cls = structs[name]
        want = C.sizeof(cls) if fld == "size" else getattr(cls, fld).offset
        assert int(val) == want, ln
        seen += 1
    assert seen == sum(len(c._fields_) + 1 for c in structs.values())
