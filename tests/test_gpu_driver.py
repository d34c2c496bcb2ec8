"""-m gpu: the complete driver pipeline (jacobianTrainingDriver.py:346-396) through the public API on a
synthetic data directory in the reference's on-disk format, with parity of the trained weights against the
oracle run on the same reduced data, initial weights and minibatch order."""
import os
import pickle

import numpy as np
import pytest
import torch

import helpers as H
from oracle import dino_oracle as O
import dino_b200 as D
from dino_b200 import synthetic as S

pytestmark = pytest.mark.gpu


def _make_data_dir(tmp_path, n=96, dM=40, dQ=10, seed=3):
    rng = np.random.default_rng(seed)
    arch = S.Arch("generic_dense", dM, dQ, hidden=[16, 16])
    tw = S.random_weights(arch, rng, teacher=True)
    m = rng.normal(size=(n, dM))
    q, J = S.teacher_eval(arch, tw, m)
    J = J.astype(np.float64)
    U, s, Vt = np.linalg.svd(J, full_matrices=False)                    # [n,dQ,dQ], [n,dQ], [n,dQ,dM]
    d = str(tmp_path) + "/"
    half = n // 2
    for r, (a, b) in enumerate([(0, half), (half, n)]):
        np.savez(d + f"mq_on_proc{r}.npz", m_data=m[a:b], q_data=q[a:b])
        np.savez(d + f"J_on_proc{r}.npz", U_data=U[a:b], sigma_data=s[a:b], V_data=np.transpose(Vt[a:b], (0, 2, 1)))
    np.save(d + "AS_input_projector.npy", rng.normal(size=(dM, 12)))
    np.save(d + "AS_output_projector.npy", rng.normal(size=(dQ, 6)))
    np.save(d + "AS_d_GN.npy", np.ones(12)); np.save(d + "AS_d_NG.npy", np.ones(6))
    return d, dict(m=m, q=q, J=J)


def _settings(data_dir, tmp_path, **kw):
    settings = D.jacobian_network_settings({})
    settings.update(data_dir=data_dir, architecture="rb_dense", depth=3, train_full_jacobian=1,
                    fixed_input_rank=12, fixed_output_rank=6, train_data_size=64, test_data_size=32,
                    reduced_input_training=True, reduced_output_training=False,
                    weights_dir=str(tmp_path) + "/trained_weights/", network_name="dino_test")
    settings["opt_parameters"].update(train_full_jacobian=1, keras_epochs=3, keras_batch_size=16, keras_verbose=False)
    settings.update(kw)
    return settings


def test_driver_end_to_end_and_weight_pickle(cuda_device, tmp_path):
    data_dir, _ = _make_data_dir(tmp_path)
    settings = _settings(data_dir, tmp_path)
    np.random.seed(7)
    regressor = D.jacobian_training_driver(settings)
    # the restitched network takes the FULL input and carries the projection layers
    names = [l.name for l in regressor.layers]
    assert "jacobian_input_proj_layer" in names and "jacobian_output_layer" in names
    assert regressor.inputs[0].shape == (None, 40) and len(regressor.outputs) == 2
    path = settings["weights_dir"] + "dino_test.pkl"
    assert os.path.isfile(path)
    with open(path, "rb") as f:
        weights = pickle.load(f)
    assert set(weights) == set(names)
    loaded = D.observable_network_loader(settings, file_name=path)
    for layer in loaded.layers:
        for a, b in zip(layer.get_weights(), weights[layer.name]):
            assert np.array_equal(a, b)
    # evaluateJacobianNetwork on the same files reports the reference's dictionary keys
    settings["full_jacobian"] = True
    out = D.evaluation.evaluateJacobianNetwork(settings, jacobian_network=loaded, data_dir=data_dir)
    assert {"l2_train", "h1_train", "l2_test", "h1_test", "train_eval", "test_eval"} <= set(out)
    assert np.isfinite(out["l2_train"]) and np.isfinite(out["h1_train"])


def test_driver_with_device_side_data_contractions(cuda_device, tmp_path):
    """settings['device_data']: get_the_data (remap_jacobian_data) and prune_the_data as device contractions (float32)
    hand the training stage the same reduced arrays as the reference's host float64 path (1e-5), and the whole driver
    runs through with them."""
    data_dir, _ = _make_data_dir(tmp_path)
    out = []
    for dev_data in (False, True):
        settings = _settings(data_dir, tmp_path)
        settings["device_data"] = dev_data
        dd = D.get_the_data(settings)
        train_dict, test_dict = dd["train_dict"], dd["test_dict"]
        proj = D.setup_reduced_bases(settings, train_dict)
        if dev_data:
            from dino_b200.construction.deviceData import prune_the_data_device as prune
        else:
            prune = D.prune_the_data
        train_dict, test_dict = prune(proj, train_dict, test_dict, input_pruning=True, output_pruning=False)
        out.append((train_dict, test_dict))
    for k in ("m_data", "q_data", "J_data"):
        assert H.rel_err(out[1][0][k], out[0][0][k]) < 1e-5, k
        assert H.rel_err(out[1][1][k], out[0][1][k]) < 1e-5, k
    settings = _settings(data_dir, tmp_path)
    settings["device_data"] = True
    reg = D.jacobian_training_driver(settings)
    assert all(np.isfinite(w).all() for l in reg.layers for w in l.get_weights())


@pytest.mark.parametrize("precision,tol", [("fp32x3", 5e-4), ("fp32", 1e-4), ("tf32", 2e-2)])
def test_driver_stages_match_oracle(cuda_device, tmp_path, precision, tol, monkeypatch):
    """get_the_data -> setup_reduced_bases -> prune_the_data -> setup_the_dino -> train_dino, then the same reduced
    problem through the oracle's fit with identical initial weights and permutations."""
    monkeypatch.setenv("DINO_B200_PRECISION", precision)
    data_dir, _ = _make_data_dir(tmp_path)
    settings = _settings(data_dir, tmp_path)
    dd = D.get_the_data(settings)
    train_dict, test_dict = dd["train_dict"], dd["test_dict"]
    proj = D.setup_reduced_bases(settings, train_dict)
    settings["reduced_input_dim"] = proj["input"].shape[1]
    train_dict, test_dict = D.prune_the_data(proj, train_dict, test_dict, input_pruning=True, output_pruning=False)
    regressor = D.setup_the_dino(settings, train_dict, proj, reduced_input_training=True)
    assert regressor.precision == precision
    # oracle network with the same initial weights (output layer frozen, like the builder's)
    net = O.ONet(regressor.core.d_in)
    core = regressor.core
    for l, name in enumerate(core.names):
        w = core.layer_weights(l)
        net.layers.append(O.OLayer(name, torch.tensor(w[0], dtype=torch.float64),
                                   torch.tensor(w[1], dtype=torch.float64) if len(w) == 2 else None,
                                   core.act[l], core.trainable[l], core.skip_from[l]))
    np.random.seed(11)
    perms = [np.random.permutation(64) for _ in range(3)]
    np.random.seed(11)                                   # model.fit draws the same permutations from the global RNG
    regressor = D.train_dino(settings, regressor, train_dict, test_dict)
    t = lambda a: torch.tensor(np.asarray(a, dtype=np.float32).astype(np.float64))
    ob = O.Batch(t(train_dict["m_data"]), t(train_dict["q_data"]), t(train_dict["J_data"]))
    O.fit(net, ob, epochs=3, batch_size=16, perms=perms)
    ow = net.get_weights()
    for layer in regressor.layers:
        for a, b in zip(layer.get_weights(), ow.get(layer.name, [])):
            assert H.rel_err(a, b) < tol, layer.name
    # reported accuracies (Keras evaluate semantics) agree with the oracle's evaluate
    regressor.compile(optimizer=D.Adam(1e-3), loss=[D.normalized_mse, D.normalized_mse_matrix],
                      loss_weights=[1.0, 1.0], metrics=[D.l2_accuracy, D.f_accuracy_matrix])
    ev = regressor.evaluate([train_dict["m_data"]], [train_dict["q_data"], train_dict["J_data"]])
    ref = O.evaluate(net, ob, batch_size=32)
    assert np.allclose(ev, ref, rtol=10 * tol, atol=1e-5)


def test_jacobian_error_metrics_from_arrays(cuda_device, tmp_path):
    """compute_jacobian_errors (jacobianTests.py:104-126) fed from arrays, prediction from the CUDA chain."""
    data_dir, raw = _make_data_dir(tmp_path)
    arch = S.Arch("generic_dense", 40, 10, hidden=[16, 16])
    w = H.student_weights(arch, 5)
    model = S.build_model(arch, w)
    truth = {"m_data": raw["m"][:6], "J_data": raw["J"][:6]}
    log = D.evaluation.compute_jacobian_errors(truth, {"net": model}, n_samples=6)
    net = H.oracle_net(arch, w)
    _, Jo = O.batch_jacobian(net, torch.tensor(raw["m"][:6].astype(np.float32).astype(np.float64)), False)
    for i in range(6):
        want = O.jacobian_errors(Jo[i].numpy(), raw["J"][i])
        for k, v in want.items():
            assert abs(log["net"][k][i] - v) < 1e-4 * max(1.0, abs(v)), k
    # InferenceOracle array methods
    io = D.inference.InferenceOracle(None, model)
    m0 = raw["m"][0]
    assert H.rel_err(io.J(m0), Jo[0].numpy()) < 1e-5
    mh, qh = np.ones(40), np.ones(10)
    assert H.rel_err(io.Jvp(m0, mh), Jo[0].numpy() @ mh) < 1e-5
    assert H.rel_err(io.Jtvp(m0, qh), Jo[0].numpy().T @ qh) < 1e-5


def test_sketched_driver_with_column_reshuffle(cuda_device, tmp_path):
    """train_full_jacobian = 0: flatten_data -> equip_model_with_sketched_jacobian -> train_h1_network on
    [m, U, V] -> [q, diag sigma], with shuffle_every_epoch re-drawing the sketch columns (jacobianTrainingDriver.py:227-246)."""
    data_dir, _ = _make_data_dir(tmp_path)
    settings = _settings(data_dir, tmp_path, architecture="generic_dense", train_full_jacobian=0, target_rank=10,
                         batch_rank=5, truncation_dimension=12, depth=2, reduced_input_training=False,
                         shuffle_every_epoch=True, outer_epochs=2, inner_epochs=1)
    settings["opt_parameters"].update(train_full_jacobian=0, keras_epochs=1, keras_batch_size=16)
    np.random.seed(3)
    regressor = D.jacobian_training_driver(settings)
    assert len(regressor.inputs) == 3 and regressor.outputs[1].shape == (None, 5, 5)
    # one more evaluation against the oracle's sketched Jacobian on flattened data
    dd = D.get_the_data(settings)
    tr = dd["train_dict"]
    assert tr["U_data"].shape == (64 * 2, 10, 5) and tr["sigma_data"].shape == (128, 5, 5)
    regressor.compile(optimizer=D.Adam(1e-3), loss=[D.normalized_mse, D.normalized_mse_matrix], loss_weights=[1.0, 1.0],
                      metrics=[D.l2_accuracy, D.f_accuracy_matrix])
    ev = regressor.evaluate([tr["m_data"], tr["U_data"], tr["V_data"]], [tr["q_data"], tr["sigma_data"]])
    net = O.ONet(regressor.core.d_in)
    core = regressor.core
    for l, name in enumerate(core.names):
        w = core.layer_weights(l)
        net.layers.append(O.OLayer(name, torch.tensor(w[0], dtype=torch.float64),
                                   torch.tensor(w[1], dtype=torch.float64) if len(w) == 2 else None,
                                   core.act[l], core.trainable[l], core.skip_from[l]))
    t = lambda a: torch.tensor(np.asarray(a, dtype=np.float32).astype(np.float64))
    ob = O.Batch(t(tr["m_data"]), t(tr["q_data"]), t(tr["sigma_data"]), t(tr["U_data"]), t(tr["V_data"]))
    ref = O.evaluate(net, ob, batch_size=32, jac_mode="sketched")
    assert np.allclose(ev, ref, rtol=1e-3, atol=1e-5)


def test_observable_training_driver_l2_only(cuda_device, tmp_path):
    """observable_training_driver (observableConstruction.py:81-116): plain L2 training of m -> q."""
    data_dir, _ = _make_data_dir(tmp_path)
    settings = D.observable_network_settings({})
    settings.update(data_dir=data_dir, architecture="rb_dense", depth=2, fixed_input_rank=12, fixed_output_rank=6,
                    train_data_size=64, test_data_size=32, reduced_input_dim=None, reduced_output_dim=None)
    settings["opt_parameters"].update(keras_epochs=2, keras_batch_size=16)
    net = D.observable_training_driver(settings)
    assert len(net.outputs) == 1
    ev = net.evaluate(np.zeros((4, 40)) + 0.1, np.ones((4, 10)))
    assert len(ev) == 2 and np.isfinite(ev).all()


def test_fit_generator_matches_train_on_batch(cuda_device):
    """The pipelined streaming call takes the same steps as one train_on_batch per minibatch."""
    import torch
    arch = H.SMALL["dense"]
    data, tw = S.make_dataset(arch, 48, 5)
    w = H.student_weights(arch, 6, tw, bias_noise=0.1)
    batches = [([data["m"][i:i + 12]], [data["q"][i:i + 12], data["J"][i:i + 12]]) for i in range(0, 48, 12)]

    def build():
        m = S.build_model(arch, w, precision="fp32")
        m.compile(optimizer=Adam(1e-3), loss=[normalized_mse, normalized_mse_matrix], loss_weights=[1.0, 1.0],
                  metrics=[l2_accuracy, f_accuracy_matrix])
        return m

    from dino_b200 import Adam
    from dino_b200.construction import normalized_mse, normalized_mse_matrix, l2_accuracy, f_accuracy_matrix
    a, b = build(), build()
    rows_a = [a.train_on_batch(x, y) for x, y in batches]
    rows_b = b.fit_generator(iter(batches))
    assert len(rows_b) == len(rows_a)
    np.testing.assert_allclose(np.array(rows_b), np.array(rows_a), rtol=1e-5, atol=1e-7)
    for la, lb in zip(a.layers, b.layers):
        for wa, wb in zip(la.get_weights(), lb.get_weights()):
            np.testing.assert_allclose(wb, wa, rtol=1e-5, atol=1e-7)


def test_sketched_driver_reshuffles_on_the_device(cuda_device, tmp_path):
    """settings['device_data'] with shuffle_every_epoch: the factors are uploaded once and every outer epoch re-flattens them
    on the device (deviceData.flatten_data_device); DinoModel.fit takes the device tensors as they are.  Same seeds ->
    the training data are the host path's element for element, so the trained weights agree to summation order."""
    data_dir, _ = _make_data_dir(tmp_path)
    out = []
    for dev_data in (False, True):
        settings = _settings(data_dir, tmp_path, architecture="generic_dense", train_full_jacobian=0, target_rank=10,
                             batch_rank=5, truncation_dimension=12, depth=2, reduced_input_training=False,
                             shuffle_every_epoch=True, outer_epochs=2, inner_epochs=1)
        settings["opt_parameters"].update(train_full_jacobian=0, keras_epochs=1, keras_batch_size=16)
        settings["device_data"] = dev_data
        np.random.seed(3)                      # the minibatch order of fit
        from dino_b200.construction.neuralNetworks import set_seed
        set_seed(11)                           # the weight initialiser
        regressor = D.jacobian_training_driver(settings)
        out.append([x for l in regressor.layers for x in l.get_weights()])
    for a, b in zip(*out):
        assert np.isfinite(b).all()
        assert H.rel_err(b, a) < 1e-4
