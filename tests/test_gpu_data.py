"""-m gpu tests of the device-side data contractions (dino_b200/construction/deviceData.py) against the golden
vectors the REFERENCE's own host functions produced (tests/golden/make_golden.py): remap_jacobian_data
(dataUtilities.py:492-510) and prune_the_data (jacobianTrainingDriver.py:93-185).  The reference computes in float64;
the device path in float32 on the fp32x3 / fp32 engines: 1e-5 relative to the largest entry."""
import os

import numpy as np
import pytest

import helpers as H
from dino_b200.construction.deviceData import remap_jacobian_data_device, prune_the_data_device

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "reference_golden.npz"))


def _svd_data():
    return {k: G["svd_" + k] for k in ("m_data", "q_data", "U_data", "sigma_data", "V_data")}


def test_remap_jacobian_data_on_the_device(cuda_device):
    out = remap_jacobian_data_device(_svd_data())
    assert out["J_data"].dtype == np.float32
    assert H.rel_err(out["J_data"], G["remap_J"]) < 1e-5


@pytest.mark.parametrize("tag,ip,op", [("in", True, False), ("inout", True, True)])
def test_prune_the_data_on_the_device(cuda_device, tag, ip, op):
    d = _svd_data()
    proj = {"input": G["prune_V"], "output": G["prune_Phi"], "last_layer_bias": G["prune_bias"]}
    trd = {"m_data": d["m_data"][:5].copy(), "q_data": d["q_data"][:5].copy(), "J_data": G["remap_J"][:5].copy()}
    ted = {"m_data": d["m_data"][5:].copy(), "q_data": d["q_data"][5:].copy(), "J_data": G["remap_J"][5:].copy()}
    trd, ted = prune_the_data_device(proj, trd, ted, input_pruning=ip, output_pruning=op)
    for k in ("m_data", "q_data", "J_data", "J_full", "m_full"):
        assert H.rel_err(trd[k], G[f"prune_{tag}_train_{k}"]) < 1e-5, k
        assert H.rel_err(ted[k], G[f"prune_{tag}_test_{k}"]) < 1e-5, k


def test_large_remap_matches_numpy(cuda_device):
    """A shape of the real problems' order (d_Q = 50, d_M = 4225, rank 50): device vs float64 numpy."""
    rng = np.random.default_rng(0)
    n, dQ, dM, k = 6, 50, 4225, 50
    d = {"m_data": rng.normal(size=(n, dM)), "q_data": rng.normal(size=(n, dQ)), "U_data": rng.normal(size=(n, dQ, k)),
         "sigma_data": np.abs(rng.normal(size=(n, k))), "V_data": rng.normal(size=(n, dM, k))}
    out = remap_jacobian_data_device(d)
    want = np.einsum("nqk,nk,nmk->nqm", d["U_data"], d["sigma_data"], d["V_data"])
    assert H.rel_err(out["J_data"], want) < 1e-5


@pytest.mark.parametrize("burn_in", [0, 3])
def test_flatten_data_on_the_device_is_the_host_function(cuda_device, burn_in):
    """flatten_data (dataUtilities.py:348-445) as device data movement: the same column permutation, grouping and sample
    replication as the host function, element for element (float32 casts of the same values)."""
    from dino_b200.construction import flatten_data
    from dino_b200.construction.deviceData import flatten_data_device, upload_factors
    d = _svd_data()
    rank = d["sigma_data"].shape[1]
    br = 2 if rank % 2 == 0 else 1
    tr = (rank // br) * br
    host = flatten_data(d, target_rank=tr, batch_rank=br, order_random=True, burn_in=burn_in)
    dev = flatten_data_device(upload_factors(d), target_rank=tr, batch_rank=br, order_random=True, burn_in=burn_in)
    assert set(dev) == set(host)
    for k in host:
        assert dev[k].is_cuda
        assert np.array_equal(dev[k].cpu().numpy(), np.asarray(host[k], dtype=np.float32)), k
