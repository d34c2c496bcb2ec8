"""-m gpu test of the data-parallel rule of SURVEY 8e on real GPUs: two ranks (NCCL, one process per GPU) that split
every global minibatch end at the parameters of one rank that takes the whole minibatch - 1e-5 on the parameters after
100 Adam steps, 1e-6 on the loss.  Needs two GPUs in the box (skipped otherwise); the same logic runs on the CPU with
gloo in tests/test_dp_gloo.py."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys
import numpy as np, torch
sys.path.insert(0, os.environ["DINO_ROOT"]); sys.path.insert(0, os.path.join(os.environ["DINO_ROOT"], "tests"))
from dino_b200 import synthetic as S, parallel, Adam
from dino_b200.construction import normalized_mse, normalized_mse_matrix, l2_accuracy, f_accuracy_matrix
dp = parallel.init_from_env("nccl")
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
arch = S.CONFIGS["cfg1"]["arch"]
n, B, epochs = 400, 16, 4                      # 25 minibatches x 4 epochs = 100 steps
data, tw = S.make_dataset(arch, n, 17)
w = S.student_weights(arch, 18, tw)
perms = [np.random.default_rng(e).permutation(n) for e in range(epochs)]
model = S.build_model(arch, w, precision=os.environ.get("DINO_TEST_PRECISION", "fp32x3"))
model.compile(optimizer=Adam(1e-3), loss=[normalized_mse, normalized_mse_matrix], loss_weights=[1.0, 1.0],
              metrics=[l2_accuracy, f_accuracy_matrix])
hist = model.fit([data["m"]], [data["q"], data["J"]], epochs=epochs, batch_size=B, perms=perms)
w_fit = [x for l in model.layers for x in l.get_weights()]
# the streaming path with HOST minibatches (fit_generator): every rank passes its shard of each global minibatch; the
# gradient and the step's statistics travel in ONE all-reduce (statistics at the tail of the gradient buffer)
def shards():
    for k in range(12):
        lo = (k * B) % (n - B)
        sl = slice(lo + dp.rank * (B // dp.world), lo + (dp.rank + 1) * (B // dp.world))
        yield [data["m"][sl]], [data["q"][sl], data["J"][sl]]
rows = model.fit_generator(shards())
if dp.rank == 0:
    np.savez(os.environ["DINO_OUT"], loss=np.array(hist.history["loss"]), fg_loss=np.array([r[0] for r in rows]),
             **{f"w{i}": a for i, a in enumerate(w_fit)},
             **{f"g{i}": a for i, a in enumerate(x for l in model.layers for x in l.get_weights())})
'''


def _run(world, out, precision):
    env = dict(os.environ, DINO_ROOT=ROOT, DINO_OUT=out, DINO_TEST_PRECISION=precision)
    if world == 1:
        cmd = [sys.executable, "-c", WORKER]
    else:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
               "--master-addr", "127.0.0.1", "--master-port", "29577", "--no-python", sys.executable, "-c", WORKER]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-3000:]
    return np.load(out)


@pytest.mark.parametrize("precision", ["fp32x3", "fp32"])
def test_two_ranks_match_one_rank_after_100_steps(cuda_device, tmp_path, precision):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs in one box")
    one = _run(1, str(tmp_path / "one.npz"), precision)
    two = _run(2, str(tmp_path / "two.npz"), precision)
    assert np.allclose(one["loss"], two["loss"], rtol=1e-6 if precision == "fp32" else 2e-6, atol=0)
    assert np.allclose(one["fg_loss"], two["fg_loss"], rtol=1e-5, atol=0)
    for k in one.files:
        if k.startswith("w") or k.startswith("g"):
            a, b = one[k], two[k]
            assert np.abs(a - b).max() <= 1e-5 * max(np.abs(a).max(), 1e-30), k
