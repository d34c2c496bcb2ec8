"""World-size-2 gloo tests of the data-parallel host logic (CPU): sharding by 1/B_global plus a SUM
all-reduce reproduces the single-process gradient; the permutation broadcast; max-over-ranks timing."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import pipeline_model as P
from dino_b200 import parallel


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _layers(rng):
    dims = [(6, 6, 1), (6, 5, 1), (5, 5, 0), (5, 4, 0)]
    out = []
    for i, (k, n, sp) in enumerate(dims):
        b = rng.normal(0, 0.1, n) if i != 2 else None
        out.append(P.Layer(rng.normal(0, 0.5, (k, n)), b, sp, trainable=(i != 3)))
    return out


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    dp = parallel.init_from_env("gloo")
    assert (dp.rank, dp.world) == (rank, world) == (parallel.context().rank, parallel.context().world)
    rng = np.random.default_rng(0)                      # same data and weights on every rank
    layers = _layers(rng)
    Bg = 7                                              # odd: shards of 4 and 3
    m, qq, J = rng.normal(size=(Bg, 6)), rng.normal(size=(Bg, 4)), rng.normal(size=(Bg, 4, 6))
    order = dp.broadcast_order(np.random.default_rng(100 + rank).permutation(Bg))   # rank 0's draw wins
    off, cnt = parallel.shard_bounds(Bg, rank, world)
    sel = order[off:off + cnt]
    out = P.step(layers, m[sel], qq[sel], J[sel], (1.0, 1.0), inv_B=1.0 / Bg)
    flat = np.concatenate([g.ravel() for g in out["dW"]] + [g.ravel() for g in out["db"] if g is not None]
                          + [np.array(out["loss3"])])
    t = torch.from_numpy(flat.copy())
    dp.all_reduce_sum(t)
    full = P.step(layers, m[order], qq[order], J[order], (1.0, 1.0))
    want = np.concatenate([g.ravel() for g in full["dW"]] + [g.ravel() for g in full["db"] if g is not None]
                          + [np.array(full["loss3"])])
    err = float(np.abs(t.numpy() - want).max() / np.abs(want).max())
    mx = dp.max_over_ranks(float(rank + 1))
    dp.barrier()
    q.put((rank, err, order.tolist(), mx))
    dist.destroy_process_group()


def test_sharded_gradient_sum_equals_full_batch_gradient():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res[0][2] == res[1][2] == np.random.default_rng(100).permutation(7).tolist()
    for rank, err, _, mx in res:
        assert err < 1e-12 and mx == 2.0
