"""Data-parallel plumbing: one process per GPU, torch.distributed (NCCL over NVLink on the GPU
box, gloo in the CPU tests).  The reference trains in a single process; sharding the minibatch
is the only parallel axis the path has: samples are independent, parameters and
optimiser state are replicated, and the one exchange per step is a SUM all-reduce of the flat
fp32 gradient, which the kernels pre-scale by 1/B_global so that the sum is the global mean.
"""
from __future__ import annotations

import os
from dataclasses import dataclass

import numpy as np


def shard_bounds(batch: int, rank: int, world: int):
    """Contiguous split of a global minibatch of ``batch`` samples: returns (offset, count) of
    ``rank``.  Counts differ by at most one; the first ``batch % world`` ranks get the extra."""
    base, rem = divmod(batch, world)
    count = base + (1 if rank < rem else 0)
    offset = rank * base + min(rank, rem)
    return offset, count


@dataclass
class DPContext:
    rank: int = 0
    world: int = 1

    def all_reduce_sum(self, tensor):
        import torch.distributed as dist
        dist.all_reduce(tensor, op=dist.ReduceOp.SUM)
        return tensor

    def broadcast_order(self, order: np.ndarray) -> np.ndarray:
        """Every rank must walk the same permutation: rank 0's draw wins."""
        import torch
        import torch.distributed as dist
        t = torch.from_numpy(np.ascontiguousarray(order, dtype=np.int64))
        if dist.get_backend() == "nccl":
            t = t.cuda()
        dist.broadcast(t, src=0)
        return t.cpu().numpy()

    def max_over_ranks(self, value: float) -> float:
        if self.world == 1:
            return value
        import torch
        import torch.distributed as dist
        t = torch.tensor([value], dtype=torch.float64)
        if dist.get_backend() == "nccl":
            t = t.cuda()
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def barrier(self):
        if self.world > 1:
            import torch.distributed as dist
            dist.barrier()


def context() -> DPContext:
    """The current data-parallel context (world 1 unless torch.distributed is initialised)."""
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            return DPContext(dist.get_rank(), dist.get_world_size())
    except ImportError:
        pass
    return DPContext()


def init_from_env(backend: str = "nccl") -> DPContext:
    """Initialise torch.distributed from RANK / WORLD_SIZE / LOCAL_RANK / MASTER_* (torchrun)."""
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world == 1:
        return DPContext()
    rank = int(os.environ["RANK"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29511")
    if backend == "nccl":
        torch.cuda.set_device(local)
    if not dist.is_initialized():
        kw = {}
        if backend == "nccl":
            kw["device_id"] = torch.device("cuda", local)
        dist.init_process_group(backend=backend, rank=rank, world_size=world, **kw)
    return DPContext(rank, world)
