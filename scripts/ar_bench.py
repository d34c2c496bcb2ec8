"""All-reduce of the cfg5 gradient (5.6 M floats) over the ranks of one box: time per call with CUDA events, max over ranks.
usage: torchrun --nproc-per-node N scripts/ar_bench.py   (NCCL_* knobs from the environment)"""
import os, torch, torch.distributed as dist
dist.init_process_group("nccl")
r = int(os.environ["LOCAL_RANK"]); torch.cuda.set_device(r)
for n in (22_448, 5_600_000):
    x = torch.ones(n, device="cuda")
    for _ in range(10):
        dist.all_reduce(x)
    torch.cuda.synchronize()
    # eager
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dist.barrier(); torch.cuda.synchronize()
    e0.record()
    for _ in range(50):
        dist.all_reduce(x)
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / 50], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    # graph replay
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        with torch.cuda.graph(g, stream=s):
            for _ in range(10):
                dist.all_reduce(x)
    torch.cuda.synchronize(); dist.barrier()
    g.replay(); torch.cuda.synchronize()
    e0.record()
    for _ in range(5):
        g.replay()
    e1.record(); torch.cuda.synchronize()
    t2 = torch.tensor([e0.elapsed_time(e1) / 50], device="cuda")
    dist.all_reduce(t2, op=dist.ReduceOp.MAX)
    if r == 0:
        print(f"n={n} ({4*n/1e6:.2f} MB) eager {1e3*t.item():.1f} us  graph {1e3*t2.item():.1f} us  env: " +
              " ".join(f"{k}={v}" for k, v in os.environ.items() if k.startswith("NCCL_")), flush=True)
dist.barrier()
