"""Experiment: is the step host-launch-bound?  Compares eager enqueue time, eager GPU time and a
CUDA-graph replay of the same step."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import helpers as H
from dino_b200 import synthetic as S, Adam
from dino_b200.construction import normalized_mse, normalized_mse_matrix, l2_accuracy, f_accuracy_matrix

for B in (32, 256, 1024):
    arch = S.CONFIGS["cfg2"]["arch"]
    data, tw = S.make_dataset(arch, 4096, 1)
    w = H.student_weights(arch, 2, tw)
    model = S.build_model(arch, w)
    model.compile(optimizer=Adam(1e-3), loss=[normalized_mse, normalized_mse_matrix], loss_weights=[1.0, 1.0],
                  metrics=[l2_accuracy, f_accuracy_matrix])
    ds = model.make_dataset([data["m"]], [data["q"], data["J"]])
    idx = torch.arange(4096, dtype=torch.int32, device="cuda")
    for i in range(10):
        model.train_on_indices(ds, idx, 0, B)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(100):
        model.train_on_indices(ds, idx, 0, B)
    t_enq = time.perf_counter() - t0
    torch.cuda.synchronize()
    t_all = time.perf_counter() - t0
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        model.train_on_indices(ds, idx, 0, B)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            model.train_on_indices(ds, idx, 0, B)
        for i in range(5):
            g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for i in range(200):
            g.replay()
        e1.record(s)
        torch.cuda.synchronize()
    print(f"B={B}: eager enqueue {t_enq*10:.3f} ms/step, eager total {t_all*10:.3f} ms/step, "
          f"graph replay {e0.elapsed_time(e1)/200:.3f} ms/step", flush=True)
