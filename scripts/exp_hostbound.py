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
    order = np.arange(4096, dtype=np.int32)
    for graph in (False, True):
        eng = model.make_engine(ds, B, use_graph=graph)
        eng.set_order(order)
        def run(k):
            for i in range(k):
                if eng.remaining() < B:
                    eng.rewind()
                eng.step(B)
        run(10)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        run(200)
        t_enq = time.perf_counter() - t0
        torch.cuda.synchronize()
        t_all = time.perf_counter() - t0
        print(f"B={B} graph={graph}: host enqueue {t_enq*5:.3f} ms/step, total {t_all*5:.3f} ms/step", flush=True)
