"""Data-parallel parity: N ranks on a global minibatch must end where one rank on the same global
minibatch ends.  Run once plainly and once under torchrun; each run writes its final parameters, the second run
compares:

    python scripts/dp_parity.py --out gpurun_out/dp_w1.npz
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 scripts/dp_parity.py \
        --out gpurun_out/dp_w2.npz --compare gpurun_out/dp_w1.npz
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch

import helpers as H
from dino_b200 import Adam, parallel, synthetic as S
from dino_b200.construction import f_accuracy_matrix, l2_accuracy, normalized_mse, normalized_mse_matrix

ap = argparse.ArgumentParser()
ap.add_argument("--out", required=True)
ap.add_argument("--compare", default=None)
ap.add_argument("--precision", default="fp32")
ap.add_argument("--epochs", type=int, default=5)
args = ap.parse_args()

dp = parallel.init_from_env("nccl")
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
arch = S.CONFIGS["cfg2"]["arch"]
n, Bg = 1280, 256                                        # 5 steps per epoch -> 25 optimiser steps
data, tw = S.make_dataset(arch, n, 41)
w = H.student_weights(arch, 42, tw, bias_noise=0.1)
model = S.build_model(arch, w, precision=args.precision)
model.compile(optimizer=Adam(1e-3), loss=[normalized_mse, normalized_mse_matrix], loss_weights=[1.0, 1.0],
              metrics=[l2_accuracy, f_accuracy_matrix])
perms = [np.random.default_rng(100 + e).permutation(n) for e in range(args.epochs)]
hist = model.fit([data["m"]], [data["q"], data["J"]], epochs=args.epochs, batch_size=Bg, perms=perms)
torch.cuda.synchronize()
if dp.rank == 0:
    flat = {f"{layer.name}/{i}": a for layer in model.layers for i, a in enumerate(layer.get_weights())}
    flat["loss"] = np.array(hist.history["loss"])
    np.savez(args.out, **flat)
    print(f"world {dp.world}: loss per epoch {hist.history['loss']}")
    if args.compare:
        ref = np.load(args.compare)
        worst = 0.0
        for k in flat:
            if k == "loss":
                continue
            worst = max(worst, H.rel_err(flat[k], ref[k]))
        dl = np.abs(flat["loss"] - ref["loss"]).max() / np.abs(ref["loss"]).max()
        print(f"world {dp.world} vs reference run: max rel parameter difference {worst:.3e}, max rel loss difference {dl:.3e}")
        tol_w, tol_l = (1e-5, 1e-6) if args.precision == "fp32" else (1e-3, 1e-4)
        assert worst < tol_w and dl < tol_l, "data-parallel run diverges from the single-rank run"
        print("dp parity ok")
if dp.world > 1:
    torch.distributed.barrier()
    torch.distributed.destroy_process_group()
