"""A few cfg2 training steps in tf32 mode without graph replay, for ncu on chain_fused_kernel."""
import os, sys
os.environ["DINO_B200_NO_GRAPH"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import helpers as H
from dino_b200 import synthetic as S, Adam
from dino_b200.construction import normalized_mse, normalized_mse_matrix, l2_accuracy, f_accuracy_matrix
arch = S.CONFIGS["cfg2"]["arch"]
data, tw = S.make_dataset(arch, 1024, 1)
model = S.build_model(arch, H.student_weights(arch, 2, tw), precision="tf32")
model.compile(optimizer=Adam(1e-3), loss=[normalized_mse, normalized_mse_matrix], loss_weights=[1.0, 1.0],
              metrics=[l2_accuracy, f_accuracy_matrix])
ds = model.make_dataset([data["m"]], [data["q"], data["J"]])
eng = model.make_engine(ds, 256, use_graph=False)
eng.set_order(np.arange(1024, dtype=np.int32))
for i in range(4):
    eng.step(256)
torch.cuda.synchronize()
print("ok")
