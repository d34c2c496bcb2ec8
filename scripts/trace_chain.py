"""Per-phase timing of the fused chain kernel (CTA 0).  Build the library with
   make -C dino_b200/csrc clean all EXTRA=-DDINO_CF_TRACE   first."""
import ctypes as C, os, sys
os.environ["DINO_B200_NO_GRAPH"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import helpers as H
from dino_b200 import synthetic as S, Adam, _lib
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
buf = (C.c_uint64 * 512)()
_lib.check(_lib.load().dino_debug_chain_trace(buf), "trace")
t = np.array(buf[:], dtype=np.int64).reshape(64, 8)
t0 = t[0, 0]
names = ["x_ready", "w_full", "commit", "w_issued", "mma_done", "epi_end", "bar_end"]
print("step  " + "  ".join(f"{n:>9s}" for n in names) + "   (ns since first stamp)")
for s in range(17):
    print(f"{s:4d}  " + "  ".join(f"{int(v - t0):9d}" for v in t[s, :7]))

print("kernel entry -> set-up done -> first x_ready -> loop end -> exit (ns):", [int(t[k, 7] - t[0, 7]) for k in range(4)], "first x_ready", int(t[0, 0] - t[0, 7]))
