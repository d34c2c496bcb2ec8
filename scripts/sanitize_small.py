"""Small end-to-end exercise of every kernel family, for compute-sanitizer (memcheck / racecheck)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import helpers as H
from dino_b200 import synthetic as S, Adam
from dino_b200.construction import (normalized_mse, normalized_mse_matrix, l2_accuracy, f_accuracy_matrix,
                                    equip_model_with_sketched_jacobian)
for precision in ("fp32", "tf32"):
    for name in ("dense_frozen", "resnet"):
        arch = H.SMALL[name]
        data, tw = S.make_dataset(arch, 21, 3)
        model = S.build_model(arch, H.student_weights(arch, 4, tw, bias_noise=0.1), precision=precision)
        model.compile(optimizer=Adam(1e-3), loss=[normalized_mse, normalized_mse_matrix], loss_weights=[1.0, 1.0],
                      metrics=[l2_accuracy, f_accuracy_matrix])
        model.fit([data["m"]], [data["q"], data["J"]], epochs=1, batch_size=8, perms=[np.arange(21)])
        model.evaluate([data["m"]], [data["q"], data["J"]], batch_size=16)
        model.predict(data["m"], batch_size=16)
        torch.cuda.synchronize()
        print(precision, name, "ok", flush=True)
arch = H.SMALL["generic"]
data, tw = S.make_dataset(arch, 10, 5)
rng = np.random.default_rng(0)
U = rng.normal(size=(10, arch.d_out, 3)).astype(np.float32); V = rng.normal(size=(10, arch.d_in, 3)).astype(np.float32)
sig = np.zeros((10, 3, 3), np.float32); sig[:, range(3), range(3)] = 1.0
sk = equip_model_with_sketched_jacobian(S.build_model(arch, H.student_weights(arch, 6, tw), jac=None), 3)
sk.compile(optimizer=Adam(1e-3), loss=[normalized_mse, normalized_mse_matrix], loss_weights=[1.0, 1.0],
           metrics=[l2_accuracy, f_accuracy_matrix])
sk.fit([data["m"], U, V], [data["q"], sig], epochs=1, batch_size=4, perms=[np.arange(10)])
torch.cuda.synchronize()
print("sketched ok")
