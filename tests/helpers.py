"""Shared test helpers: oracle <-> product conversions and seeded problems."""
import numpy as np
import torch

from oracle import dino_oracle as O
from dino_b200 import synthetic as S


def oracle_net(arch, weights, dtype=torch.float64):
    net = O.ONet(arch.d_in)
    for (name, k, n, sp, bias, skip) in S.layer_table(arch):
        frozen = (name == "output_layer" and arch.frozen_output)
        net.layers.append(O.OLayer(name, torch.tensor(weights[name][0], dtype=dtype),
                                   torch.tensor(weights[name][1], dtype=dtype) if bias else None,
                                   O.ACT_SOFTPLUS if sp else O.ACT_LINEAR, not frozen, skip))
    return net


def oracle_batch(data, idx=None, dtype=torch.float64, U=None, V=None):
    sel = (lambda a: a) if idx is None else (lambda a: a[idx])
    t = lambda a: None if a is None else torch.tensor(sel(np.asarray(a)), dtype=dtype)
    return O.Batch(t(data["m"]), t(data["q"]), t(data["J"]), t(U), t(V))


student_weights = S.student_weights


def rel_err(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / (np.abs(b).max() + 1e-30))


SMALL = {
    "dense": S.Arch("rb_dense", 12, 7, hidden=[9, 9, 7], frozen_output=False, r_out=7),
    "dense_frozen": S.Arch("rb_dense", 13, 10, hidden=[8, 6], frozen_output=True, r_out=6),
    "resnet": S.Arch("rb_resnet", 11, 9, ranks=[4, 5, 3], frozen_output=True, r_out=6),
    "resnet_square": S.Arch("rb_resnet", 6, 6, ranks=[8, 8], frozen_output=False, r_out=6),
    "generic": S.Arch("generic_dense", 37, 5, hidden=[10, 10]),
}
