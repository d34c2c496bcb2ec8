"""Experiment: how does the tensor core accumulate?  C = A B^T with same-sign operands: a round-toward-zero accumulator
shows up as a systematic negative relative error that grows with the number of MMA accumulation steps."""
import ctypes as C
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
from dino_b200 import _lib
from dino_b200.model import _stream

lib = _lib.load()
rng = np.random.default_rng(0)
for signed in (False, True):
    for K in (64, 256, 1024, 4096, 16384):
        M, N = 256, 64
        A = rng.uniform(0.5, 1.5, size=(M, K)).astype(np.float32)
        B = rng.uniform(0.5, 1.5, size=(N, K)).astype(np.float32)
        if signed:
            A *= rng.choice([-1.0, 1.0], size=A.shape).astype(np.float32)
        want = A.astype(np.float64) @ B.astype(np.float64).T
        dA, dB = torch.from_numpy(A).cuda(), torch.from_numpy(B).cuda()
        row = []
        for eng in (_lib.PREC_FP32, _lib.PREC_TF32, _lib.PREC_FP32X3):
            out = torch.zeros(M, N, device="cuda")
            _lib.check(lib.dino_gemm_nt(eng, C.c_void_p(dA.data_ptr()), K, C.c_void_p(dB.data_ptr()), K,
                                        C.c_void_p(out.data_ptr()), N, M, N, K, _stream()), "gemm")
            torch.cuda.synchronize()
            got = out.cpu().numpy().astype(np.float64)
            scale = np.abs(want).max()
            row.append((float(((got - want) / scale).mean()), float(np.abs(got - want).max() / scale)))
        print(f"signed={signed} K={K:6d}  fp32 mean/max {row[0][0]:+.2e} {row[0][1]:.2e} | tf32 {row[1][0]:+.2e} {row[1][1]:.2e} | "
              f"fp32x3 {row[2][0]:+.2e} {row[2][1]:.2e}", flush=True)
