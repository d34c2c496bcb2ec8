"""Micro-benchmark of the chain contraction C = A Bt^T: SIMT fp32 engine vs tcgen05 TF32 engine vs
cuBLAS TF32 (torch.matmul, the measured TF32 peak of this GPU).  CUDA-event timed."""
import ctypes as C, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from dino_b200 import _lib
from dino_b200.model import _stream

lib = _lib.load()
torch.backends.cuda.matmul.allow_tf32 = True


def time_ms(fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


rows = []
shapes = [(12800, 50, 50), (12800, 100, 50), (12800, 50, 100), (25600, 100, 200), (25600, 400, 400),
          (25600, 1024, 1024), (102400, 1024, 1024), (8192, 8192, 8192)]
for (M, N, K) in shapes:
    r4 = lambda n: (n + 3) & ~3
    A = torch.randn(M, r4(K), device="cuda")
    Bt = torch.randn(N, r4(K), device="cuda")
    Cm = torch.empty(M, r4(N), device="cuda")
    fl = 2.0 * M * N * K
    iters = 20 if fl > 1e10 else 100
    out = {"M": M, "N": N, "K": K}
    for name, eng in (("simt_fp32", _lib.PREC_FP32), ("tcgen05_tf32", _lib.PREC_TF32)):
        if eng == _lib.PREC_FP32 and fl > 3e11:
            continue
        def run():
            _lib.check(lib.dino_gemm_nt(eng, C.c_void_p(A.data_ptr()), A.shape[1], C.c_void_p(Bt.data_ptr()), Bt.shape[1],
                                        C.c_void_p(Cm.data_ptr()), Cm.shape[1], M, N, K, _stream()), "gemm")
        ms = time_ms(run, iters)
        out[name + "_us"] = round(ms * 1e3, 2)
        out[name + "_tflops"] = round(fl / ms / 1e9, 2)
    Ad, Bd = A[:, :K].contiguous(), Bt[:, :K].contiguous()
    ms = time_ms(lambda: torch.matmul(Ad, Bd.t()), iters)
    out["cublas_tf32_us"] = round(ms * 1e3, 2)
    out["cublas_tf32_tflops"] = round(fl / ms / 1e9, 2)
    rows.append(out)
    print(json.dumps(out), flush=True)
