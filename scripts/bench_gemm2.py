"""Micro-benchmark of the tcgen05 engine in both arithmetic modes next to cuBLAS TF32 (torch.matmul): the chain
contraction C = A Bt^T (dino_gemm_nt) and the weight-gradient reduction C += X^T Y (dino_gemm_tn_acc).  CUDA-event timed.
DINO_TC_VARIANT selects the tile variant of the fp32x3 contraction (read once per process)."""
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


var = os.environ.get("DINO_TC_VARIANT", "0")
for (M, N, K) in [(25600, 1024, 1024), (25600, 400, 400), (102400, 1024, 1024), (25600, 200, 1024), (25600, 1024, 200)]:
    A = torch.randn(M, K, device="cuda"); Bt = torch.randn(N, K, device="cuda"); Cm = torch.empty(M, N, device="cuda")
    fl = 2.0 * M * N * K
    out = {"op": "nt", "variant": var, "M": M, "N": N, "K": K}
    for name, eng in (("tf32", _lib.PREC_TF32), ("fp32x3", _lib.PREC_FP32X3)):
        ms = time_ms(lambda: _lib.check(lib.dino_gemm_nt(eng, C.c_void_p(A.data_ptr()), K, C.c_void_p(Bt.data_ptr()), K,
                                                         C.c_void_p(Cm.data_ptr()), N, M, N, K, _stream()), "gemm"), 20)
        out[name + "_us"] = round(ms * 1e3, 1); out[name + "_tflops"] = round(fl / ms / 1e9, 1)
    ms = time_ms(lambda: torch.matmul(A, Bt.t()), 20)
    out["cublas_tf32_us"] = round(ms * 1e3, 1); out["cublas_tf32_tflops"] = round(fl / ms / 1e9, 1)
    print(json.dumps(out), flush=True)
    del A, Bt, Cm
for (rows, n_in, n_out) in [(25600, 1024, 1024), (25600, 400, 400), (25600, 100, 52)]:
    X = torch.randn(rows, n_in, device="cuda"); Y = torch.randn(rows, n_out, device="cuda"); G = torch.zeros(n_in, n_out, device="cuda")
    fl = 2.0 * rows * n_in * n_out
    out = {"op": "tn_acc", "rows": rows, "n_in": n_in, "n_out": n_out}
    for name, eng in (("tf32", _lib.PREC_TF32), ("fp32x3", _lib.PREC_FP32X3)):
        ms = time_ms(lambda: _lib.check(lib.dino_gemm_tn_acc(eng, C.c_void_p(X.data_ptr()), n_in, C.c_void_p(Y.data_ptr()), n_out,
                                                             C.c_void_p(G.data_ptr()), rows, n_in, n_out, _stream()), "wgrad"), 20)
        out[name + "_us"] = round(ms * 1e3, 1); out[name + "_tflops"] = round(fl / ms / 1e9, 1)
    ms = time_ms(lambda: torch.matmul(X.t(), Y), 20)
    out["cublas_tf32_us"] = round(ms * 1e3, 1); out["cublas_tf32_tflops"] = round(fl / ms / 1e9, 1)
    print(json.dumps(out), flush=True)
