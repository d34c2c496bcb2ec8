"""One tcgen05 chain-shaped GEMM (cfg5: [25600,1024] x [1024,1024]) for ncu."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from dino_b200 import _lib
from dino_b200.model import _stream
lib = _lib.load()
M, N, K = 25600, 1024, 1024
A = torch.randn(M, K, device="cuda"); Bt = torch.randn(N, K, device="cuda"); Cm = torch.empty(M, N, device="cuda")
for _ in range(6):
    _lib.check(lib.dino_gemm_nt((_lib.PREC_FP32X3 if __import__("os").environ.get("PROF_X3") else _lib.PREC_TF32), C.c_void_p(A.data_ptr()), K, C.c_void_p(Bt.data_ptr()), K,
                                C.c_void_p(Cm.data_ptr()), N, M, N, K, _stream()), "gemm")
torch.cuda.synchronize()
print("ok", float(Cm[0, 0]))
