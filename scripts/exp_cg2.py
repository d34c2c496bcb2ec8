"""Correctness + timing of the contraction engine (DINO_TC_CG2 selects the CTA-pair kernel) against fp64 on sampled rows."""
import ctypes as C, os, sys
sys.path.insert(0, ".")
import torch
from dino_b200 import _lib
from dino_b200.model import _stream
lib = _lib.load()
torch.manual_seed(0)
for (M, N, K) in [(25600, 1024, 1024), (20000, 1000, 520), (102400, 1024, 1024)]:
    ldk, ldn = (K + 3) & ~3, (N + 3) & ~3
    A = torch.randn(M, ldk, device="cuda"); Bt = torch.randn(N, ldk, device="cuda")
    rows = torch.randint(0, M, (64,), device="cuda")
    want = A[rows, :K].double() @ Bt[:, :K].double().t()
    for name, eng in (("tf32", _lib.PREC_TF32), ("fp32x3", _lib.PREC_FP32X3)):
        Cm = torch.full((M, ldn), float("nan"), device="cuda")
        def run():
            _lib.check(lib.dino_gemm_nt(eng, C.c_void_p(A.data_ptr()), ldk, C.c_void_p(Bt.data_ptr()), ldk,
                                        C.c_void_p(Cm.data_ptr()), ldn, M, N, K, _stream()), "gemm")
        run(); torch.cuda.synchronize()
        err = (Cm[rows, :N].double() - want).abs().max().item() / want.abs().max().item()
        fin = bool(torch.isfinite(Cm[:, :N]).all())
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for _ in range(3): run()
        e0.record()
        for _ in range(10): run()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        print(f"cg2={os.environ.get('DINO_TC_CG2','0')} {M}x{N}x{K} {name}: rel err {err:.2e} finite {fin} {ms*1e3:.1f} us {2.0*M*N*K/ms/1e9:.1f} TFLOP/s", flush=True)
