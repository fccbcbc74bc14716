#!/usr/bin/env python
"""Does alternating kernels with different shared-memory carve-outs cost time?  Times the same tcgen05 GEMM launched
back to back vs interleaved with a tiny elementwise kernel (0 B shared) vs a flush of L2."""
import ctypes, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mdgen_finetune_b200 import _lib
L = _lib.lib()
p = lambda t: ctypes.c_void_p(t.data_ptr())
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
M, N, K = 32000, 1152, 384
A = torch.randn(M, K, device="cuda"); W = torch.randn(N, K, device="cuda") / K ** 0.5; b = torch.randn(N, device="cuda")
C = torch.zeros(M, N, device="cuda")
tiny = torch.zeros(1024, device="cuda")
big = torch.zeros(64 << 20, device="cuda")
def gemm(bn): _lib.check(L.mdgen_test_gemm(M, N, K, p(A), p(W), p(b), p(C), 2, bn, st))
for bn in (128, 256):
    for mode in ("back-to-back", "tiny kernel between", "big elementwise between"):
        for _ in range(3): gemm(bn)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 20
        # time the interleaved kernels alone first
        e0.record()
        for _ in range(n):
            if mode == "tiny kernel between": tiny.add_(1.0)
            elif mode == "big elementwise between": big.add_(1.0)
        e1.record(); torch.cuda.synchronize()
        other = e0.elapsed_time(e1) * 1e3 / n
        e0.record()
        for _ in range(n):
            gemm(bn)
            if mode == "tiny kernel between": tiny.add_(1.0)
            elif mode == "big elementwise between": big.add_(1.0)
        e1.record(); torch.cuda.synchronize()
        tot = e0.elapsed_time(e1) * 1e3 / n
        print(f"bn={bn} {mode:26s}: {tot:7.1f} us per iteration, other kernel alone {other:6.1f} us -> gemm {tot - other:6.1f} us", flush=True)
