#!/usr/bin/env python
"""Times the fused projection + residual + LayerNorm kernel (gemm_ln.cuh, mdgen_test_gemm mode 5) on the config-3 shapes."""
import ctypes, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mdgen_finetune_b200 import _lib
L = _lib.lib()
p = lambda t: ctypes.c_void_p(t.data_ptr())
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
M, N = 32000, 384
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for K in (384, 1536):
    A = torch.randn(M, K, device="cuda"); W = torch.randn(N, K, device="cuda") / K ** 0.5; b = torch.randn(2 * N, device="cuda")
    C = torch.zeros(2 * M, N, device="cuda")
    for _ in range(2):
        _lib.check(L.mdgen_test_gemm(M, N, K, p(A), p(W), p(b), p(C), 5, 0, st))
    ts = []
    for _ in range(5):
        if not os.environ.get('NO_FLUSH'): flush.sum()
        _lib.profile_begin()
        _lib.check(L.mdgen_test_gemm(M, N, K, p(A), p(W), p(b), p(C), 5, 0, st))
        ts.append(_lib.profile_end()["gemm_tcgen05"][0] * 1e3)
    ts.sort()
    print(f"fused resid+LN  M={M} N={N} K={K}: {ts[2]:7.1f} us", flush=True)
