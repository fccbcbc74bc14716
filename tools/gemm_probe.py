#!/usr/bin/env python
"""Times the tcgen05 GEMM alone (mdgen_test_gemm) over the config-3 projection shapes and epilogue variants.
use_tc modes: 1 fp32 store, 2 bf16 store, 3 bf16 GELU, 4 gated residual RMW."""
import ctypes, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mdgen_finetune_b200 import _lib

L = _lib.lib()
p = lambda t: ctypes.c_void_p(t.data_ptr())
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
M = int(sys.argv[1]) if len(sys.argv) > 1 else 32000
shapes = [(1152, 384), (384, 384), (1536, 384), (384, 1536)]
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for N, K in shapes:
    A = torch.randn(M, K, device="cuda"); W = torch.randn(N, K, device="cuda") / K ** 0.5; b = torch.randn(N, device="cuda")
    C = torch.zeros(M, N, device="cuda")
    for mode in (2, 3, 4):
        for bn in (128, 256):
            for _ in range(2):
                _lib.check(L.mdgen_test_gemm(M, N, K, p(A), p(W), p(b), p(C), mode, bn, st))
            ts = []
            for _ in range(5):
                flush.sum()          # read-only L2 flush: leaves clean lines behind, no dirty write-backs billed to the GEMM
                _lib.profile_begin()
                _lib.check(L.mdgen_test_gemm(M, N, K, p(A), p(W), p(b), p(C), mode, bn, st))
                pr = _lib.profile_end()
                ts.append(pr["gemm_tcgen05"][0] * 1e3)
            ts.sort()
            t = ts[len(ts) // 2]
            print(f"M={M} N={N:5d} K={K:5d} mode={mode} bn={bn:3d} : {t:7.1f} us  {2.0 * M * N * K / t / 1e6:7.1f} TFLOP/s", flush=True)
