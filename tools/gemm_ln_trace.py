#!/usr/bin/env python
"""Clock stamps of one CTA of the fused projection + residual + LayerNorm kernel (gemm_ln.cuh)."""
import ctypes, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mdgen_finetune_b200 import _lib
L = _lib.lib()
p = lambda t: ctypes.c_void_p(t.data_ptr())
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
M, N = 32000, 384
K = int(sys.argv[1]) if len(sys.argv) > 1 else 384
cta = int(sys.argv[2]) if len(sys.argv) > 2 else 40
A = torch.randn(M, K, device="cuda"); W = torch.randn(N, K, device="cuda") / K ** 0.5; b = torch.randn(2 * N, device="cuda")
C = torch.zeros(2 * M, N, device="cuda")
for _ in range(2):
    _lib.check(L.mdgen_test_gemm(M, N, K, p(A), p(W), p(b), p(C), 5, 0, st))
dbg = torch.zeros(1024, dtype=torch.int64, device="cuda")
L.mdgen_debug_set_buffer(p(dbg), cta)
_lib.check(L.mdgen_test_gemm(M, N, K, p(A), p(W), p(b), p(C), 5, 0, st))
L.mdgen_debug_set_buffer(None, 0)
torch.cuda.synchronize()
d = dbg.cpu()[:256].view(16, 16).tolist()
t0 = min(v for r in d for v in r if v)
names = ["mma start", "mma issued", "epi top", "acc ready", "pass1 done", "pass2 done", "pass3 done"]
print("tile " + " ".join(f"{n:>11s}" for n in names))
for i, r in enumerate(d):
    if any(r):
        print(f"{i:4d} " + " ".join(f"{(v - t0) if v else -1:11d}" for v in r[:7]))
