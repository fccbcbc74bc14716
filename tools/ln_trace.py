#!/usr/bin/env python
"""Clock stamps of one leader CTA of gemm_ln_kernel (projection + residual + LayerNorm), per tile.
usage: ln_trace.py [M] [K] [cta]"""
import ctypes, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mdgen_finetune_b200 import _lib
L = _lib.lib()
p = lambda t: ctypes.c_void_p(t.data_ptr())
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
M = int(sys.argv[1]) if len(sys.argv) > 1 else 32000
K = int(sys.argv[2]) if len(sys.argv) > 2 else 384
cta = int(sys.argv[3]) if len(sys.argv) > 3 else 40
N = 384
A = torch.randn(M, K, device="cuda"); W = torch.randn(N, K, device="cuda") / K ** 0.5
v = torch.randn(7 * N, device="cuda") * 0.3
X = torch.zeros(2 * M, N, device="cuda")
for _ in range(2):
    _lib.check(L.mdgen_test_gemm(M, N, K, p(A), p(W), p(v), p(X), 5, 0, st))
dbg = torch.zeros(1024, dtype=torch.int64, device="cuda")
L.mdgen_debug_set_buffer(p(dbg), cta)
_lib.profile_begin()
_lib.check(L.mdgen_test_gemm(M, N, K, p(A), p(W), p(v), p(X), 5, 0, st))
prof = _lib.profile_end()
L.mdgen_debug_set_buffer(None, 0)
torch.cuda.synchronize()
d = dbg.cpu()[:256].view(16, 16).tolist()
t0 = min(x for r in d for x in r[:7] if x)
print("kernel time (profile scope): %.1f us" % (prof["gemm_tcgen05"][0] * 1e3))
print("tile  mma start  mma issued  epi start  acc seen  pass1 done  stats done  pass3 done   (clk)")
for i, r in enumerate(d):
    if any(r[:7]):
        print("%4d " % i + " ".join("%10d" % ((x - t0) if x else -1) for x in r[:7]))
