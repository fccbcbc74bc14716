#!/usr/bin/env python
"""Clock stamps of one CTA of the tcgen05 GEMM: producer / MMA / epilogue timeline per tile."""
import ctypes, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mdgen_finetune_b200 import _lib
L = _lib.lib()
p = lambda t: ctypes.c_void_p(t.data_ptr())
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
M, N, K, mode = [int(v) for v in sys.argv[1:5]]
A = torch.randn(M, K, device="cuda"); W = torch.randn(N, K, device="cuda") / K ** 0.5; b = torch.randn(N, device="cuda")
C = torch.zeros(M, N, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for _ in range(2):
    _lib.check(L.mdgen_test_gemm(M, N, K, p(A), p(W), p(b), p(C), mode, 0, st))
dbg = torch.zeros(96, dtype=torch.int64, device="cuda")
flush.zero_()
L.mdgen_debug_set_buffer(p(dbg), int(os.environ.get("DBG_CTA", "70")))
_lib.check(L.mdgen_test_gemm(M, N, K, p(A), p(W), p(b), p(C), mode, 0, st))
L.mdgen_debug_set_buffer(None, 0)
torch.cuda.synchronize()
d = dbg.cpu().tolist()
t0 = min(v for v in d if v)
names = {0: "prod start", 8: "mma start", 16: "mma issued", 24: "epi top", 32: "epi acc ready", 40: "epi tmem read", 48: "epi done"}
ev = []
for t, n in ev:
    print(f"{t:8d} clk  {n}")
