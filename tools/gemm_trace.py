#!/usr/bin/env python
"""Clock stamps of one CTA of the CTA-pair tcgen05 GEMM (gemm_tc2.cuh): MMA / epilogue / producer timeline per tile.
usage: gemm_trace.py M N K mode [cta]   (mode: 2 bf16 store, 3 GELU, 4 gated residual)"""
import ctypes, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mdgen_finetune_b200 import _lib
L = _lib.lib()
p = lambda t: ctypes.c_void_p(t.data_ptr())
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
M, N, K, mode = [int(v) for v in sys.argv[1:5]]
cta = int(sys.argv[5]) if len(sys.argv) > 5 else 40
A = torch.randn(M, K, device="cuda"); W = torch.randn(N, K, device="cuda") / K ** 0.5; b = torch.randn(N, device="cuda")
C = torch.zeros(M, N, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for _ in range(2):
    _lib.check(L.mdgen_test_gemm(M, N, K, p(A), p(W), p(b), p(C), mode, 256, st))
dbg = torch.zeros(256, dtype=torch.int64, device="cuda")
flush.zero_()
L.mdgen_debug_set_buffer(p(dbg), cta)
_lib.check(L.mdgen_test_gemm(M, N, K, p(A), p(W), p(b), p(C), mode, 256, st))
L.mdgen_debug_set_buffer(None, 0)
torch.cuda.synchronize()
d = dbg.cpu().view(16, 16).tolist()
t0 = min(v for r in d for v in r if v)
names = ["mma top", "acc free", "kb0", "kb1", "kb2", "kb3", "kb4", "kb5", "mma issued", "epi acc ready", "epi tmem read", "epi done",
         "prod kb0 slot free", "prod issued"]
print("tile " + " ".join(f"{n[:9]:>9s}" for n in names))
for i, r in enumerate(d):
    if not any(r):
        continue
    print(f"{i:4d} " + " ".join(f"{(v - t0) if v else -1:9d}" for v in r[:14]))
