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
dbg = torch.zeros(1024, dtype=torch.int64, device="cuda")
flush.zero_()
L.mdgen_debug_set_buffer(p(dbg), cta)
_lib.profile_begin()
_lib.check(L.mdgen_test_gemm(M, N, K, p(A), p(W), p(b), p(C), mode, 256, st))
prof = _lib.profile_end()
L.mdgen_debug_set_buffer(None, 0)
torch.cuda.synchronize()
allc = dbg.cpu()[256:256 + 2 * 148].view(148, 2)
ent, ext = allc[:, 0], allc[:, 1]
g0 = int(ent[ent > 0].min())
print("CTA entry  (us after first): min %.1f median %.1f max %.1f" % (0.0, float((ent[ent > 0] - g0).double().median()) / 1e3, float((ent[ent > 0] - g0).max()) / 1e3))
print("CTA exit   (us after first entry): min %.1f median %.1f max %.1f" % (float((ext[ext > 0] - g0).min()) / 1e3, float((ext[ext > 0] - g0).double().median()) / 1e3, float((ext[ext > 0] - g0).max()) / 1e3))
d = dbg.cpu()[:256].view(16, 16).tolist()
ns = d[15][4] - d[15][3]
d[15][3] = d[15][4] = 0
t0 = min(v for r in d for v in r if v)
print(f"in-kernel wall time {ns / 1e3:.1f} us for {d[15][2] - d[15][0]} clk -> SM clock {(d[15][2] - d[15][0]) / ns * 1e3:.0f} MHz")
names = ["mma top", "acc free", "kb0", "kb1", "kb2", "kb3", "kb4", "kb5", "mma issued", "epi acc ready", "epi tmem read", "epi done",
         "prod kb0 slot free", "prod issued"]
print("tile " + " ".join(f"{n[:9]:>9s}" for n in names))
print(f"entry {d[15][0] - t0}, setup done {d[15][1] - t0}, exit {d[15][2] - t0} clk; profile-scope kernel time:", end=" ")
print(f"{prof['gemm_tcgen05'][0] * 1e3:.1f} us")
d[15] = [0] * 16
for i, r in enumerate(d):
    if not any(r):
        continue
    print(f"{i:4d} " + " ".join(f"{(v - t0) if v else -1:9d}" for v in r[:14]))
