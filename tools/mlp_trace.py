#!/usr/bin/env python
"""Clock stamps of the second tile of one (leader) CTA of the one-kernel MLP block (gemm_mlp.cuh).
usage: mlp_trace.py [M] [cta]"""
import ctypes, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mdgen_finetune_b200 import _lib
L = _lib.lib()
p = lambda t: ctypes.c_void_p(t.data_ptr())
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
M = int(sys.argv[1]) if len(sys.argv) > 1 else 32000
cta = int(sys.argv[2]) if len(sys.argv) > 2 else 40
C, F = 384, 1536
A = torch.randn(M, C, device="cuda")
W = torch.randn(2 * F * C, device="cuda") / 20
v = torch.randn(F + 7 * C, device="cuda") * 0.3
X = torch.zeros(2 * M, C, device="cuda")
for _ in range(2):
    _lib.check(L.mdgen_test_gemm(M, C, C, p(A), p(W), p(v), p(X), 6, 0, st))
dbg = torch.zeros(1024, dtype=torch.int64, device="cuda")
L.mdgen_debug_set_buffer(p(dbg), cta)
_lib.profile_begin()
_lib.check(L.mdgen_test_gemm(M, C, C, p(A), p(W), p(v), p(X), 6, 0, st))
prof = _lib.profile_end()
L.mdgen_debug_set_buffer(None, 0)
torch.cuda.synchronize()
d = dbg.cpu()[:192].view(12, 16).tolist()
t0 = d[4][0]
print("kernel time (profile scope): %.1f us" % (prof["gemm_tcgen05"][0] * 1e3))
print("h rows landed at 0; MMAs all issued %d; acc2 seen by epilogue %d; LayerNorm epilogue done %d clk" % (d[4][1] - t0, d[4][2] - t0, d[4][3] - t0))
print("  j  fc1 issue..end  fc2 issue..end (MMA thread)")
for j in range(12):
    print("%3d %10d %6d %10d %6d" % (j, d[0][j] - t0, d[10][j] - t0, d[1][j] - t0, d[11][j] - t0))
print("  j  fc1 issue   fc2 issue | GELU: acc1 seen  math done  stored+fenced  arrived  bias stored   (clk after the h rows landed)")
for j in range(12):
    print("%3d %10d %11d | %14d %8d %10d %7d %11d" % (j, d[0][j] - t0, d[1][j] - t0, d[2][j] - t0, d[7][j] - t0, d[5][j] - t0, d[6][j] - t0, d[3][j] - t0))
print("LayerNorm epilogue: start %d, acc2 seen %d, pass 1 done %d, pass 2 done %d, pass 3 done %d" % tuple(d[9][k] - t0 for k in (2, 3, 4, 5, 6)))
