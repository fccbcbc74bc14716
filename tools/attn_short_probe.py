#!/usr/bin/env python
"""Times the bf16 residue-axis attention kernels alone on the config-3 shape (nb = 8000 sequences of S = 4 residues)
through mdgen_test_attention(use_tc=2).  MDGEN_ASB_SEQ / MDGEN_NO_ATTN_BULK select the variant."""
import ctypes, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mdgen_finetune_b200 import _lib

nb, S, inner = [int(v) for v in (sys.argv[1:4] if len(sys.argv) > 3 else (8000, 4, 1))]
H, hd = 16, 24
C = H * hd
L = _lib.lib()
g = torch.Generator().manual_seed(0)
qkv = (torch.randn(nb * S, 3 * C, generator=g) * 0.7).cuda()
bk, bv = torch.randn(C, generator=g).cuda(), torch.randn(C, generator=g).cuda()
valid = torch.ones(nb * S).cuda()
out = torch.empty(nb * S, C, device="cuda")
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
p = lambda t: ctypes.c_void_p(t.data_ptr())
def run():
    _lib.check(L.mdgen_test_attention(nb, S, inner, H, hd, p(qkv), p(bk), p(bv), p(valid), p(out), 2, 1, st))
for _ in range(3):
    run()
_lib.profile_begin()
for _ in range(10):
    run()
prof = _lib.profile_end()
print({k: (round(v[0] / v[1] * 1e3, 1), v[1]) for k, v in prof.items()}, "us/launch")
