#!/usr/bin/env python
"""Times the tcgen05 attention kernel alone on the config-3 frame-axis shape (nb=32 sequences, S=1000,
inner=4) through mdgen_test_attention and prints the per-role clock64 stamps of one CTA."""
import ctypes, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mdgen_finetune_b200 import _lib

nb, S, inner = [int(v) for v in (sys.argv[1:4] if len(sys.argv) > 3 else (32, 1000, 4))]
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
dbg = torch.zeros(256, dtype=torch.int64, device="cuda")
cta = int(os.environ.get("DBG_CTA", "5"))
def run():
    _lib.check(L.mdgen_test_attention(nb, S, inner, H, hd, p(qkv), p(bk), p(bv), p(valid), p(out), 1, 1, st))
for _ in range(3):
    run()
_lib.profile_begin()
for _ in range(5):
    run()
prof = _lib.profile_end()
print({k: (round(v[0] / v[1] * 1e3, 1), v[1]) for k, v in prof.items()}, "us/launch")
if os.environ.get("NOSTAMP"): sys.exit(0)
L.mdgen_debug_set_buffer(p(dbg), cta)
run()
L.mdgen_debug_set_buffer(None, 0)
torch.cuda.synchronize()
d = dbg.cpu().tolist()
t0 = d[0]
names = {0: "entry", 1: "init done", 2: "t0: V cp.async issued", 3: "t0: K staged", 4: "t0: Q0 staged", 5: "t0: masks", 6: "t0: cp.async landed", 7: "prologue barrier", 13: "all Q staged", 14: "mma first", 15: "mma last"}
for w in range(2):
    for i in range(16):
        names[16 + w * 40 + 2 * i] = f"wg{w} blk{i} S ready"
        names[17 + w * 40 + 2 * i] = f"wg{w} blk{i} P done"
    names[48 + w * 40] = f"wg{w} tile0 epilogue"
names.update({88: "  blk4: A loaded", 89: "  blk4: chunk A done", 90: "  blk4: B loaded", 91: "  blk4: s_empty+prefetch issued", 92: "  blk4: chunk B done", 93: "  blk4: pv waited", 94: "  blk4: P stored"})
for i in range(48):
    names[96 + 2 * i] = f"    producer batch {i} start"
    names[97 + 2 * i] = f"    producer batch {i} published"
ev = sorted((d[i] - t0, names.get(i, str(i))) for i in range(256) if d[i])
for t, n in ev:
    print(f"{t:8d} clk  {t / 1.965e3:8.2f} us  {n}")
