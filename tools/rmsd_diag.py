import os, sys, torch
sys.path.insert(0, '/root/repo')
from tests.helpers import *
from tests.test_gpu_parity import build, gpu_kwargs
from mdgen_finetune_b200 import decode_samples
g, c, args, sd, inp = load_case("sim_full")
m = build(args, c["D"], sd, "bf16")
kw = gpu_kwargs(inp)
with torch.no_grad():
    out = m.euler_rollout(inp["x0"].cuda(), g["euler_steps"], **kw).cpu()
ref = g["euler_final"]
tables = torch.load('/root/repo/tests/golden/decode.pt')["tables"]
first = (inp["start_rot"].cuda(), inp["start_trans"].cuda()); seq = inp["aatype"].cuda()
a_out, _ = decode_samples(args, out.cuda(), first, seq, tables)
a_ref, _ = decode_samples(args, ref.cuda(), first, seq, tables)
msk = tables["atom_mask"][inp["aatype"]][:, None].expand(-1, out.shape[1], -1, -1)
d2 = ((a_out - a_ref).cpu().pow(2).sum(-1) * msk).sum((-1, -2)) / msk.sum((-1, -2))
r = d2.sqrt().flatten()
tn = ref[..., 7:21].reshape(1, -1, 4, 7, 2).norm(dim=-1)        # torsion pair norms [B,T,L,7]
mn = tn.amin((-1, -2)).flatten()
qn = ref[..., :4].norm(dim=-1).amin(-1).flatten()
print("frames", r.numel(), "median", float(r.median()), "p99", float(r.quantile(0.99)), "max", float(r.max()), "n>0.05", int((r > 0.05).sum()))
idx = torch.argsort(r, descending=True)[:8]
for i in idx: print(int(i), float(r[i]), "min torsion norm", float(mn[i]), "min quat norm", float(qn[i]))
print("torsion norm stats", float(tn.min()), float(tn.median()), "quat norm", float(qn.min()))
