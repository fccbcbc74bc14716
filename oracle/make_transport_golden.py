"""TEST INFRASTRUCTURE ONLY -- goldens for Transport.training_losses / ICPlan.plan / GVPCPlan.plan
(src/mdgen/transport/transport.py:125-222, path.py:112-190) from the UNMODIFIED reference.

The model inside training_losses is a cheap analytic stand-in (same callable on both sides), so the fixture pins the
interpolant, the noise / time draws (torch.manual_seed, CPU generator), the masked loss and the design branch (Dirichlet
noise on the sequence channels + cross-entropy).  Run:  python -m oracle.make_transport_golden  ->  tests/golden/transport.pt"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mdgen_finetune_b200 import configs as cfgs  # noqa: E402
from oracle.ref_loader import load_reference  # noqa: E402

CASES = {"gvp_sim": (1, "GVP"), "linear_sim": (1, "Linear"), "gvp_design": (4, "GVP")}


def standin_model(xt, t, **kw):
    return torch.tanh(xt) * (1 + t.view(-1, 1, 1, 1)) + 0.1 * kw.get("shift", 0.0)


def case_inputs(cfg_id):
    c = cfgs.small_variant(cfg_id, B=3, T=5, L=4)
    args = cfgs.make_args(c)
    g = torch.Generator().manual_seed(100 + cfg_id)
    Dx = c["D"] - 20 if args.design else c["D"]
    x1 = torch.randn(3, 5, 4, Dx, generator=g)
    mask = (torch.rand(3, 5, 4, Dx, generator=g) > 0.2).float()
    aatype1 = torch.randint(0, 20, (3, 4), generator=g)
    return args, x1, mask, aatype1


def main():
    ref = load_reference()
    out = {}
    for name, (cfg_id, path) in CASES.items():
        args, x1, mask, aatype1 = case_inputs(cfg_id)
        tr = ref.create_transport(args, path, "velocity")
        torch.manual_seed(11)
        terms = tr.training_losses(standin_model, x1, aatype1=aatype1 if args.design else None, mask=mask,
                                   model_kwargs={"shift": 0.5})
        out[name] = {k: (v.clone() if torch.is_tensor(v) else v) for k, v in terms.items()}
        print(name, {k: (tuple(v.shape) if torch.is_tensor(v) else v) for k, v in terms.items()})
    path = os.path.join(ROOT, "tests", "golden", "transport.pt")
    torch.save(out, path)
    print("wrote", path, os.path.getsize(path))


if __name__ == "__main__":
    main()
