"""TEST INFRASTRUCTURE ONLY -- gradient goldens for row N4 from the UNMODIFIED reference's autograd.

For each case: loss = sum(LatentMDGenModel.forward(x, t, **kw) * R) with a seeded random cotangent R and a per-sample t
(the training / RTB call convention); the reference's autograd gives dLoss/dx and dLoss/dparam.  A full set of parameter
gradients is 137 MB per case, so the fixture holds dx in full, every parameter gradient with <= 4096 elements in full, and
for ALL parameters the norm and 8 seeded random projections (helpers in tests/helpers.py rebuild the same vectors).
Run in the build container:  python -m oracle.make_grad_golden   ->  tests/golden/grads.pt"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mdgen_finetune_b200 import configs as cfgs  # noqa: E402
from oracle.make_golden import CASES, build_reference_model, case_config, ref_kwargs  # noqa: E402
from oracle.ref_loader import load_reference  # noqa: E402

GRAD_CASES = ["sim_small", "tps_small", "design_small", "atlas_small"]
NPROJ, FULL_MAX = 8, 4096


def cotangent(shape):
    return torch.randn(shape, generator=torch.Generator().manual_seed(123))


def projections(name_index, g):
    """8 projections of a gradient on seeded N(0, 1) vectors (seed from the parameter's index in the state_dict)."""
    gen = torch.Generator().manual_seed(7919 * (name_index + 1))
    r = torch.randn(NPROJ, g.numel(), generator=gen)
    return r @ g.reshape(-1).double().float()


def summarize(names, grads):
    out = {}
    for i, (n, g) in enumerate(zip(names, grads)):
        if g is None:
            continue
        e = {"norm": float(g.double().norm()), "proj": projections(i, g), "index": i}
        if g.numel() <= FULL_MAX:
            e["full"] = g.clone()
        out[n] = e
    return out


def main():
    ref = load_reference()
    out = {}
    for name in GRAD_CASES:
        c, pad = case_config(name)
        model, args, sd = build_reference_model(ref, c)
        inp = cfgs.synthetic_inputs(c, seed=0, pad_last=pad)
        kw = ref_kwargs(ref, inp)
        B = c["B"]
        x = inp["x0"].clone().requires_grad_(True)
        t = torch.linspace(0.1, 0.9, B) if B > 1 else torch.full((1,), 0.3)
        y = model.forward(x, t, **kw)
        loss = (y * cotangent(y.shape)).sum()
        params = [(n, p) for n, p in model.named_parameters()]
        grads = torch.autograd.grad(loss, [x] + [p for _, p in params], allow_unused=True)
        out[name] = {"t": t, "y": y.detach().clone(), "dx": grads[0].clone(),
                     "params": summarize([n for n, _ in params], grads[1:]),
                     "unused": [n for (n, _), g in zip(params, grads[1:]) if g is None]}
        print(name, "loss", float(loss), "|dx|", float(grads[0].norm()), "params", len(out[name]["params"]), "unused", out[name]["unused"])
    path = os.path.join(ROOT, "tests", "golden", "grads.pt")
    torch.save(out, path)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
