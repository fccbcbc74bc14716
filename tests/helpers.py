"""Shared test helpers: load a golden case and rebuild its inputs / weights."""
import os

import torch

from mdgen_finetune_b200 import configs as cfgs
from mdgen_finetune_b200.weights import random_state_dict

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ["sim_small", "tps_small", "upsample_small", "design_small", "atlas_small", "atlas_mid"]
# BASELINE configs 2-5 at FULL size (oracle/make_golden.py FULL): the shapes that select the kernels bench.py times
FULL_CASES = ["tps_full", "upsample_full", "design_full", "atlas_full"]


def load_case(name, seed=0):
    g = torch.load(os.path.join(GOLD, name + ".pt"), weights_only=False)
    c = g["config"]
    args = cfgs.make_args(c)
    sd = random_state_dict(args, c["D"], seed=0)
    inp = cfgs.synthetic_inputs(c, seed=seed, pad_last=g["pad_last"])
    return g, c, args, sd, inp


def frames_of(inp):
    start = (inp["start_rot"], inp["start_trans"])
    end = (inp["end_rot"], inp["end_trans"]) if "end_rot" in inp else None
    return start, end


def rel_l2(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm())


def oracle_velocity(O, sd, args, inp, x, t):
    start, end = frames_of(inp)
    return O.forward_inference(sd, args, x, t, inp["mask"], start, end, inp["x_cond"],
                               inp["x_cond_mask"], inp["aatype"])


# ---- gradient goldens (oracle/make_grad_golden.py) -----------------------------------------------------------------
GRAD_CASES = ["sim_small", "tps_small", "design_small", "atlas_small"]


def grad_cotangent(shape):
    return torch.randn(shape, generator=torch.Generator().manual_seed(123))


def grad_projections(name_index, g, nproj=8):
    gen = torch.Generator().manual_seed(7919 * (name_index + 1))
    r = torch.randn(nproj, g.numel(), generator=gen)
    return r @ g.reshape(-1).double().float()


def check_grads(dx, named_grads, gold, tol):
    """dx and {name: grad} against one case of tests/golden/grads.pt.  Every parameter: norm and 8 random projections
    (a gradient error e changes a projection by ~|e|, so the bound is tol * |g| with a 4-sigma margin); parameters with
    <= 4096 elements and dx: full rel-L2.  Returns the worst relative error seen."""
    worst = rel_l2(dx, gold["dx"])
    assert worst < tol, ("dx", worst)
    assert set(named_grads) == set(gold["params"]), set(named_grads) ^ set(gold["params"])
    for name, e in gold["params"].items():
        g = named_grads[name].detach().cpu().float()
        n = e["norm"]
        if n == 0.0:
            assert float(g.norm()) == 0.0, name
            continue
        err_n = abs(float(g.double().norm()) - n) / n
        err_p = float((grad_projections(e["index"], g) - e["proj"]).abs().max()) / n
        assert err_n < tol and err_p < 4 * tol, (name, err_n, err_p)
        worst = max(worst, err_n, err_p / 4)
        if "full" in e:
            err_f = float((g.reshape(-1).double() - e["full"].reshape(-1).double()).norm()) / n
            assert err_f < tol, (name, err_f)
            worst = max(worst, err_f)
    return worst
