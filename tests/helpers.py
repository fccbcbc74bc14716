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
