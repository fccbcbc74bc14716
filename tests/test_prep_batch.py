"""prep_batch (wrapper.py:283-365, row N1): the oracle restatement against goldens produced by the reference's own
Rigid / get_offsets code (oracle/make_prep_golden.py), and the CUDA path (mdgen_prep_batch) against the same goldens."""
import argparse
import os

import pytest
import torch

from oracle import mdgen_oracle as O

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "prep_batch.pt")
FLAGS = ("sim_condition", "tps_condition", "inpainting", "design", "dynamic_mpnn", "mpnn", "no_offsets", "no_torsion",
         "no_design_torsion", "design_key_frames", "supervise_all_torsions", "supervise_no_torsions", "no_frames")
CASES = ["sim", "upsample", "tps", "design", "design_key_frames", "no_offsets"]


def load(name):
    g = torch.load(GOLDEN)["cases"][name]
    a = argparse.Namespace(**{f: False for f in FLAGS}, cond_interval=None)
    for k, v in g["args"].items():
        setattr(a, k, v)
    return a, g["batch"], g["out"]


def check(got, ref, tol):
    kw = got["model_kwargs"]
    assert got["latents"].shape == ref["latents"].shape
    assert (got["latents"].cpu() - ref["latents"]).abs().max() < tol
    assert torch.equal(got["loss_mask"].cpu().float(), ref["loss_mask"].float())
    assert torch.equal(kw["x_cond_mask"].cpu(), ref["x_cond_mask"])
    assert torch.equal(kw["aatype"].cpu(), ref["aatype"])
    assert torch.equal(kw["mask"].cpu().float(), ref["mask"].float())
    assert (kw["x_cond"].cpu() - ref["x_cond"]).abs().max() < tol
    for k, v in (("start_frames", ("start_rot", "start_trans")), ("end_frames", ("end_rot", "end_trans"))):
        assert torch.equal(kw[k][0].cpu(), ref[v[0]]) and torch.equal(kw[k][1].cpu(), ref[v[1]])


@pytest.mark.parametrize("name", CASES)
def test_oracle_prep_batch_vs_reference_golden(name):
    args, batch, ref = load(name)
    check(O.prep_batch(args, batch), ref, 2e-6)
    lat = ref["latents"]
    assert (lat[..., 0] >= 0).all()                                    # quaternion real part (wrapper.py:308-309)
    assert torch.allclose(lat[:, 0, :, :4], torch.tensor([1.0, 0, 0, 0]).expand_as(lat[:, 0, :, :4]), atol=1e-5) or args.no_offsets


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_cuda_prep_batch_vs_reference_golden(name):
    from mdgen_finetune_b200 import prep_batch
    args, batch, ref = load(name)
    got = prep_batch(args, {k: v.cuda() for k, v in batch.items()})
    assert got["latents"].is_cuda
    check(got, ref, 2e-5)     # fp32 rounding of R^T (t - t0) on coordinates of tens of Angstrom
