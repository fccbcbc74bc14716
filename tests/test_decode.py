"""decode (wrapper.py:487-508, row N2): sampled latents -> atom14.  The oracle restatement against goldens replayed on the
reference's Rigid / geometry.frames_torsions_to_atom14 (oracle/make_decode_golden.py), and the CUDA path
(mdgen_decode_atom14) against the same goldens.  The residue tables are inputs (stored in the fixture)."""
import argparse
import os

import pytest
import torch

from oracle import mdgen_oracle as O

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "decode.pt")
CASES = ["sim", "tps", "design", "no_offsets", "oracle_torsions"]


def load(name):
    g = torch.load(GOLDEN)
    c = g["cases"][name]
    a = argparse.Namespace(tps_condition=False, inpainting=False, design=False, no_offsets=False, oracle=False, no_frames=False)
    for k, v in c["args"].items():
        setattr(a, k, v)
    return a, c, g["tables"]


@pytest.mark.parametrize("name", CASES)
def test_oracle_decode_vs_reference_golden(name):
    args, c, tables = load(name)
    atom14, aa = O.decode_samples(args, c["samples"], c["rot0"], c["trans0"], c["seqres"], tables)
    assert atom14.shape == c["atom14"].shape
    assert (atom14 - c["atom14"]).abs().max() < 2e-4           # Angstrom; coordinates of tens of Angstrom in fp32
    assert torch.equal(aa, c["aa_out"])


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_cuda_decode_vs_reference_golden(name):
    from mdgen_finetune_b200 import decode_samples
    args, c, tables = load(name)
    dev = torch.device("cuda")
    atom14, aa = decode_samples(args, c["samples"].to(dev), (c["rot0"].to(dev), c["trans0"].to(dev)), c["seqres"].to(dev),
                                {k: v.to(dev) for k, v in tables.items()})
    assert atom14.is_cuda and atom14.shape == c["atom14"].shape
    assert (atom14.cpu() - c["atom14"]).abs().max() < 2e-4
    assert torch.equal(aa.cpu(), c["aa_out"])
