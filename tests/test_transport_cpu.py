"""Transport.training_losses + the interpolant plans (transport.py:125-222, path.py:112-190) against goldens produced by
the UNMODIFIED reference's Transport (oracle/make_transport_golden.py).  CPU only: the model is an analytic stand-in."""
import os

import pytest
import torch

from mdgen_finetune_b200 import create_transport
from oracle.make_transport_golden import CASES, case_inputs, standin_model

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "transport.pt")


@pytest.mark.parametrize("name", list(CASES))
def test_training_losses_match_reference(name):
    gold = torch.load(GOLD, weights_only=False)[name]
    cfg_id, path = CASES[name]
    args, x1, mask, aatype1 = case_inputs(cfg_id)
    tr = create_transport(args, path, "velocity")
    torch.manual_seed(11)
    terms = tr.training_losses(standin_model, x1, aatype1=aatype1 if args.design else None, mask=mask,
                               model_kwargs={"shift": 0.5})
    assert set(terms) == set(gold)
    for k, v in gold.items():
        assert torch.allclose(terms[k], v, rtol=1e-6, atol=1e-7), k
    assert terms["loss"].shape == (3,) if not args.design else terms["loss"].dim() == 1


def test_plans():
    from mdgen_finetune_b200.transport import GVPCPlan, ICPlan
    t = torch.tensor([0.0, 0.25, 1.0])
    x0, x1 = torch.randn(3, 2, 2), torch.randn(3, 2, 2)
    _, xt, ut = ICPlan().plan(t, x0, x1)
    assert torch.allclose(xt[0], x0[0]) and torch.allclose(xt[2], x1[2]) and torch.allclose(ut, x1 - x0)
    _, xt, ut = GVPCPlan().plan(t, x0, x1)
    assert torch.allclose(xt[0], x0[0], atol=1e-6) and torch.allclose(xt[2], x1[2], atol=1e-6)
    assert torch.allclose(ut[0], torch.pi / 2 * x1[0], atol=1e-6)
