"""CPU checks of the boundary: the C-ABI library builds, loads and exports every symbol that
include/mdgen_b200.h declares; the host module keeps the reference's state_dict layout."""
import ctypes
import os
import re

import pytest
import torch

from mdgen_finetune_b200 import _lib, configs as cfgs
from mdgen_finetune_b200.weights import param_shapes, random_state_dict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    path = _lib.build()
    lib = ctypes.CDLL(path)
    header = open(os.path.join(ROOT, "include", "mdgen_b200.h")).read()
    declared = set(re.findall(r"\b(mdgen_[a-z_0-9]+)\s*\(", header))
    assert declared == set(_lib.EXPORTS)
    for name in declared:
        assert hasattr(lib, name), name
    lib.mdgen_abi_version.restype = ctypes.c_int
    assert lib.mdgen_abi_version() == 1


@pytest.mark.parametrize("cfg", [1, 2, 3, 4, 5])
def test_state_dict_layout_and_strict_load(cfg):
    from mdgen_finetune_b200 import LatentMDGenModel
    args = cfgs.make_args(cfg)
    D = cfgs.latent_dim_of(cfg)
    m = LatentMDGenModel(args, D)
    sd = random_state_dict(args, D, seed=1)
    assert list(sd.keys()) == list(param_shapes(args, D).keys())
    m.load_state_dict(sd, strict=True)
    n = sum(p.numel() for p in m.parameters())
    assert abs(n / 1e6 - {1: 34.15, 2: 34.17, 3: 34.15, 4: 34.63, 5: 34.15}[cfg]) < 0.01
    # fresh model follows the reference init: adaLN-Zero => zero-initialised output head (Q2)
    fresh = LatentMDGenModel(args, D)
    assert not fresh.emb_to_latent.linear.weight.any()
    assert not fresh.layers[0].adaLN_modulation[-1].weight.any()


def test_cpu_call_fails_loudly():
    from mdgen_finetune_b200 import LatentMDGenModel
    args = cfgs.make_args(cfgs.small_variant(1, B=1, T=2, L=4))
    m = LatentMDGenModel(args, 21)
    x = torch.zeros(1, 2, 4, 21)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m.forward(x, torch.zeros(1), torch.ones(1, 2, 4))


def test_forward_refuses_to_return_a_gradless_output_when_gradients_are_expected(monkeypatch):
    """RTB call convention (samplers.py:136-139): parameters require grad, x does not.  The C ABI output has no grad_fn,
    so forward() must refuse instead of silently training nothing."""
    from mdgen_finetune_b200 import LatentMDGenModel
    from mdgen_finetune_b200 import model as M
    args = cfgs.make_args(cfgs.small_variant(1, B=1, T=2, L=4))
    m = LatentMDGenModel(args, 21)
    monkeypatch.setattr(M.LatentMDGenModel, "device", property(lambda self: torch.device("cuda", 0)))
    x = torch.zeros(1, 2, 4, 21)
    assert not x.requires_grad and any(p.requires_grad for p in m.parameters())
    with pytest.raises(NotImplementedError, match="row N4"):
        m.forward(x, torch.zeros(1), torch.ones(1, 2, 4))


def test_sampler_generic_callable_matches_oracle_euler():
    """Sampler.sample_ode('euler') on a plain callable == oracle euler_sample (torchdiffeq restatement)."""
    from mdgen_finetune_b200 import Sampler, create_transport
    from oracle import mdgen_oracle as O
    args = cfgs.make_args(1)
    f = lambda x, t: torch.sin(x) * (1 + t.view(-1, 1, 1, 1))
    x0 = torch.randn(2, 3, 4, 5)
    traj = Sampler(create_transport(args)).sample_ode(sampling_method="euler", num_steps=50)(x0, f)
    ref = O.euler_sample(f, x0, 50, return_all=True)
    assert traj.shape == (50, 2, 3, 4, 5)
    assert torch.equal(traj, ref)
    with pytest.raises(NotImplementedError):
        Sampler(create_transport(args)).sample_ode(sampling_method="dopri5")
