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



def test_variant_names_and_switch_docs_match_the_sources():
    """_lib.VARIANTS mirrors `enum Variant` of csrc/common.cuh (index = what mdgen_variant_launches takes, as the header says),
    and every MDGEN_* switch read_switches() reads is described in INTEGRATION.md or DESIGN.md."""
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    common = open(os.path.join(root, "mdgen_finetune_b200", "csrc", "common.cuh")).read()
    m = re.search(r"enum Variant : int \{([^}]*)\}", common)
    names = [v.strip().split("=")[0].strip() for v in m.group(1).split(",")]
    assert names[-1] == "V_COUNT" and len(names) - 1 == len(_lib.VARIANTS), (names, _lib.VARIANTS)
    header = open(os.path.join(root, "include", "mdgen_b200.h")).read()
    for i, v in enumerate(_lib.VARIANTS):
        assert re.search(r"\b%d %s\b" % (i, v.replace("graph_replays", "whole-rollout graph replays")), header), (i, v)
    model = open(os.path.join(root, "mdgen_finetune_b200", "csrc", "model.cu")).read()
    body = model[model.index("void read_switches()"):]
    body = body[:body.index("\n}\n")]
    switches = set(re.findall(r'"(MDGEN_[A-Z0-9_]+)"', body))
    docs = open(os.path.join(root, "INTEGRATION.md")).read() + open(os.path.join(root, "DESIGN.md")).read()
    missing = sorted(v for v in switches if v not in docs)
    assert not missing, missing


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


def test_forward_routes_gradient_calls_through_the_autograd_function(monkeypatch):
    """RTB call convention (samplers.py:136-139): parameters require grad, x does not.  The output must carry a grad_fn
    (autograd.VelocityFn -> mdgen_backward); frozen modules / no_grad return plain tensors.  CPU: the C-ABI call is
    replaced by a stub, only the dispatch is checked."""
    from mdgen_finetune_b200 import LatentMDGenModel
    from mdgen_finetune_b200 import model as M
    args = cfgs.make_args(cfgs.small_variant(1, B=1, T=2, L=4))
    m = LatentMDGenModel(args, 21)
    monkeypatch.setattr(M.LatentMDGenModel, "device", property(lambda self: torch.device("cuda", 0)))
    monkeypatch.setattr(M.LatentMDGenModel, "_forward_no_grad", lambda self, x, t, cond: torch.zeros_like(x))
    x = torch.zeros(1, 2, 4, 21)
    y = m.forward(x, torch.zeros(1), torch.ones(1, 2, 4), peptide=["AAAA"])
    assert y.requires_grad and "VelocityFn" in type(y.grad_fn).__name__
    with torch.no_grad():
        assert not m.forward(x, torch.zeros(1), torch.ones(1, 2, 4)).requires_grad
    m.requires_grad_(False)
    assert m.forward(x, torch.zeros(1), torch.ones(1, 2, 4)).grad_fn is None
    assert m.forward(x.clone().requires_grad_(), torch.zeros(1), torch.ones(1, 2, 4)).grad_fn is not None


class _LoraLinear(torch.nn.Module):
    """PEFT's LoraLayer layout (peft is not in this image): base_layer + lora_A / lora_B ModuleDicts + scaling."""

    def __init__(self, base, r=4, alpha=8.0):
        super().__init__()
        self.base_layer = base
        self.lora_A = torch.nn.ModuleDict({"default": torch.nn.Linear(base.in_features, r, bias=False)})
        self.lora_B = torch.nn.ModuleDict({"default": torch.nn.Linear(r, base.out_features, bias=False)})
        self.scaling = {"default": alpha / r}
        self.active_adapters = ["default"]
        self.merged = False


def test_lora_wrapped_linears_contribute_merged_weights():
    """gfn_diffusion.py:58-83 attaches LoRA to the named Linear children; the CUDA path must see W + scaling * B @ A and
    autograd must route the weight gradient to A and B."""
    from mdgen_finetune_b200 import LatentMDGenModel
    from mdgen_finetune_b200.autograd import effective_tensors
    args = cfgs.make_args(cfgs.small_variant(1, B=1, T=2, L=4))
    m = LatentMDGenModel(args, 21)
    base = m.layers[0].mha_t.attn.q_proj
    w0 = base.weight.detach().clone()
    m.requires_grad_(False)
    lora = _LoraLinear(base)
    torch.nn.init.normal_(lora.lora_B["default"].weight, std=0.1)
    m.layers[0].mha_t.attn.q_proj = lora
    names = m._ref_names
    assert "layers.0.mha_t.attn.q_proj.weight" in names              # reference names are kept
    eff = effective_tensors(m, names)
    A, B = lora.lora_A["default"].weight, lora.lora_B["default"].weight
    assert torch.allclose(eff["layers.0.mha_t.attn.q_proj.weight"], w0 + 2.0 * (B @ A))
    assert eff["layers.0.mha_t.attn.q_proj.bias"] is base.bias
    assert eff["layers.0.mha_t.attn.k_proj.weight"] is m.layers[0].mha_t.attn.k_proj.weight
    g = torch.randn_like(w0)
    (eff["layers.0.mha_t.attn.q_proj.weight"] * g).sum().backward()
    assert torch.allclose(A.grad, 2.0 * B.detach().T @ g) and torch.allclose(B.grad, 2.0 * g @ A.detach().T)
    assert base.weight.grad is None
    # the weight cache key sees the adapter parameters
    k0 = m._raw_key()
    with torch.no_grad():
        A.add_(1.0)
    assert m._raw_key() != k0


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
