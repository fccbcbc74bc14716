"""Row N4: gradients through the CUDA denoiser (mdgen_backward behind autograd.VelocityFn) against
  * the gradients of the UNMODIFIED reference's autograd (tests/golden/grads.pt, oracle/make_grad_golden.py), and
  * the oracle's autograd in full, parameter by parameter (the oracle is pinned to those goldens on CPU),
and the reference's call conventions around it: HGFNode.run_policy (int64 t, torch.utils.checkpoint, stray kwargs;
samplers.py:120-146), LoRA on the named Linear children (gfn_diffusion.py:58-83), an optimiser step.
Tolerances: north_star's 1e-4 (fp32 check mode) / 2e-2 (bf16 handle: the forward is bf16, the gradient the fp32 one)."""
import os

import pytest
import torch

from oracle import mdgen_oracle as O
from tests.helpers import GRAD_CASES, check_grads, frames_of, grad_cotangent, load_case, rel_l2
from tests.test_gpu_parity import build, gpu_kwargs

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "grads.pt")


def _cuda_grads(m, inp, kw, t, need_x=True):
    x = inp["x0"].cuda().clone().requires_grad_(need_x)
    y = m.forward(x, t.cuda(), **kw)
    loss = (y * grad_cotangent(y.shape).cuda()).sum()
    m.zero_grad(set_to_none=True)
    loss.backward()
    grads = {n: p.grad for n, p in m.named_parameters() if p.grad is not None}
    return y.detach().cpu(), (x.grad.cpu() if need_x else None), grads


@pytest.mark.parametrize("prec", ["fp32", "bf16"])
@pytest.mark.parametrize("name", GRAD_CASES)
def test_gradients_vs_reference_autograd(name, prec):
    g, c, args, sd, inp = load_case(name)
    gold = torch.load(GOLD, weights_only=False)[name]
    m = build(args, c["D"], sd, prec).train()
    kw = gpu_kwargs(inp)
    y, dx, grads = _cuda_grads(m, inp, kw, gold["t"])
    assert rel_l2(y, gold["y"]) < (1e-4 if prec == "fp32" else 2e-2)
    worst = check_grads(dx, grads, gold, 1e-4)          # the gradient itself is fp32 in both modes
    print(name, prec, "worst relative gradient error vs reference autograd", worst)


@pytest.mark.parametrize("name", ["sim_small", "design_small"])
def test_gradients_vs_oracle_autograd_in_full(name):
    """Every element of every parameter gradient (the goldens hold projections for the large ones)."""
    g, c, args, sd, inp = load_case(name)
    gold = torch.load(GOLD, weights_only=False)[name]
    m = build(args, c["D"], sd, "fp32").train()
    _, dx, grads = _cuda_grads(m, inp, gpu_kwargs(inp), gold["t"])
    names = list(gold["params"].keys())
    sdg = {k: (v.clone().requires_grad_(True) if k in gold["params"] else v) for k, v in sd.items()}
    x = inp["x0"].clone().requires_grad_(True)
    start, end = frames_of(inp)
    y = O.forward(sdg, args, x, gold["t"], inp["mask"], start, end, inp["x_cond"], inp["x_cond_mask"], inp["aatype"])
    ref = torch.autograd.grad((y * grad_cotangent(y.shape)).sum(), [x] + [sdg[n] for n in names])
    assert rel_l2(dx, ref[0]) < 1e-4
    for n, r in zip(names, ref[1:]):
        if float(r.norm()) == 0.0:
            assert float(grads[n].norm()) == 0.0, n
            continue
        err = rel_l2(grads[n].cpu().reshape(r.shape), r)
        assert err < 1e-4, (n, err)


def test_run_policy_call_convention_checkpoint_and_frozen_x():
    """HGFNode.run_policy (samplers.py:120-146): long timesteps, **condition with a stray `peptide` key, optional
    torch.utils.checkpoint(use_reentrant=False), x without requires_grad, several forwards before ONE backward
    (PosteriorPriorDGFN.sample_fwd accumulates over the trajectory, samplers.py:379-486)."""
    import torch.utils.checkpoint
    g, c, args, sd, inp = load_case("design_small")
    m = build(args, c["D"], sd, "fp32").train()
    condition = dict(gpu_kwargs(inp), peptide=["AAAA"] * c["B"])
    x = inp["x0"].cuda()
    t_long = torch.full((c["B"],), 0, dtype=torch.long).cuda()

    def trajectory_loss(use_ckpt):
        m.zero_grad(set_to_none=True)
        tot = 0.0
        xx = x
        for k in range(3):                       # conditioning identical, state changes: the tape is rebuilt per backward
            tk = t_long + k
            res = (torch.utils.checkpoint.checkpoint(m, xx, tk, **condition, use_reentrant=False) if use_ckpt
                   else m(xx, tk, **condition))
            assert res.requires_grad and not xx.requires_grad
            tot = tot + (res * res).mean()
            xx = (xx + 0.1 * res.detach())
            if args.design:                      # keep the simplex channels on the simplex
                xx = torch.cat([xx[..., :-20], x[..., -20:]], -1)
        tot.backward()
        return float(tot.detach()), {n: p.grad.clone() for n, p in m.named_parameters()}

    l0, g0 = trajectory_loss(False)
    l1, g1 = trajectory_loss(True)
    assert abs(l0 - l1) < 1e-6 * max(1.0, abs(l0))
    for n in g0:       # atomics make the summation order (not the value) differ run to run: compare in norm
        if float(g0[n].norm()) > 0:
            assert rel_l2(g1[n].cpu(), g0[n].cpu()) < 1e-5, (n, rel_l2(g1[n].cpu(), g0[n].cpu()))
    assert all(torch.isfinite(v).all() for v in g0.values())
    assert sum(float(v.abs().sum()) for v in g0.values()) > 0


def test_lora_adapters_train_through_the_cuda_path():
    """gfn_diffusion.py:58-83: LoRA on the named Linear children, base weights frozen.  Gradients of the adapters against
    the oracle's autograd on the merged weights; one SGD step changes the CUDA output."""
    from tests.test_abi_cpu import _LoraLinear
    g, c, args, sd, inp = load_case("sim_small")
    m = build(args, c["D"], sd, "fp32").train()
    m.requires_grad_(False)
    torch.manual_seed(0)
    targets = {"layers.1.mha_t.attn.q_proj": m.layers[1].mha_t.attn, "layers.0.fc1": m.layers[0],
               "ipa_layers.2.ipa.linear_kv": m.ipa_layers[2].ipa, "t_embedder.mlp.2": m.t_embedder.mlp}
    loras = {}
    for full, parent in targets.items():
        leaf = full.split(".")[-1]
        base = parent[int(leaf)] if leaf.isdigit() else getattr(parent, leaf)
        lo = _LoraLinear(base, r=4, alpha=8.0).cuda()
        torch.nn.init.normal_(lo.lora_B["default"].weight, std=0.05)
        if leaf.isdigit():
            parent[int(leaf)] = lo
        else:
            setattr(parent, leaf, lo)
        loras[full] = lo
    kw = gpu_kwargs(inp)
    t = torch.linspace(0.1, 0.9, c["B"])
    y, _, _ = _cuda_grads(m, inp, kw, t, need_x=False)
    # oracle on the merged weights, adapters as leaves
    sdo = dict(sd)
    leaves = {}
    for full, lo in loras.items():
        A = lo.lora_A["default"].weight.detach().cpu().clone().requires_grad_(True)
        Bm = lo.lora_B["default"].weight.detach().cpu().clone().requires_grad_(True)
        leaves[full] = (A, Bm)
        sdo[full + ".weight"] = sd[full + ".weight"] + 2.0 * (Bm @ A)
    start, end = frames_of(inp)
    yo = O.forward(sdo, args, inp["x0"], t, inp["mask"], start, end, inp["x_cond"], inp["x_cond_mask"], inp["aatype"])
    assert rel_l2(y, yo.detach()) < 1e-4
    flat = [v for ab in leaves.values() for v in ab]
    ref = torch.autograd.grad((yo * grad_cotangent(yo.shape)).sum(), flat)
    k = 0
    for full, lo in loras.items():
        for mod in (lo.lora_A["default"], lo.lora_B["default"]):
            assert mod.weight.grad is not None, full
            assert rel_l2(mod.weight.grad.cpu(), ref[k]) < 1e-4, (full, rel_l2(mod.weight.grad.cpu(), ref[k]))
            k += 1
        assert lo.base_layer.weight.grad is None
    opt = torch.optim.SGD([p for p in m.parameters() if p.requires_grad], lr=0.5)
    opt.step()
    with torch.no_grad():
        y2 = m.forward(inp["x0"].cuda(), t.cuda(), **kw).cpu()
    assert rel_l2(y2, y) > 1e-4                       # the updated adapters reached the packed CUDA weights


def test_full_parameter_training_step_reduces_a_regression_loss():
    g, c, args, sd, inp = load_case("tps_small")
    m = build(args, c["D"], sd, "bf16").train()
    kw = gpu_kwargs(inp)
    x, t = inp["x0"].cuda(), torch.linspace(0.1, 0.9, c["B"]).cuda()
    target = torch.randn(x.shape, generator=torch.Generator().manual_seed(5)).cuda() * 0.1
    opt = torch.optim.Adam(m.parameters(), lr=2e-4)
    losses = []
    for _ in range(6):
        opt.zero_grad(set_to_none=True)
        loss = ((m(x, t, **kw) - target) ** 2).mean()
        loss.backward()
        torch.nn.utils.clip_grad_norm_(m.parameters(), 1.0)       # gfn_diffusion.py:216
        opt.step()
        losses.append(float(loss))
    assert all(torch.isfinite(torch.tensor(losses)))
    assert losses[-1] < losses[0], losses
