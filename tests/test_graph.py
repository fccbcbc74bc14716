"""CUDA-graph behaviour of the rollout (include/mdgen_b200.h header comment) and the per-device launcher
state (re-entrant per handle + device; the reference's nn.DataParallel path, samplers.py:75-79)."""
import pytest
import torch

from tests.helpers import load_case, rel_l2
from tests.test_gpu_parity import build, gpu_kwargs

pytestmark = pytest.mark.gpu


def _graph_replays():
    from mdgen_finetune_b200 import _lib
    return int(_lib.lib().mdgen_variant_launches(_lib.VARIANTS.index("graph_replays")))


@pytest.mark.parametrize("name", ["sim_small", "tps_small", "design_small"])
def test_rollout_graph_replay_matches_eager(name, monkeypatch):
    """The internally captured whole-rollout graph against the plain launch sequence (MDGEN_NO_GRAPH=1): bit-identical
    (same kernels, same order, same arguments), over repeated replays and after a change of the initial state."""
    g, c, args, sd, inp = load_case(name)
    m = build(args, c["D"], sd, "bf16")
    kw = gpu_kwargs(inp)
    x0 = inp["x0"].cuda()
    with torch.no_grad():
        r0 = _graph_replays()
        a1 = m.euler_rollout(x0, g["euler_steps"], **kw)
        a2 = m.euler_rollout(x0, g["euler_steps"], **kw)
        tr = m.euler_rollout(x0, g["euler_steps"], return_trajectory=True, **kw)
        x1 = x0 * 0.5
        if args.design:
            x1[..., -20:] = x0[..., -20:]
        b1 = m.euler_rollout(x1, g["euler_steps"], **kw)
        assert _graph_replays() - r0 == 4
        monkeypatch.setenv("MDGEN_NO_GRAPH", "1")
        e1 = m.euler_rollout(x0, g["euler_steps"], **kw)
        etr = m.euler_rollout(x0, g["euler_steps"], return_trajectory=True, **kw)
        f1 = m.euler_rollout(x1, g["euler_steps"], **kw)
        assert _graph_replays() - r0 == 4
    assert torch.equal(a1, a2)
    assert torch.equal(a1, e1) and torch.equal(b1, f1) and torch.equal(tr, etr)
    assert torch.equal(tr[-1], a1) and torch.equal(tr[0], x0)
    assert rel_l2(a1.cpu(), g["euler_final"]) < 4e-2


def test_rollout_graph_rebuilds_after_weight_and_conditioning_changes():
    g, c, args, sd, inp = load_case("sim_small")
    m = build(args, c["D"], sd, "bf16")
    kw = gpu_kwargs(inp)
    x0 = inp["x0"].cuda()
    with torch.no_grad():
        a = m.euler_rollout(x0, g["euler_steps"], **kw)
        # new conditioning values at the same shape: buffers are refilled in place, the graph is reused
        kw2 = dict(kw)
        kw2["x_cond"] = kw["x_cond"] * 2.0
        b = m.euler_rollout(x0, g["euler_steps"], **kw2)
        assert not torch.equal(a, b)
        a_again = m.euler_rollout(x0, g["euler_steps"], **kw)
        assert torch.equal(a, a_again)
        # weight update: packed buffers are re-created, the graph must be rebuilt
        sd2 = {k: (v * 1.01 if v.dtype.is_floating_point and "inv_freq" not in k and "pos_embed" not in k else v) for k, v in sd.items()}
        m.load_state_dict(sd2)
        w = m.euler_rollout(x0, g["euler_steps"], **kw)
        assert not torch.equal(a, w)
        m.load_state_dict(sd)
        assert torch.equal(m.euler_rollout(x0, g["euler_steps"], **kw), a)
        # different step count
        s8 = m.euler_rollout(x0, 8, **kw)
        assert torch.isfinite(s8).all() and not torch.equal(s8, a)


def test_rollout_is_capturable_into_a_callers_graph():
    """torch.cuda.graph around the C-ABI call: no allocation, host copy or synchronisation inside (after one un-captured
    call of the same shape); replaying the caller's graph reproduces the direct call and follows the input buffer."""
    g, c, args, sd, inp = load_case("sim_small")
    m = build(args, c["D"], sd, "bf16")
    kw = gpu_kwargs(inp)
    x_static = inp["x0"].cuda().clone()
    with torch.no_grad():
        ref = m.euler_rollout(x_static, g["euler_steps"], **kw)
        v_ref = m.forward(x_static, torch.full((c["B"],), 0.3).cuda(), **kw)
        t_static = torch.full((c["B"],), 0.3).cuda()
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        side = torch.cuda.Stream()
        with torch.cuda.stream(side):
            with torch.cuda.graph(graph, stream=side):
                out = m.euler_rollout(x_static, g["euler_steps"], **kw)
                v = m.forward(x_static, t_static, **kw)
        graph.replay()
        torch.cuda.synchronize()
        assert torch.equal(out, ref) and torch.equal(v, v_ref)
        x_static.mul_(0.5)
        graph.replay()
        torch.cuda.synchronize()
        ref2 = m.euler_rollout(x_static, g["euler_steps"], **kw)
        assert torch.equal(out, ref2) and not torch.equal(ref2, ref)


@pytest.mark.skipif(torch.cuda.is_available() and torch.cuda.device_count() < 2, reason="needs two GPUs (gpurun --gpus 2)")
def test_two_handles_on_two_devices_in_one_process():
    """One process, two devices, one handle each (what nn.DataParallel(policy.unet) does, samplers.py:75-79): every
    kernel's dynamic-shared-memory attribute and the SM count are configured per device."""
    g, c, args, sd, inp = load_case("upsample_small")
    outs = []
    for dev in (0, 1):
        with torch.cuda.device(dev):
            from mdgen_finetune_b200 import LatentMDGenModel
            m = LatentMDGenModel(args, c["D"], precision="bf16", quat_sign="eigh_cpu")
            m.load_state_dict(sd)
            m = m.to(f"cuda:{dev}").eval()
            kw = {k: (tuple(t.to(f"cuda:{dev}") for t in v) if isinstance(v, tuple) else v.to(f"cuda:{dev}"))
                  for k, v in gpu_kwargs(inp).items()}
            with torch.no_grad():
                v = m.forward(inp["x0"].to(f"cuda:{dev}"), torch.full((c["B"],), 0.3, device=f"cuda:{dev}"), **kw)
                r = m.euler_rollout(inp["x0"].to(f"cuda:{dev}"), 6, **kw)
            outs.append((v.cpu(), r.cpu(), m))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
    assert rel_l2(outs[0][0], g["v_t1"]) < 2e-2
