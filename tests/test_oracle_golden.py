"""Pins the oracle restatement (oracle/mdgen_oracle.py) against golden vectors produced by
the UNMODIFIED reference (oracle/make_golden.py).  CPU only."""
import pytest
import torch

from oracle import mdgen_oracle as O
from tests.helpers import CASES, load_case, frames_of, rel_l2, oracle_velocity

TOL = 2e-6   # fp32 CPU, same math in a different op order


@pytest.mark.parametrize("name", CASES)
def test_velocity_matches_reference(name):
    g, c, args, sd, inp = load_case(name)
    with torch.no_grad():
        for i, tv in enumerate(g["t_values"]):
            v = oracle_velocity(O, sd, args, inp, inp["x0"], torch.full((c["B"],), tv))
            assert v.shape == g[f"v_t{i}"].shape
            assert rel_l2(v, g[f"v_t{i}"]) < TOL, (name, tv)
        if "v_tvec" in g:
            v = oracle_velocity(O, sd, args, inp, inp["x0"], torch.linspace(0.1, 0.9, c["B"]))
            assert rel_l2(v, g["v_tvec"]) < TOL


@pytest.mark.parametrize("name", ["sim_small", "tps_small", "design_small", "atlas_small"])
def test_euler_matches_reference_sampler(name):
    g, c, args, sd, inp = load_case(name)
    with torch.no_grad():
        f = lambda x, t: oracle_velocity(O, sd, args, inp, x, t)
        if "euler_traj" in g:
            traj = O.euler_sample(f, inp["x0"], g["euler_steps"], return_all=True)
            assert traj.shape == g["euler_traj"].shape
            assert rel_l2(traj, g["euler_traj"]) < TOL
        y = O.euler_sample(f, inp["x0"], g["euler_steps"])
        assert rel_l2(y, g["euler_final"]) < TOL


@pytest.mark.parametrize("name", ["sim_small", "atlas_small"])
def test_submodules_match_reference(name):
    g, c, args, sd, inp = load_case(name)
    start, _ = frames_of(inp)
    m = inp["mask"][:, 0]
    B, T, L, C = c["B"], c["T"], c["L"], args.embed_dim
    with torch.no_grad():
        p = "ipa_layers.1"
        n = torch.nn.functional.layer_norm(g["sub_s"], (C,), sd[p + ".ipa_norm.weight"],
                                           sd[p + ".ipa_norm.bias"], 1e-5)
        assert rel_l2(O.ipa(sd, p + ".ipa", n, *start, m), g["sub_ipa"]) < TOL
        assert rel_l2(O.ipa_layer(sd, p, g["sub_s"], g["sub_te"], m, *start), g["sub_ipalayer"]) < TOL
        assert rel_l2(O.main_layer(sd, "layers.2", g["sub_x"], g["sub_te"][:, None], inp["mask"]),
                      g["sub_mainlayer"]) < TOL
        y = O.mha(sd, "layers.2.mha_t", g["sub_x"].transpose(1, 2).reshape(B * L, T, C),
                  inp["mask"].transpose(1, 2).reshape(B * L, T), args.mha_heads)
        assert rel_l2(y, g["sub_mha_t"]) < TOL


def test_full_size_config1_velocity():
    import os
    from tests.helpers import GOLD
    if not os.path.exists(os.path.join(GOLD, "sim_full.pt")):
        pytest.skip("sim_full golden not generated")
    g, c, args, sd, inp = load_case("sim_full")
    with torch.no_grad():
        v = oracle_velocity(O, sd, args, inp, inp["x0"], torch.full((c["B"],), g["t_values"][0]))
    assert rel_l2(v, g["v_t0"]) < TOL


# ---- row N4: the oracle's autograd against the reference's autograd ----------------------------------------------
from tests.helpers import GRAD_CASES, check_grads, grad_cotangent  # noqa: E402


@pytest.mark.parametrize("name", GRAD_CASES)
def test_oracle_gradients_match_reference_autograd(name):
    """The oracle is a functional restatement in torch, so torch's autograd through it is the gradient of the pinned
    function; this pins it to the gradients of the UNMODIFIED reference (tests/golden/grads.pt)."""
    import os
    g, c, args, sd, inp = load_case(name)
    gold = torch.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "grads.pt"), weights_only=False)[name]
    names = list(gold["params"].keys())
    sdg = {k: (v.clone().requires_grad_(True) if k in gold["params"] else v) for k, v in sd.items()}
    x = inp["x0"].clone().requires_grad_(True)
    start, end = frames_of(inp)
    y = O.forward(sdg, args, x, gold["t"], inp["mask"], start, end, inp["x_cond"], inp["x_cond_mask"], inp["aatype"])
    assert rel_l2(y.detach(), gold["y"]) < TOL
    loss = (y * grad_cotangent(y.shape)).sum()
    grads = torch.autograd.grad(loss, [x] + [sdg[n] for n in names])
    worst = check_grads(grads[0], dict(zip(names, grads[1:])), gold, 2e-5)
    print(name, "worst relative gradient error", worst)
