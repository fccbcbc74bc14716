"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/*.pt by running the UNMODIFIED
reference (/root/reference/src, via oracle/ref_loader.py) on CPU in fp32.

Run in the build container:   python oracle/make_golden.py
The GPU box has no /root/reference; tests read only the committed .pt files.

Every case is regenerable from (config variant, seed): weights come from
mdgen_finetune_b200.weights.random_state_dict (loaded into the reference model with
strict=True -- which also pins the state_dict layout of SURVEY 8b), inputs from
mdgen_finetune_b200.configs.synthetic_inputs.  Goldens hold only reference OUTPUTS.
"""
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from mdgen_finetune_b200 import configs as cfgs          # noqa: E402
from mdgen_finetune_b200.weights import random_state_dict  # noqa: E402
from oracle.ref_loader import load_reference             # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")

# name -> (base config, shape overrides, pad_last)
CASES = {
    "sim_small": (1, dict(B=2, T=12, L=4), 0),
    "tps_small": (2, dict(B=3, T=10, L=4), 0),
    "upsample_small": (3, dict(B=2, T=20, L=4), 0),
    "design_small": (4, dict(B=2, T=8, L=4), 0),
    "atlas_small": (5, dict(B=1, T=6, L=24), 5),
    "atlas_mid": (5, dict(B=1, T=20, L=72), 9),
    "sim_full": (1, dict(), 0),
    # BASELINE configs at FULL size: the shapes that select the kernels bench.py times (fused projection + LayerNorm
    # from 8192 rows, CTA-pair GEMM from 74 tiles of 256 x 256, tcgen05 attention on both axes of ATLAS, IPA at L = 256)
    "tps_full": (2, dict(), 0),
    "upsample_full": (3, dict(), 0),
    "design_full": (4, dict(), 0),
    "atlas_full": (5, dict(), 32),
}
T_VALUES = [0.0, 0.3, 48.0 / 49.0]
# full-size cases: (velocity times, Euler grid points of the short reference-sampler rollout)
FULL = {"sim_full": ([0.3], 50), "tps_full": ([0.3], 5), "upsample_full": ([0.3, 48.0 / 49.0], 5),
        "design_full": ([0.3], 4), "atlas_full": ([0.3], 3)}


def case_config(name):
    base, ov, pad = CASES[name]
    c = cfgs.small_variant(base, **ov) if ov else dict(cfgs.CONFIGS[base])
    c["id"] = base
    return c, pad


def build_reference_model(ref, c, seed=0):
    args = cfgs.make_args(c)
    model = ref.LatentMDGenModel(args, c["D"])
    sd = random_state_dict(args, c["D"], seed=seed)
    model.load_state_dict(sd, strict=True)
    return model.eval(), args, sd


def ref_kwargs(ref, inp):
    def rigid(rot, trans):
        return ref.Rigid(ref.Rotation(rot_mats=rot), trans)
    kw = dict(mask=inp["mask"], start_frames=rigid(inp["start_rot"], inp["start_trans"]),
              x_cond=inp["x_cond"], x_cond_mask=inp["x_cond_mask"], aatype=inp["aatype"])
    if "end_rot" in inp:
        kw["end_frames"] = rigid(inp["end_rot"], inp["end_trans"])
    return kw


@torch.no_grad()
def main(only=None):
    ref = load_reference()
    os.makedirs(GOLD, exist_ok=True)
    for name in CASES:
        if only and name not in only:
            continue
        t0 = time.time()
        c, pad = case_config(name)
        model, args, sd = build_reference_model(ref, c)
        inp = cfgs.synthetic_inputs(c, seed=0, pad_last=pad)
        kw = ref_kwargs(ref, inp)
        B = c["B"]
        out = {"case": name, "config": c, "pad_last": pad}
        tvals = FULL[name][0] if name in FULL else T_VALUES
        for i, tv in enumerate(tvals):
            t = torch.full((B,), tv)
            out[f"v_t{i}"] = model.forward_inference(inp["x0"], t, **kw).clone()
        out["t_values"] = tvals
        if B > 1 and name not in FULL:  # per-sample t (training / RTB call convention)
            t = torch.linspace(0.1, 0.9, B)
            out["v_tvec"] = model.forward_inference(inp["x0"], t, **kw).clone()
        # sampler: the reference's own Sampler / ode classes (transport.py:407-450)
        transport = ref.create_transport(args, args.path_type, args.prediction)
        steps = FULL[name][1] if name in FULL else 6
        if name in ("sim_small", "tps_small", "design_small", "atlas_small") or name in FULL:
            fn = ref.Sampler(transport).sample_ode(sampling_method="euler", num_steps=steps)
            traj = fn(inp["x0"], lambda x, t, **k: model.forward_inference(x, t, **k), **kw)
            out["euler_steps"] = steps
            out["euler_final"] = traj[-1].clone()
            if name == "sim_small":
                out["euler_traj"] = traj.clone()
        # sub-module goldens (pin the oracle's pieces)
        if name in ("sim_small", "atlas_small"):
            g = torch.Generator().manual_seed(7)
            C = args.embed_dim
            te = torch.randn(B, C, generator=g)
            s = torch.randn(B, c["L"], C, generator=g)
            m = inp["mask"][:, 0]
            lay = model.ipa_layers[1]
            out["sub_te"], out["sub_s"] = te, s
            out["sub_ipa"] = lay.ipa(lay.ipa_norm(s), kw["start_frames"], frame_mask=m).clone()
            out["sub_ipalayer"] = lay(s, te, m, frames=kw["start_frames"]).clone()
            xx = torch.randn(B, c["T"], c["L"], C, generator=g)
            out["sub_x"] = xx
            ml = model.layers[2]
            out["sub_mainlayer"] = ml(xx, te[:, None], inp["mask"], kw["start_frames"]).clone()
            out["sub_mha_t"] = ml.mha_t(xx.transpose(1, 2).reshape(B * c["L"], c["T"], C),
                                        mask=inp["mask"].transpose(1, 2).reshape(B * c["L"], c["T"])).clone()
        torch.save(out, os.path.join(GOLD, name + ".pt"))
        print(f"{name}: {time.time() - t0:.1f}s  |v|={out['v_t0'].norm():.4f}", flush=True)


if __name__ == "__main__":
    main(sys.argv[1:] or None)
