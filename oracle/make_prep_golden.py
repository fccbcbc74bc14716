"""TEST INFRASTRUCTURE ONLY -- golden vectors for prep_batch (wrapper.py:283-365), row N1.

The goldens are the outputs of the reference's OWN `NewMDGenWrapper.prep_batch` (mdgen/wrapper.py, unmodified),
imported through oracle/ref_loader.load_wrapper() (pytorch_lightning / wandb stubbed: absent from this image, no
arithmetic).  Control flow (flag handling, key-frame remapping, mask assembly) and arithmetic (invert, compose,
rot_to_quat through eigh, to_tensor_7) are therefore both the reference's.
Run in the build container:  python -m oracle.make_prep_golden   ->  tests/golden/prep_batch.pt"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mdgen_finetune_b200 import configs as cfgs  # noqa: E402
from oracle.ref_loader import load_wrapper  # noqa: E402

# name -> (BASELINE config whose model flags the wrapper is built with, prep_batch flags)
CASES = {
    "sim": (1, dict()),
    "upsample": (1, dict(cond_interval=3)),
    "tps": (2, dict()),
    "design": (4, dict(no_torsion=True)),
    "design_key_frames": (4, dict(design_key_frames=True, no_design_torsion=True)),
    "no_offsets": (1, dict(no_offsets=True, supervise_all_torsions=True)),
}
WRAPPER_FLAGS = dict(no_torsion=False, supervise_all_torsions=False, supervise_no_torsions=False, design_key_frames=False,
                     no_design_torsion=False, cond_interval=None, no_offsets=False, no_frames=False, oracle=False, ema=False)


def wrapper_args(cfg_id, T, **kw):
    """Namespace for NewMDGenWrapper(args): the model flags of BASELINE config `cfg_id` at T frames x 4 residues + the
    wrapper's own flags (wrapper.py:178-194)."""
    c = cfgs.small_variant(cfg_id, T=T, L=4)
    a = cfgs.make_args(c, **{**WRAPPER_FLAGS, **kw})
    return a, c


def synthetic_batch(B=2, T=7, L=4, seed=0):
    g = torch.Generator().manual_seed(seed)
    q = torch.nn.functional.normalize(torch.randn(B, T, L, 4, generator=g), dim=-1)
    w, x, y, z = q.unbind(-1)
    rots = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                        2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                        2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], -1).reshape(B, T, L, 3, 3)
    step = torch.nn.functional.normalize(torch.randn(B, T, L, 3, generator=g), dim=-1) * 3.8
    trans = step.cumsum(2) + torch.randn(B, T, 1, 3, generator=g)
    tors = torch.nn.functional.normalize(torch.randn(B, T, L, 7, 2, generator=g), dim=-1)
    return {"rots": rots, "trans": trans, "torsions": tors,
            "torsion_mask": (torch.rand(B, L, 7, generator=g) > 0.3).float(),
            "mask": torch.ones(B, L), "seqres": torch.randint(0, 20, (B, L), generator=g)}


def flags_of(args):
    """the prep_batch-relevant flags as a plain dict (what the tests rebuild their Namespace from)"""
    keys = ("sim_condition", "tps_condition", "inpainting", "design", "dynamic_mpnn", "mpnn", "no_offsets", "no_torsion",
            "no_design_torsion", "design_key_frames", "supervise_all_torsions", "supervise_no_torsions", "cond_interval")
    return {k: getattr(args, k) for k in keys if getattr(args, k, None)}


def run_prep(wrapper, batch):
    out = wrapper.prep_batch({k: v.clone() for k, v in batch.items()})
    kw, rigids = out["model_kwargs"], out["rigids"]
    return {
        "latents": out["latents"].clone(), "loss_mask": out["loss_mask"].clone(),
        "start_rot": kw["start_frames"].get_rots().get_rot_mats().clone(), "start_trans": kw["start_frames"].get_trans().clone(),
        "end_rot": kw["end_frames"].get_rots().get_rot_mats().clone(), "end_trans": kw["end_frames"].get_trans().clone(),
        "rigids_rot": rigids.get_rots().get_rot_mats().clone(), "rigids_trans": rigids.get_trans().clone(),
        "mask": kw["mask"].clone(), "aatype": kw["aatype"].clone(), "x_cond": kw["x_cond"].clone(),
        "x_cond_mask": kw["x_cond_mask"].clone(),
    }


@torch.no_grad()
def main():
    ref = load_wrapper()
    out = {"cases": {}, "source": "NewMDGenWrapper.prep_batch (reference, unmodified)"}
    for name, (cfg_id, kw) in CASES.items():
        args, _ = wrapper_args(cfg_id, T=7, **kw)
        w = ref.NewMDGenWrapper(args).eval()
        batch = synthetic_batch(seed=len(name))
        out["cases"][name] = {"args": flags_of(args), "batch": batch, "out": run_prep(w, batch)}
        print(name, flags_of(args), tuple(out["cases"][name]["out"]["latents"].shape))
    path = os.path.join(ROOT, "tests", "golden", "prep_batch.pt")
    torch.save(out, path)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
