"""TEST INFRASTRUCTURE ONLY -- golden vectors for prep_batch (wrapper.py:283-365), row N1.

`mdgen.wrapper` itself cannot be imported here (pytorch_lightning is absent), so the body of
NewMDGenWrapper.prep_batch is replayed line by line on the reference's OWN classes and helpers
(`mdgen.rigid_utils.Rigid / Rotation`, `mdgen.utils.get_offsets`, the module constants of wrapper.py:41-43):
everything numeric (invert, compose, rot_to_quat through eigh, to_tensor_7) is the reference's code.
Run in the build container:  python -m oracle.make_prep_golden   ->  tests/golden/prep_batch.pt"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.ref_loader import load_reference  # noqa: E402

DESIGN_IDX = [1, 2]                 # wrapper.py:41-43
COND_IDX = [0, 3]
DESIGN_MAP_TO_COND = [0, 0, 3, 3]

CASES = {
    "sim": dict(sim_condition=True),
    "upsample": dict(sim_condition=True, cond_interval=3),
    "tps": dict(tps_condition=True),
    "design": dict(inpainting=True, design=True, no_torsion=True),
    "design_key_frames": dict(inpainting=True, design=True, design_key_frames=True, no_design_torsion=True),
    "no_offsets": dict(sim_condition=True, no_offsets=True, supervise_all_torsions=True),
}
FLAGS = ("sim_condition", "tps_condition", "inpainting", "design", "dynamic_mpnn", "mpnn", "no_offsets", "no_torsion",
         "no_design_torsion", "design_key_frames", "supervise_all_torsions", "supervise_no_torsions", "no_frames")


def make_args(**kw):
    a = argparse.Namespace(**{f: False for f in FLAGS}, cond_interval=None)
    for k, v in kw.items():
        setattr(a, k, v)
    return a


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


def reference_prep_batch(ref, utils, args, batch):
    """wrapper.py:283-365 replayed on the reference's classes (frames path)."""
    Rigid, Rotation = ref.Rigid, ref.Rotation
    rigids = Rigid(trans=batch["trans"], rots=Rotation(rot_mats=batch["rots"]))
    B, T, L = rigids.shape
    if args.design_key_frames:
        rigids = Rigid.cat([rigids[:, :1, DESIGN_MAP_TO_COND], rigids[:, 1:-1], rigids[:, -1:, DESIGN_MAP_TO_COND]], 1)
    if args.no_offsets:
        offsets = rigids.to_tensor_7()
    else:
        offsets = utils.get_offsets(rigids[:, 0:1], rigids)
    offsets[..., :4] *= torch.where(offsets[:, :, :, 0:1] < 0, -1, 1)
    frame_loss_mask = batch["mask"].unsqueeze(-1).expand(-1, -1, 7)
    torsion_loss_mask = batch["torsion_mask"].unsqueeze(-1).expand(-1, -1, -1, 2).reshape(B, L, 14)
    if args.tps_condition or args.inpainting or args.dynamic_mpnn:
        offsets_r = utils.get_offsets(rigids[:, -1:], rigids)
        offsets_r[..., :4] *= torch.where(offsets_r[:, :, :, 0:1] < 0, -1, 1)
        offsets = torch.cat([offsets, offsets_r], -1)
        frame_loss_mask = torch.cat([frame_loss_mask, frame_loss_mask], -1)
    if args.no_torsion:
        latents = torch.cat([offsets, torch.zeros_like(batch["torsions"].view(B, T, L, 14))], -1)
    elif args.no_design_torsion:
        torsions_ = batch["torsions"].clone()
        torsions_[:, :, DESIGN_IDX] = 0
        latents = torch.cat([offsets, torsions_.view(B, T, L, 14)], -1)
    else:
        latents = torch.cat([offsets, batch["torsions"].view(B, T, L, 14)], -1)
    if args.supervise_all_torsions:
        torsion_loss_mask = torch.ones_like(torsion_loss_mask)
    elif args.supervise_no_torsions:
        torsion_loss_mask = torch.zeros_like(torsion_loss_mask)
    loss_mask = torch.cat([frame_loss_mask, torsion_loss_mask], -1)
    loss_mask = loss_mask.unsqueeze(1).expand(-1, T, -1, -1)
    cond_mask = torch.zeros(B, T, L, dtype=int, device=offsets.device)
    if args.sim_condition:
        cond_mask[:, 0] = 1
    if args.tps_condition:
        cond_mask[:, 0] = cond_mask[:, -1] = 1
    if args.cond_interval:
        cond_mask[:, ::args.cond_interval] = 1
    if args.inpainting or args.dynamic_mpnn or args.mpnn:
        cond_mask[:, :, COND_IDX] = 1
    aatype_mask = torch.ones_like(batch["seqres"])
    if args.design:
        aatype_mask[:, DESIGN_IDX] = 0
    return {
        "latents": latents.clone(), "loss_mask": loss_mask.clone(),
        "start_rot": rigids[:, 0].get_rots().get_rot_mats().clone(), "start_trans": rigids[:, 0].get_trans().clone(),
        "end_rot": rigids[:, -1].get_rots().get_rot_mats().clone(), "end_trans": rigids[:, -1].get_trans().clone(),
        "mask": batch["mask"].unsqueeze(1).expand(-1, T, -1).clone(),
        "aatype": torch.where(aatype_mask.bool(), batch["seqres"], 20),
        "x_cond": torch.where(cond_mask.unsqueeze(-1).bool(), latents, 0.0), "x_cond_mask": cond_mask,
    }


def main():
    ref = load_reference()
    import importlib
    utils = importlib.import_module("mdgen.utils")
    out = {"cases": {}}
    for name, kw in CASES.items():
        args = make_args(**kw)
        batch = synthetic_batch(seed=len(name))
        out["cases"][name] = {"args": kw, "batch": batch, "out": reference_prep_batch(ref, utils, args, batch)}
    path = os.path.join(ROOT, "tests", "golden", "prep_batch.pt")
    torch.save(out, path)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
