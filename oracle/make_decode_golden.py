"""TEST INFRASTRUCTURE ONLY -- golden vectors for the decode step (wrapper.py:487-508), row N2.

`mdgen.wrapper` cannot be imported (pytorch_lightning absent); the decode lines of NewMDGenWrapper.inference are
replayed on the reference's OWN `Rigid` class and `mdgen.geometry.frames_torsions_to_atom14`.  The four residue tables
the decode needs (mdgen.residue_constants: restype_rigid_group_default_frame, restype_atom14_to_rigid_group,
restype_atom14_mask, restype_atom14_rigid_group_positions) are stored in the fixture as TEST INPUTS: the product takes
them as arguments from its caller, exactly like geometry.frames_torsions_to_atom14 takes them from `rc`.
Run in the build container:  python -m oracle.make_decode_golden   ->  tests/golden/decode.pt"""
import argparse
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.ref_loader import load_reference  # noqa: E402
from oracle.make_prep_golden import synthetic_batch  # noqa: E402

CASES = {"sim": dict(D=21), "tps": dict(D=28, tps_condition=True), "design": dict(D=48, inpainting=True, design=True),
         "no_offsets": dict(D=21, no_offsets=True), "oracle_torsions": dict(D=21, oracle=True)}


def main():
    ref = load_reference()
    geometry = importlib.import_module("mdgen.geometry")
    rc = importlib.import_module("mdgen.residue_constants")
    Rigid, Rotation = ref.Rigid, ref.Rotation
    tables = {"default_frames": torch.from_numpy(rc.restype_rigid_group_default_frame).float(),
              "group_idx": torch.from_numpy(rc.restype_atom14_to_rigid_group).long(),
              "atom_mask": torch.from_numpy(rc.restype_atom14_mask).float(),
              "lit_positions": torch.from_numpy(rc.restype_atom14_rigid_group_positions).float()}
    out = {"tables": tables, "cases": {}}
    for name, kw in CASES.items():
        kw = dict(kw)
        D = kw.pop("D")
        args = argparse.Namespace(tps_condition=False, inpainting=False, design=False, no_offsets=False, oracle=False, no_frames=False)
        for k, v in kw.items():
            setattr(args, k, v)
        batch = synthetic_batch(B=2, T=5, L=4, seed=10 + len(name))
        g = torch.Generator().manual_seed(len(name))
        samples = torch.randn(2, 5, 4, D, generator=g)
        samples[..., 4:7] *= 3.0
        rigids = Rigid(trans=batch["trans"], rots=Rotation(rot_mats=batch["rots"]))
        B, T, L = rigids.shape
        # ---- wrapper.py:487-508
        offsets = samples[..., :7]
        if args.tps_condition or args.inpainting:
            torsions = samples[..., 14:28]
            logits = samples[..., -20:]
        else:
            torsions = samples[..., 7:21]
            logits = samples[..., -20:]
        if args.no_offsets:
            frames = Rigid.from_tensor_7(offsets, normalize_quats=True)
        else:
            frames = rigids[:, 0:1].compose(Rigid.from_tensor_7(offsets, normalize_quats=True))
        if args.design:
            trans = frames.get_trans()
            rots = frames.get_rots().get_rot_mats()
            frames = Rigid(trans=trans, rots=Rotation(rot_mats=rots))
        torsions = torsions.reshape(B, T, L, 7, 2)
        if not args.oracle:
            torsions = torsions / torch.linalg.norm(torsions, dim=-1, keepdims=True)
        atom14 = geometry.frames_torsions_to_atom14(frames, torsions.view(B, T, L, 7, 2), batch["seqres"][:, None].expand(B, T, L))
        aa_out = torch.argmax(logits, -1) if args.design else batch["seqres"][:, None].expand(B, T, L)
        out["cases"][name] = {"args": kw, "samples": samples, "rot0": batch["rots"][:, 0:1].clone(), "trans0": batch["trans"][:, 0:1].clone(),
                              "seqres": batch["seqres"], "atom14": atom14.clone(), "aa_out": aa_out.clone()}
    path = os.path.join(ROOT, "tests", "golden", "decode.pt")
    torch.save(out, path)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
