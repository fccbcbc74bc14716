"""TEST INFRASTRUCTURE ONLY -- golden vectors for the decode step (wrapper.py:487-508), row N2, and for the whole
`NewMDGenWrapper.inference` call chain (prep_batch -> 49-step Euler sampler -> decode, wrapper.py:436-514).

The goldens are produced by the reference's OWN `NewMDGenWrapper.inference(batch, zs0)` (unmodified, imported through
oracle/ref_loader.load_wrapper()).  The sampler's output (`samples`, the tensor the decode lines consume) is recorded
by wrapping `transport_sampler.sample_ode` -- the wrapper's code is not touched.  The four residue tables the decode
needs (mdgen.residue_constants) are stored in the fixture as TEST INPUTS: the product takes them as arguments from its
caller, exactly like geometry.frames_torsions_to_atom14 takes them from `rc`.
Run in the build container:  python -m oracle.make_decode_golden   ->  tests/golden/decode.pt"""
import importlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from mdgen_finetune_b200.weights import random_state_dict  # noqa: E402
from oracle.ref_loader import load_wrapper  # noqa: E402
from oracle.make_prep_golden import synthetic_batch, wrapper_args, flags_of  # noqa: E402

CASES = {"sim": (1, dict()), "tps": (2, dict()), "design": (4, dict(allow_nan_cfactor=True)),
         "no_offsets": (1, dict(no_offsets=True)), "oracle_torsions": (1, dict(oracle=True))}
B, T, L = 2, 5, 4


@torch.no_grad()
def main():
    ref = load_wrapper()
    rc = importlib.import_module("mdgen.residue_constants")
    tables = {"default_frames": torch.from_numpy(rc.restype_rigid_group_default_frame).float(),
              "group_idx": torch.from_numpy(rc.restype_atom14_to_rigid_group).long(),
              "atom_mask": torch.from_numpy(rc.restype_atom14_mask).float(),
              "lit_positions": torch.from_numpy(rc.restype_atom14_rigid_group_positions).float()}
    out = {"tables": tables, "cases": {}, "source": "NewMDGenWrapper.inference (reference, unmodified)"}
    for name, (cfg_id, kw) in CASES.items():
        args, c = wrapper_args(cfg_id, T=T, **kw)
        w = ref.NewMDGenWrapper(args).eval()
        # Q2: a freshly constructed model outputs exactly 0; re-draw the zero-initialised parameters
        w.model.load_state_dict(random_state_dict(args, w.latent_dim, seed=0), strict=True)
        batch = synthetic_batch(B=B, T=T, L=L, seed=10 + len(name))
        torch.manual_seed(len(name))
        zs0 = w.sample_prior_latent(B, T, L, device="cpu")
        rec = {}
        orig = w.transport_sampler.sample_ode

        def recording_sample_ode(**k):
            fn = orig(**k)

            def run(x, model, **mk):
                rec["traj"] = fn(x, model, **mk)
                return rec["traj"]
            return run
        w.transport_sampler.sample_ode = recording_sample_ode
        atom14, aa_out = w.inference({k: v.clone() for k, v in batch.items()}, zs0=zs0.clone())
        samples = rec["traj"][-1]
        assert torch.isfinite(atom14).all(), name
        out["cases"][name] = {"args": {**flags_of(args), **({"oracle": True} if args.oracle else {})}, "config": c,
                              "latent_dim": w.latent_dim, "batch": batch, "zs0": zs0, "samples": samples.clone(),
                              "rot0": batch["rots"][:, 0:1].clone(), "trans0": batch["trans"][:, 0:1].clone(),
                              "seqres": batch["seqres"], "atom14": atom14.clone(), "aa_out": aa_out.clone()}
        print(name, tuple(samples.shape), float(atom14.abs().max()))
    path = os.path.join(ROOT, "tests", "golden", "decode.pt")
    torch.save(out, path)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
