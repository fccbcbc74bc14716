"""prep_batch -- drop-in for NewMDGenWrapper.prep_batch (src/mdgen/wrapper.py:283-365), frames path: trajectory frames
and torsions -> the sampler's latents, conditioning tensors and loss mask.  The arithmetic (relative rigid transforms,
rotation -> quaternion with real part >= 0, latent / x_cond / x_cond_mask assembly) is one CUDA kernel behind
mdgen_prep_batch; what stays here is tensor bookkeeping (mask broadcasting, index remapping of the key frames)."""
import ctypes

import torch

from . import _lib

DESIGN_IDX = [1, 2]                 # wrapper.py:41-43
COND_IDX = [0, 3]
DESIGN_MAP_TO_COND = [0, 0, 3, 3]

_TWO_REF, _NO_OFFSETS, _NO_TORSION, _NO_DESIGN_TORSION, _KEY_FRAMES, _SIM, _TPS, _INPAINT = (1 << i for i in range(8))


def prep_batch(args, batch):
    """batch: 'trans' [B,T,L,3], 'rots' [B,T,L,3,3], 'torsions' [B,T,L,7,2], 'torsion_mask' [B,L,7], 'mask' [B,L],
    'seqres' [B,L] on a CUDA device.  Returns the reference's dict; 'rigids', 'start_frames' and 'end_frames' are
    (rot_mats, trans) tuples, which LatentMDGenModel accepts in place of Rigid objects."""
    g = lambda name: bool(getattr(args, name, False))
    if g("no_frames") or "latents" in batch:
        raise NotImplementedError("--no_frames / hyena latents are outside the B200 hot path (DESIGN.md, out of scope)")
    trans, rots = batch["trans"], batch["rots"]
    if not trans.is_cuda:
        raise RuntimeError("mdgen_finetune_b200.prep_batch needs CUDA tensors: there is no CPU fallback")
    B, T, L = trans.shape[:3]
    dev = trans.device
    f32 = lambda v: v.detach().to(dev, torch.float32).contiguous()
    trans, rots = f32(trans), f32(rots)
    tors = f32(batch["torsions"]).reshape(B, T, L, 14)
    two = g("tps_condition") or g("inpainting") or g("dynamic_mpnn")
    flags = ((_TWO_REF if two else 0) | (_NO_OFFSETS if g("no_offsets") else 0) | (_NO_TORSION if g("no_torsion") else 0) |
             (_NO_DESIGN_TORSION if g("no_design_torsion") and not g("no_torsion") else 0) |
             (_KEY_FRAMES if g("design_key_frames") else 0) | (_SIM if g("sim_condition") else 0) |
             (_TPS if g("tps_condition") else 0) | (_INPAINT if g("inpainting") or g("dynamic_mpnn") or g("mpnn") else 0))
    D = (14 if two else 7) + 14
    latents = torch.empty(B, T, L, D, device=dev)
    x_cond = torch.empty_like(latents)
    cond_mask = torch.empty(B, T, L, dtype=torch.int64, device=dev)
    p = lambda v: ctypes.c_void_p(v.data_ptr())
    with torch.cuda.device(dev):
        st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        _lib.check(_lib.lib().mdgen_prep_batch(B, T, L, p(trans), p(rots), p(tors), flags, int(getattr(args, "cond_interval", 0) or 0),
                                               p(latents), p(x_cond), p(cond_mask), st), "mdgen_prep_batch")
    if g("design_key_frames"):                                                        # :296-302
        idx = torch.tensor(DESIGN_MAP_TO_COND, device=dev)
        rots = torch.cat([rots[:, :1, idx], rots[:, 1:-1], rots[:, -1:, idx]], 1)
        trans = torch.cat([trans[:, :1, idx], trans[:, 1:-1], trans[:, -1:, idx]], 1)
    mask = batch["mask"].to(dev)
    frame_loss_mask = mask.unsqueeze(-1).expand(-1, -1, 14 if two else 7)             # :311, 314-318
    torsion_loss_mask = batch["torsion_mask"].to(dev).unsqueeze(-1).expand(-1, -1, -1, 2).reshape(B, L, 14)
    if g("supervise_all_torsions"):                                                   # :329-332
        torsion_loss_mask = torch.ones_like(torsion_loss_mask)
    elif g("supervise_no_torsions"):
        torsion_loss_mask = torch.zeros_like(torsion_loss_mask)
    loss_mask = torch.cat([frame_loss_mask, torsion_loss_mask], -1).unsqueeze(1).expand(-1, T, -1, -1)
    seqres = batch["seqres"].to(dev)
    aatype_mask = torch.ones_like(seqres)
    if g("design"):
        aatype_mask[:, DESIGN_IDX] = 0
    return {
        "rigids": (rots, trans),
        "latents": latents,
        "loss_mask": loss_mask,
        "model_kwargs": {
            "start_frames": (rots[:, 0], trans[:, 0]),
            "end_frames": (rots[:, -1], trans[:, -1]),
            "mask": mask.unsqueeze(1).expand(-1, T, -1),
            "aatype": torch.where(aatype_mask.bool(), seqres, 20),
            "x_cond": x_cond,
            "x_cond_mask": cond_mask,
        },
    }


def decode_samples(args, samples, first_frames, seqres, tables):
    """The decode lines of NewMDGenWrapper.inference (wrapper.py:487-508): samples [B,T,L,D] (the sampler's last state),
    first_frames = rigids[:, 0:1] as (rot_mats [B,1,L,3,3] or [B,L,3,3], trans), seqres [B,L] ->
    (atom14 [B,T,L,14,3], aa_out [B,T,L]).  `tables`: dict with the caller's mdgen.residue_constants arrays
    'default_frames' (restype_rigid_group_default_frame), 'group_idx' (restype_atom14_to_rigid_group), 'atom_mask'
    (restype_atom14_mask), 'lit_positions' (restype_atom14_rigid_group_positions) as tensors."""
    g = lambda name: bool(getattr(args, name, False))
    if g("no_frames"):
        raise NotImplementedError("--no_frames is outside the B200 hot path (DESIGN.md, out of scope)")
    if not samples.is_cuda:
        raise RuntimeError("mdgen_finetune_b200.decode_samples needs CUDA tensors: there is no CPU fallback")
    dev = samples.device
    B, T, L, D = samples.shape
    f32 = lambda v: v.detach().to(dev, torch.float32).contiguous()
    i64 = lambda v: v.detach().to(dev, torch.int64).contiguous()
    xs = f32(samples)
    no_off = g("no_offsets")
    rot0 = trans0 = None
    if not no_off:
        rot0, trans0 = f32(first_frames[0]).reshape(B, L, 3, 3), f32(first_frames[1]).reshape(B, L, 3)
    tb = {"default_frames": f32(tables["default_frames"]), "group_idx": i64(tables["group_idx"]),
          "atom_mask": f32(tables["atom_mask"]), "lit_positions": f32(tables["lit_positions"])}
    atom14 = torch.empty(B, T, L, 14, 3, device=dev)
    aa_out = torch.empty(B, T, L, dtype=torch.int64, device=dev)
    flags = (1 if no_off else 0) | (2 if g("oracle") else 0) | (4 if g("design") else 0)
    tors_off = 14 if (g("tps_condition") or g("inpainting")) else 7
    p = lambda v: None if v is None else ctypes.c_void_p(v.data_ptr())
    seq = i64(seqres)
    with torch.cuda.device(dev):
        st = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        _lib.check(_lib.lib().mdgen_decode_atom14(B, T, L, D, p(xs), tors_off, p(rot0), p(trans0), p(seq), p(tb["default_frames"]),
                                                  p(tb["group_idx"]), p(tb["atom_mask"]), p(tb["lit_positions"]), flags, p(atom14),
                                                  p(aa_out), st), "mdgen_decode_atom14")
    return atom14, aa_out
