"""The five BASELINE.json configurations: model flags + seeded synthetic inputs.

The flag names are the reference's argparse names (src/mdgen/parsing.py:5-125); the
Namespace returned by `make_args` can be handed unchanged to the reference's
`LatentMDGenModel(args, latent_dim)` (latent_model.py:43-48) and to ours.
Synthetic inputs follow SURVEY.md section 8(d).
"""
import argparse
import math

import torch

# (B, T, L, D) per BASELINE.json config, plus task flags
CONFIGS = {
    1: dict(name="4AA forward-sim", B=1, T=1000, L=4, D=21, task="sim", crop=4, abs_pos_emb=True),
    2: dict(name="4AA TPS", B=16, T=100, L=4, D=28, task="tps", crop=4, abs_pos_emb=True),
    3: dict(name="4AA upsampling", B=8, T=1000, L=4, D=21, task="sim", crop=4, abs_pos_emb=True,
            cond_interval=100),
    4: dict(name="4AA inpainting/design", B=32, T=100, L=4, D=48, task="design", crop=4,
            abs_pos_emb=True),
    5: dict(name="ATLAS forward-sim", B=1, T=250, L=256, D=21, task="sim", crop=256,
            abs_pos_emb=False),
}


def make_args(cfg, **overrides):
    """argparse.Namespace with every flag LatentMDGenModel reads."""
    c = CONFIGS[cfg] if isinstance(cfg, int) else cfg
    a = argparse.Namespace(
        embed_dim=384, num_layers=5, mha_heads=16, ipa_heads=4, ipa_head_dim=32,
        ipa_qk=8, ipa_v=8, dropout=0.0, no_rope=False, hyena=False,
        num_frames=c["T"], crop=c["crop"], prepend_ipa=True, interleave_ipa=False,
        abs_pos_emb=c["abs_pos_emb"], abs_time_emb=False, no_aa_emb=False,
        time_multiplier=100.0, grad_checkpointing=False,
        sim_condition=False, tps_condition=False, inpainting=False, design=False,
        dynamic_mpnn=False, mpnn=False, alpha_max=8, dirichlet_flow_temp=1.0,
        allow_nan_cfactor=False, flow_matching=False, prior_variance=1.0,
        # transport flags read by Transport.__init__ (transport.py:67-83)
        path_type="GVP", prediction="velocity", sampling_method="euler",
        design_variance=1.0, weight_loss_var_x0=0.0, dirichlet_flow_matching=False, discrete_loss_weight=0.5,
    )
    task = c["task"]
    if task == "sim":
        a.sim_condition = True
    elif task == "tps":
        a.tps_condition = True
    elif task == "design":
        a.inpainting = True
        a.design = True
        a.no_aa_emb = True
    for k, v in overrides.items():
        setattr(a, k, v)
    return a


def latent_dim_of(cfg):
    return (CONFIGS[cfg] if isinstance(cfg, int) else cfg)["D"]


def small_variant(cfg, B=None, T=None, L=None):
    """Same task flags as config `cfg` at a reduced shape (parity tests / goldens)."""
    c = dict(CONFIGS[cfg])
    if B is not None:
        c["B"] = B
    if T is not None:
        c["T"] = T
    if L is not None:
        c["L"] = L
        c["crop"] = L
    if "cond_interval" in c:
        c["cond_interval"] = max(1, min(c["cond_interval"], c["T"] // 2))
    return c


def quat_to_rot(q):
    """[*,4] unit quaternion (w,x,y,z) -> [*,3,3].  Same convention as rigid_utils.py:168."""
    a, b, c, d = q.unbind(-1)
    rows = [
        [a * a + b * b - c * c - d * d, 2 * (b * c - a * d), 2 * (b * d + a * c)],
        [2 * (b * c + a * d), a * a - b * b + c * c - d * d, 2 * (c * d - a * b)],
        [2 * (b * d - a * c), 2 * (c * d + a * b), a * a - b * b - c * c + d * d],
    ]
    return torch.stack([torch.stack(r, -1) for r in rows], -2)


def _random_frames(g, B, L):
    q = torch.randn(B, L, 4, generator=g)
    q = q / q.norm(dim=-1, keepdim=True)
    q = torch.where(q[..., :1] < 0, -q, q)
    rot = quat_to_rot(q)
    step = torch.randn(B, L, 3, generator=g)
    step = 3.8 * step / step.norm(dim=-1, keepdim=True)
    trans = torch.cumsum(step, dim=1)
    return rot.contiguous(), trans.contiguous()


def synthetic_inputs(cfg, seed=0, pad_last=0):
    """Seeded synthetic inputs of SURVEY.md 8(d).  All CPU fp32 / int64 tensors.

    Returns dict: x0 [B,T,L,D], mask [B,T,L], start_rot/start_trans, (end_rot/end_trans),
    x_cond [B,T,L,Dc], x_cond_mask [B,T,L] i64, aatype [B,L] i64.
    """
    c = CONFIGS[cfg] if isinstance(cfg, int) else cfg
    idx = cfg if isinstance(cfg, int) else c.get("id", 9)
    B, T, L, D = c["B"], c["T"], c["L"], c["D"]
    task = c["task"]
    g = torch.Generator().manual_seed(1000 * idx + seed)
    out = {}
    x0 = torch.randn(B, T, L, D, generator=g)
    if task == "design":
        e = -torch.log(torch.rand(B, 1, L, 20, generator=g).clamp_min(1e-12))
        x0[..., -20:] = (e / e.sum(-1, keepdim=True)).expand(B, T, L, 20)
    out["x0"] = x0.contiguous()
    out["start_rot"], out["start_trans"] = _random_frames(g, B, L)
    if task in ("tps", "design"):
        out["end_rot"], out["end_trans"] = _random_frames(g, B, L)
    aatype = torch.randint(0, 20, (B, L), generator=g)
    mask = torch.ones(B, T, L)
    if pad_last:
        mask[:, :, L - pad_last:] = 0
    out["mask"] = mask
    cm = torch.zeros(B, T, L, dtype=torch.long)
    if task == "sim":
        ci = c.get("cond_interval")
        if ci:
            cm[:, ::ci] = 1
        else:
            cm[:, 0] = 1
    elif task == "tps":
        cm[:, 0] = 1
        cm[:, -1] = 1
    elif task == "design":
        cm[:, :, 0] = 1
        cm[:, :, L - 1] = 1
        if L > 2:
            aatype[:, 1:L - 1] = 20
    cm = cm * mask.long()
    out["x_cond_mask"] = cm
    out["aatype"] = aatype
    Dc = D - 20 if task == "design" else D
    xc = 0.5 * torch.randn(B, T, L, Dc, generator=g)
    n_off = 14 if task in ("tps",) else 7          # quaternion(4)+trans(3) per offset block
    for o in range(0, min(n_off, Dc), 7):
        qq = xc[..., o:o + 4]
        qq = qq / qq.norm(dim=-1, keepdim=True)
        xc[..., o:o + 4] = torch.where(qq[..., :1] < 0, -qq, qq)
    for o in range(min(n_off, Dc), Dc - 1, 2):      # torsion (sin, cos) pairs
        p = xc[..., o:o + 2]
        xc[..., o:o + 2] = p / p.norm(dim=-1, keepdim=True)
    out["x_cond"] = (xc * cm[..., None].float()).contiguous()
    return out


def flops_per_forward(cfg):
    """Algorithmic FLOPs of one velocity evaluation (BASELINE.md section 3 formula)."""
    c = CONFIGS[cfg] if isinstance(cfg, int) else cfg
    B, T, L, D = c["B"], c["T"], c["L"], c["D"]
    C, F = 384, 1536
    N = B * T * L
    task = c["task"]
    Dc = D - 20 if task == "design" else D
    f = 2 * N * C * (D + Dc)
    f += 5 * (32 * N * C * C + 4 * B * T * L * (L + 1) * C + 4 * B * L * T * (T + 1) * C)
    M = B * L * (2 if task in ("tps", "design") else 1)
    f += 5 * (2 * M * C * (128 + 256 + 96 + 192) + 2 * M * 256 * C
              + M * L * 4 * (2 * 32 + 3 * 8 * 3 + 2 * 32 + 2 * 3 * 8)
              + 8 * M * C * C + 4 * M * (L + 1) * C + 4 * M * C * F)
    streams = 2 if task in ("tps", "design") else 1
    f += 2 * B * (256 * C + C * C) + 5 * 2 * B * C * 9 * C + streams * 5 * 2 * B * C * 6 * C + 2 * B * C * 2 * C
    f += 2 * N * C * D
    if task == "design":
        f += 4 * N * C * C + 2 * B * L * C * C + 2 * B * L * C * 20
    return float(f)
