"""TEST INFRASTRUCTURE ONLY -- CPU fp32 restatement of the reference's denoising hot path.

This file is the parity checker.  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import it; the product package
(mdgen_finetune_b200/) never does.

It restates, as plain functional PyTorch on CPU tensors driven by a reference-layout
state_dict, what /root/reference/src computes on the path SURVEY.md section 8(a) names.
Each function cites the reference file:line it follows.  PARITY PIN: the restatement is
checked (tests/test_oracle_golden.py) against golden vectors produced by running the
UNMODIFIED reference modules in the build container (oracle/make_golden.py ->
tests/golden/*.pt).  The two third-party arithmetic pieces (fair-esm RoPE, torchdiffeq
Euler) have no tests in the reference: at those two boundaries parity is "unpinned" and
rests on the published algorithms restated in `rope_tables` / `euler_sample`.

Frames are passed as (rot [B,L,3,3], trans [B,L,3]) tuples instead of Rigid objects.
"""
import math

import numpy as np
import torch
import torch.nn.functional as F


# ------------------------------------------------------------------------------------
# small helpers
def _lin(sd, p, x):
    """nn.Linear / primitives.Linear forward (primitives.py:194-206)."""
    return F.linear(x, sd[p + ".weight"], sd.get(p + ".bias"))


def gelu_erf(x):
    """layers.py:78-85."""
    return x * 0.5 * (1.0 + torch.erf(x / math.sqrt(2.0)))


def modulate(x, shift, scale):
    """layers.py:14-15 (broadcast over dim 1)."""
    return x * (1 + scale.unsqueeze(1)) + shift.unsqueeze(1)


def ln_plain(x):
    """nn.LayerNorm(C, elementwise_affine=False, eps=1e-6) (latent_model.py:371,376,449,454)."""
    return F.layer_norm(x, x.shape[-1:], None, None, 1e-6)


def timestep_embedding(t, dim=256, max_period=10000):
    """layers.py:30-50: [cos | sin]."""
    half = dim // 2
    freqs = torch.exp(-math.log(max_period) * torch.arange(half, dtype=torch.float32) / half)
    a = t[:, None].float() * freqs[None]
    return torch.cat([torch.cos(a), torch.sin(a)], -1)


def t_embedder(sd, t):
    """layers.py:52-55: Linear(256,C) -> SiLU -> Linear(C,C)."""
    h = _lin(sd, "t_embedder.mlp.0", timestep_embedding(t))
    return _lin(sd, "t_embedder.mlp.2", F.silu(h))


# ------------------------------------------------------------------------------------
# fair-esm rotary embedding (third-party; call sites mha.py:129-130,356-357)
def rope_tables(n, dim):
    inv_freq = 1.0 / (10000 ** (torch.arange(0, dim, 2).float() / dim))
    t = torch.arange(n).float()
    fr = torch.einsum("i,j->ij", t, inv_freq)
    emb = torch.cat((fr, fr), -1)
    return emb.cos(), emb.sin()


def _rot_half(x):
    x1, x2 = x.chunk(2, -1)
    return torch.cat((-x2, x1), -1)


def mha(sd, p, x, mask, heads, rope=True):
    """AttentionWithRoPE + MultiheadAttention as reached from the model.

    latent_model.py:334-338, mha.py:258-286,356-397.  x [Nb,S,C] batch-first,
    mask [Nb,S] (1 = valid).  Returns [Nb,S,C].
    """
    Nb, S, C = x.shape
    d = C // heads
    q = _lin(sd, p + ".attn.q_proj", x) * d ** -0.5                     # mha.py:260-263
    k = _lin(sd, p + ".attn.k_proj", x)
    v = _lin(sd, p + ".attn.v_proj", x)
    bk = sd[p + ".attn.bias_k"].reshape(1, 1, C).expand(Nb, 1, C)       # mha.py:265-268
    bv = sd[p + ".attn.bias_v"].reshape(1, 1, C).expand(Nb, 1, C)
    k = torch.cat([k, bk], 1)
    v = torch.cat([v, bv], 1)
    pad = torch.cat([1 - mask, mask.new_zeros(Nb, 1)], 1) != 0          # mha.py:273-280, 374
    q = q.view(Nb, S, heads, d).transpose(1, 2)                         # [Nb,H,S,d]
    k = k.view(Nb, S + 1, heads, d).transpose(1, 2)
    v = v.view(Nb, S + 1, heads, d).transpose(1, 2)
    if rope:                                                            # mha.py:356-357
        cos, sin = rope_tables(S + 1, d)
        q = q * cos[:S] + _rot_half(q) * sin[:S]
        k = k * cos + _rot_half(k) * sin
    a = q @ k.transpose(-1, -2)                                         # mha.py:359
    a = a.masked_fill(pad[:, None, None, :], float("-inf"))             # mha.py:370-376
    a = F.softmax(a, -1, dtype=torch.float32)                           # mha.py:381
    o = (a @ v).transpose(1, 2).reshape(Nb, S, C)                       # mha.py:389-396
    return _lin(sd, p + ".attn.out_proj", o)


# ------------------------------------------------------------------------------------
# rigid helpers (rigid_utils.py)
def rot_apply(rot, pts):
    """rot_vec_mul (rigid_utils.py:64-86); rot [*,3,3], pts [*,3]."""
    return (rot * pts[..., None, :]).sum(-1)


def rigid_invert(rot, trans):
    """Rigid.invert (rigid_utils.py:1075-1085)."""
    ri = rot.transpose(-1, -2)
    return ri, -rot_apply(ri, trans)


def rigid_compose(r1, t1, r2, t2):
    """Rigid.compose (rigid_utils.py:1031-1045)."""
    return r1 @ r2, rot_apply(r1, t2) + t1


def rot_to_quat(rot):
    """rigid_utils.py:191-215: top eigenvector of the 4x4 K matrix (torch.linalg.eigh)."""
    xx, xy, xz = rot[..., 0, 0], rot[..., 0, 1], rot[..., 0, 2]
    yx, yy, yz = rot[..., 1, 0], rot[..., 1, 1], rot[..., 1, 2]
    zx, zy, zz = rot[..., 2, 0], rot[..., 2, 1], rot[..., 2, 2]
    k = [[xx + yy + zz, zy - yz, xz - zx, yx - xy],
         [zy - yz, xx - yy - zz, xy + yx, xz + zx],
         [xz - zx, xy + yx, yy - xx - zz, yz + zy],
         [yx - xy, xz + zx, yz + zy, zz - xx - yy]]
    k = (1.0 / 3.0) * torch.stack([torch.stack(r, -1) for r in k], -2)
    _, vec = torch.linalg.eigh(k)
    return vec[..., -1]


def to_tensor_7(rot, trans):
    """Rigid.to_tensor_7 (rigid_utils.py:1143-1155)."""
    return torch.cat([rot_to_quat(rot), trans], -1)


def relative_7(start, end):
    """latent_model.py:198-199: start^-1 o end and end^-1 o start as 7-vectors."""
    si = rigid_invert(*start)
    ei = rigid_invert(*end)
    x_f = to_tensor_7(*rigid_compose(*si, *end))
    x_r = to_tensor_7(*rigid_compose(*ei, *start))
    return x_f, x_r


# ------------------------------------------------------------------------------------
# invariant point attention, c_z = 0  (ipa.py:92-255)
def ipa(sd, p, s, rot, trans, mask, H=4, c=32, Pq=8, Pv=8, inf=1e5, eps=1e-8):
    """s [B,L,C]; rot [B,L,3,3]; trans [B,L,3]; mask [B,L] -> [B,L,C]."""
    B, L, _ = s.shape
    q = _lin(sd, p + ".linear_q", s).view(B, L, H, c)                    # ipa.py:113-117
    kv = _lin(sd, p + ".linear_kv", s).view(B, L, H, 2 * c)              # ipa.py:114-123
    k, v = kv[..., :c], kv[..., c:]

    def points(name, n):
        x = _lin(sd, p + name, s)                                        # [B,L,3*n], laid out (xyz, n)
        x = x.view(B, L, 3, n).transpose(-1, -2)                         # ipa.py:130-131 -> [B,L,n,3]
        return rot_apply(rot[:, :, None], x) + trans[:, :, None]         # ipa.py:132 Rigid.apply

    q_pts = points(".linear_q_points", H * Pq).view(B, L, H, Pq, 3)      # ipa.py:126-135
    kv_pts = points(".linear_kv_points", H * (Pq + Pv)).view(B, L, H, Pq + Pv, 3)
    k_pts, v_pts = kv_pts[..., :Pq, :], kv_pts[..., Pq:, :]              # ipa.py:149-151

    a = torch.einsum("bihc,bjhc->bhij", q, k) * math.sqrt(1.0 / (3 * c))  # ipa.py:161-166
    d2 = ((q_pts[:, :, None] - k_pts[:, None]) ** 2).sum(-1)             # [B,i,j,H,Pq]  ipa.py:171-175
    hw = F.softplus(sd[p + ".head_weights"]) * math.sqrt(1.0 / (3 * (Pq * 9.0 / 2)))  # ipa.py:176-181
    pt = (d2 * hw.view(1, 1, 1, H, 1)).sum(-1) * (-0.5)                  # ipa.py:182-185
    a = a + pt.permute(0, 3, 1, 2)                                       # ipa.py:193-196
    sq = inf * (mask[:, :, None] * mask[:, None, :] - 1)                 # ipa.py:188-190
    a = a + sq[:, None]                                                  # ipa.py:198
    a = torch.softmax(a, -1)                                             # ipa.py:203

    o = torch.einsum("bhij,bjhc->bihc", a, v).reshape(B, L, H * c)       # ipa.py:209-212
    o_pt = torch.einsum("bhij,bjhpx->bihpx", a, v_pts)                   # ipa.py:216-225
    o_pt = o_pt - trans[:, :, None, None]                                # ipa.py:226 invert_apply
    o_pt = rot_apply(rot.transpose(-1, -2)[:, :, None, None], o_pt)
    o_norm = torch.sqrt((o_pt ** 2).sum(-1) + eps).reshape(B, L, H * Pv)  # ipa.py:229-231
    o_pt = o_pt.reshape(B, L, H * Pv, 3)                                 # ipa.py:234
    cat = torch.cat([o, o_pt[..., 0], o_pt[..., 1], o_pt[..., 2], o_norm], -1)   # ipa.py:250-251
    return _lin(sd, p + ".linear_out", cat)


def ipa_layer(sd, p, s, t, mask, rot, trans, heads=16):
    """IPALayer.forward (latent_model.py:378-394).  s [B,L,C], t [B,C], mask [B,L]."""
    mod = _lin(sd, p + ".adaLN_modulation.1", F.silu(t))
    sh_a, sc_a, g_a, sh_m, sc_m, g_m = mod.chunk(6, -1)
    C = s.shape[-1]
    n = F.layer_norm(s, (C,), sd[p + ".ipa_norm.weight"], sd[p + ".ipa_norm.bias"], 1e-5)
    s = s + ipa(sd, p + ".ipa", n, rot, trans, mask)
    h = modulate(ln_plain(s), sh_a, sc_a)
    s = s + g_a.unsqueeze(1) * mha(sd, p + ".mha_l", h, mask, heads)
    h = modulate(ln_plain(s), sh_m, sc_m)
    s = s + g_m.unsqueeze(1) * _lin(sd, p + ".fc2", gelu_erf(_lin(sd, p + ".fc1", h)))
    return s


def main_layer(sd, p, x, t, mask, heads=16):
    """LatentMDGenLayer.forward (latent_model.py:456-493).  x [B,T,L,C], t [B,1,C], mask [B,T,L]."""
    B, T, L, C = x.shape
    mod = _lin(sd, p + ".adaLN_modulation.1", F.silu(t))
    sh_l, sc_l, g_l, sh_t, sc_t, g_t, sh_m, sc_m, g_m = mod.chunk(9, -1)
    h = modulate(ln_plain(x), sh_l, sc_l)
    y = mha(sd, p + ".mha_l", h.reshape(B * T, L, C), mask.reshape(B * T, L), heads)
    x = x + g_l.unsqueeze(1) * y.reshape(B, T, L, C)
    h = modulate(ln_plain(x), sh_t, sc_t)
    y = mha(sd, p + ".mha_t", h.transpose(1, 2).reshape(B * L, T, C),
            mask.transpose(1, 2).reshape(B * L, T), heads)
    x = x + g_t.unsqueeze(1) * y.reshape(B, L, T, C).transpose(1, 2)
    h = modulate(ln_plain(x), sh_m, sc_m)
    x = x + g_m.unsqueeze(1) * _lin(sd, p + ".fc2", gelu_erf(_lin(sd, p + ".fc1", h)))
    return x


def run_ipa(sd, args, t, mask, start, end, aatype, x_d=None):
    """LatentMDGenModel.run_ipa (latent_model.py:179-214); t [B,C], mask [B,L] (Q1 semantics)."""
    B, L = mask.shape
    n = args.num_layers
    if args.sim_condition or args.mpnn:
        s = torch.zeros(B, L, args.embed_dim)
        if aatype is not None and not args.no_aa_emb:
            s = s + sd["aatype_to_emb.weight"][aatype]
        if args.design:
            s = s + _lin(sd, "x_d_to_emb", x_d)
        for i in range(n):
            s = ipa_layer(sd, f"ipa_layers.{i}", s, t, mask, *start, heads=args.mha_heads)
        return s
    x_f7, x_r7 = relative_7(start, end)
    x_f = _lin(sd, "latent_to_emb_f", x_f7)
    x_r = _lin(sd, "latent_to_emb_r", x_r7)
    if aatype is not None and not args.no_aa_emb:
        e = sd["aatype_to_emb.weight"][aatype]
        x_f, x_r = x_f + e, x_r + e
    if args.design:
        e = _lin(sd, "x_d_to_emb", x_d)
        x_f, x_r = x_f + e, x_r + e
    for i in range(n):
        x_r = ipa_layer(sd, f"ipa_layers.{i}", x_r, t, mask, *start, heads=args.mha_heads)
        x_f = ipa_layer(sd, f"ipa_layers.{i}", x_f, t, mask, *end, heads=args.mha_heads)
    return x_r + x_f


def forward(sd, args, x, t, mask, start=None, end=None, x_cond=None, x_cond_mask=None, aatype=None):
    """LatentMDGenModel.forward (latent_model.py:216-269), Q1 = mask[:, 0] for run_ipa."""
    if mask.dim() == 2:
        mask = mask.repeat(x.shape[0], 1, 1)                              # latent_model.py:233-234
    x_d = x[..., -20:].mean(1) if args.design else None                   # :237-240
    h = _lin(sd, "latent_to_emb", x)                                      # :242
    if args.abs_pos_emb:
        h = h + sd["pos_embed"]                                           # :243-244
    if getattr(args, "abs_time_emb", False):
        h = h + sd["time_embed"][:, :, None]
    if x_cond is not None:
        h = h + _lin(sd, "cond_to_emb", x_cond) + sd["mask_to_emb.weight"][x_cond_mask]   # :249-250
    te = t_embedder(sd, t.float() * args.time_multiplier)                 # :252
    if args.prepend_ipa:
        h = h + run_ipa(sd, args, te, mask[:, 0], start, end, aatype, x_d)[:, None]       # :254-255
    for i in range(args.num_layers):
        h = main_layer(sd, f"layers.{i}", h, te[:, None], mask, args.mha_heads)           # :257-258
    mod = _lin(sd, "emb_to_latent.adaLN_modulation.1", F.silu(te[:, None]))               # layers.py:71-75
    shift, scale = mod.chunk(2, -1)
    latent = _lin(sd, "emb_to_latent.linear", modulate(ln_plain(h), shift, scale))
    if args.design:                                                       # :262-268
        xl = _lin(sd, "fc2", gelu_erf(_lin(sd, "fc1", h))).mean(1)
        logits = _lin(sd, "emb_to_logits", gelu_erf(_lin(sd, "fc3", xl)))
        latent = torch.cat([latent[..., :-20], latent[..., -20:] + logits[:, None]], -1)
    return latent


# ------------------------------------------------------------------------------------
# design-mode Dirichlet conditional flow (utils.py:33-57, latent_model.py:279-326)
class DirichletFlowTables:
    """DirichletConditionalFlow (utils.py:33-43) with the model's constructor arguments
    (latent_model.py:125-126: K=20, alpha_spacing=0.001, alpha_max=args.alpha_max)."""

    def __init__(self, K=20, alpha_min=1, alpha_max=8, alpha_spacing=0.001):
        import scipy.special
        self.alphas = np.arange(alpha_min, alpha_max + alpha_spacing, alpha_spacing)
        self.bs = np.linspace(0, 1, 1000)
        cdfs = np.array([scipy.special.betainc(a, K - 1, self.bs) for a in self.alphas])
        self.deriv = np.diff(cdfs, axis=0) / alpha_spacing
        self.alpha_spacing = alpha_spacing
        self.K = K

    def c_factor(self, bs, alpha):
        """utils.py:45-57."""
        import scipy.special
        beta = scipy.special.beta(alpha, self.K - 1)
        with np.errstate(all="ignore"):
            beta_div = np.where(bs < 1, beta / ((1 - bs) ** (self.K - 1)), 0)
            beta_div_full = np.where((bs ** (alpha - 1)) > 0, beta_div / (bs ** (alpha - 1)), 0)
        I_func = self.deriv[np.argmin(np.abs(alpha - self.alphas))]
        return -np.interp(bs, self.bs, I_func) * beta_div_full


_tables = {}


def forward_inference(sd, args, x, t, mask, start=None, end=None, x_cond=None, x_cond_mask=None,
                      aatype=None):
    """LatentMDGenModel.forward_inference (latent_model.py:272-326)."""
    if not args.design or args.dynamic_mpnn or args.mpnn:
        return forward(sd, args, x, t, mask, start, end, x_cond, x_cond_mask, aatype)
    xd = x[..., -20:]
    latent = forward(sd, args, x, t, mask, start, end, x_cond, x_cond_mask, aatype)
    probs = F.softmax(latent[..., -20:] / args.dirichlet_flow_temp, -1)        # :296
    alpha = float(1 * (1 - t[0]) + t[0] * args.alpha_max)                      # transport.py:52-57
    dalpha = args.alpha_max - 1
    key = float(args.alpha_max)
    if key not in _tables:
        _tables[key] = DirichletFlowTables(alpha_max=args.alpha_max)
    tab = _tables[key]
    if alpha > args.alpha_max:                                                 # :308-309
        alpha = args.alpha_max - tab.alpha_spacing
    cf = torch.from_numpy(tab.c_factor(xd.numpy(), alpha)).to(xd)              # :310-311
    if torch.isnan(cf).any():
        if args.allow_nan_cfactor:
            cf = torch.nan_to_num(cf)
        else:
            raise RuntimeError("NAN cfactor")
    eye = torch.eye(20)
    cond_flows = (eye - xd.unsqueeze(-1)) * cf.unsqueeze(-2)                   # :322-323
    flow = (probs.unsqueeze(-2) * cond_flows).sum(-1) * dalpha                 # :324
    return torch.cat([latent[..., :-20], flow], -1)


# ------------------------------------------------------------------------------------
# sampler: Sampler.sample_ode(sampling_method='euler') -> ode.sample -> torchdiffeq Euler
def euler_sample(velocity, x0, num_steps=50, return_all=False):
    """transport.py:407-450 + integrators.py:74-114 + torchdiffeq fixed-grid Euler.

    grid = linspace(0,1,num_steps) (fp32); per step t_vec = ones(B)*t_i,
    y_{i+1} = y_i + (t_{i+1}-t_i) * f(t_i, y_i).  Q3: 49 evaluations, the last at t=48/49.
    """
    grid = torch.linspace(0, 1, num_steps)
    y = x0
    ys = [y]
    for i in range(num_steps - 1):
        tv = torch.ones(x0.shape[0]) * grid[i]
        y = y + (grid[i + 1] - grid[i]) * velocity(y, tv)
        if return_all:
            ys.append(y)
    return torch.stack(ys, 0) if return_all else y


def rerandomize_zero_params(sd, seed=0, std=0.02):
    """Q2: adaLN-Zero / final / IPA linear_out are zero-initialised
    (latent_model.py:144-177) so a fresh model outputs exactly 0; redraw them."""
    g = torch.Generator().manual_seed(seed)
    for k in sorted(sd.keys()):
        v = sd[k]
        if v.dtype.is_floating_point and v.numel() > 1 and not v.any() and "pos_embed" not in k:
            sd[k] = torch.randn(v.shape, generator=g) * std
    return sd


# ------------------------------------------------------------------------------------
# prep_batch (wrapper.py:283-365): trajectory frames -> latents + conditioning, the step before the sampler
DESIGN_IDX = [1, 2]                 # wrapper.py:41-43
COND_IDX = [0, 3]
DESIGN_MAP_TO_COND = [0, 0, 3, 3]


def get_offsets(ref_rot, ref_trans, rot, trans):
    """utils.get_offsets (utils.py:7-14): ref^-1 o rigids as 7-vectors, then the sign fix of wrapper.py:308-309
    (quaternion real part >= 0), which also removes the sign ambiguity of the eigenvector in rot_to_quat."""
    ri, ti = rigid_invert(ref_rot, ref_trans)
    off = to_tensor_7(*rigid_compose(ri, ti, rot, trans))
    sign = torch.where(off[..., 0:1] < 0, -1.0, 1.0)
    return torch.cat([off[..., :4] * sign, off[..., 4:]], -1)


def prep_batch(args, batch):
    """NewMDGenWrapper.prep_batch (wrapper.py:283-365), frames path (no --no_frames, no hyena latents).
    batch: trans [B,T,L,3], rots [B,T,L,3,3], torsions [B,T,L,7,2], torsion_mask [B,L,7], mask [B,L], seqres [B,L]."""
    rot, trans = batch["rots"], batch["trans"]
    B, T, L = trans.shape[:3]
    if getattr(args, "design_key_frames", False):                                     # :296-302
        idx = torch.tensor(DESIGN_MAP_TO_COND)
        rot = torch.cat([rot[:, :1, idx], rot[:, 1:-1], rot[:, -1:, idx]], 1)
        trans = torch.cat([trans[:, :1, idx], trans[:, 1:-1], trans[:, -1:, idx]], 1)
    if getattr(args, "no_offsets", False):                                            # :304-305
        offsets = get_offsets(torch.eye(3).expand(B, 1, L, 3, 3), torch.zeros(B, 1, L, 3), rot, trans)
    else:
        offsets = get_offsets(rot[:, 0:1], trans[:, 0:1], rot, trans)                 # :307-309
    frame_loss_mask = batch["mask"].unsqueeze(-1).expand(-1, -1, 7)
    torsion_loss_mask = batch["torsion_mask"].unsqueeze(-1).expand(-1, -1, -1, 2).reshape(B, L, 14)
    if args.tps_condition or args.inpainting or getattr(args, "dynamic_mpnn", False):  # :314-318
        offsets = torch.cat([offsets, get_offsets(rot[:, -1:], trans[:, -1:], rot, trans)], -1)
        frame_loss_mask = torch.cat([frame_loss_mask, frame_loss_mask], -1)
    tors = batch["torsions"].reshape(B, T, L, 14)
    if getattr(args, "no_torsion", False):                                            # :320-327
        tors = torch.zeros_like(tors)
    elif getattr(args, "no_design_torsion", False):
        tors = tors.clone()
        tors[:, :, DESIGN_IDX] = 0
    latents = torch.cat([offsets, tors], -1)
    if getattr(args, "supervise_all_torsions", False):                                # :329-332
        torsion_loss_mask = torch.ones_like(torsion_loss_mask)
    elif getattr(args, "supervise_no_torsions", False):
        torsion_loss_mask = torch.zeros_like(torsion_loss_mask)
    loss_mask = torch.cat([frame_loss_mask, torsion_loss_mask], -1).unsqueeze(1).expand(-1, T, -1, -1)
    cond_mask = torch.zeros(B, T, L, dtype=torch.int64)                               # :338-346
    if args.sim_condition:
        cond_mask[:, 0] = 1
    if args.tps_condition:
        cond_mask[:, 0] = cond_mask[:, -1] = 1
    if getattr(args, "cond_interval", None):
        cond_mask[:, ::args.cond_interval] = 1
    if args.inpainting or getattr(args, "dynamic_mpnn", False) or getattr(args, "mpnn", False):
        cond_mask[:, :, COND_IDX] = 1
    aatype_mask = torch.ones_like(batch["seqres"])
    if args.design:
        aatype_mask[:, DESIGN_IDX] = 0
    return {
        "rigids": (rot, trans),
        "latents": latents,
        "loss_mask": loss_mask,
        "model_kwargs": {
            "start_frames": (rot[:, 0], trans[:, 0]),
            "end_frames": (rot[:, -1], trans[:, -1]),
            "mask": batch["mask"].unsqueeze(1).expand(-1, T, -1),
            "aatype": torch.where(aatype_mask.bool(), batch["seqres"], 20),
            "x_cond": torch.where(cond_mask.unsqueeze(-1).bool(), latents, 0.0),
            "x_cond_mask": cond_mask,
        },
    }


# ------------------------------------------------------------------------------------
# decode (wrapper.py:487-508): sampled latents -> atom14 coordinates, the step after the sampler
def quat_to_rot(q):
    """rigid_utils.quat_to_rot (rigid_utils.py:129-165) for a (normalised) quaternion [*, 4] -> [*, 3, 3]."""
    w, x, y, z = q.unbind(-1)
    return torch.stack([w * w + x * x - y * y - z * z, 2 * (x * y - w * z), 2 * (x * z + w * y),
                        2 * (x * y + w * z), w * w - x * x + y * y - z * z, 2 * (y * z - w * x),
                        2 * (x * z - w * y), 2 * (y * z + w * x), w * w - x * x - y * y + z * z], -1).reshape(*q.shape[:-1], 3, 3)


def torsion_angles_to_frames(rot, trans, alpha, aatype, rrgdf):
    """geometry.torsion_angles_to_frames (geometry.py:273-334): rot/trans [*, N], alpha [*, N, 7, 2] -> 8 global frames."""
    d4 = rrgdf[aatype]                                               # [*, N, 8, 4, 4]
    d_rot, d_trans = d4[..., :3, :3], d4[..., :3, 3]
    bb = alpha.new_zeros(*alpha.shape[:-2], 1, 2)
    bb[..., 1] = 1
    alpha = torch.cat([bb, alpha], -2)                               # [*, N, 8, 2]
    rx = alpha.new_zeros(*alpha.shape[:-1], 3, 3)
    rx[..., 0, 0] = 1
    rx[..., 1, 1] = alpha[..., 1]
    rx[..., 1, 2] = -alpha[..., 0]
    rx[..., 2, 1] = alpha[..., 0]
    rx[..., 2, 2] = alpha[..., 1]
    f_rot, f_trans = d_rot @ rx, d_trans                              # default_r.compose(all_rots)
    rots, transs = [f_rot[..., g, :, :] for g in range(8)], [f_trans[..., g, :] for g in range(8)]
    for g in (5, 6, 7):                                               # chi2..chi4 chained onto chi1 (:314-319)
        rots[g], transs[g] = rigid_compose(rots[g - 1], transs[g - 1], rots[g], transs[g])
    to_bb_rot, to_bb_trans = torch.stack(rots, -3), torch.stack(transs, -2)
    return rigid_compose(rot[..., None, :, :], trans[..., None, :], to_bb_rot, to_bb_trans)


def frames_torsions_to_atom14(rot, trans, torsions, aatype, tables):
    """geometry.frames_torsions_to_atom14 (geometry.py:61-79, 236-270).  tables = the four residue_constants arrays:
    default_frames [21,8,4,4], group_idx [21,14], atom_mask [21,14], lit_positions [21,14,3]."""
    g_rot, g_trans = torsion_angles_to_frames(rot, trans, torsions, aatype, tables["default_frames"])
    gi = tables["group_idx"][aatype].long()                           # [*, N, 14]
    a_rot = torch.gather(g_rot, -3, gi[..., None, None].expand(*gi.shape, 3, 3))
    a_trans = torch.gather(g_trans, -2, gi[..., None].expand(*gi.shape, 3))
    pos = rot_apply(a_rot, tables["lit_positions"][aatype]) + a_trans
    return pos * tables["atom_mask"][aatype].unsqueeze(-1)


def decode_samples(args, samples, rot0, trans0, seqres, tables):
    """wrapper.py:487-508: samples [B,T,L,D], first-frame rigids rot0/trans0 [B,1,L], seqres [B,L] -> atom14 [B,T,L,14,3],
    aa_out [B,T,L]."""
    B, T, L = samples.shape[:3]
    offsets = samples[..., :7]
    two = args.tps_condition or args.inpainting
    torsions = samples[..., 14:28] if two else samples[..., 7:21]
    logits = samples[..., -20:]
    q = offsets[..., :4]
    q = q / torch.linalg.norm(q, dim=-1, keepdim=True)               # Rigid.from_tensor_7(normalize_quats=True)
    rot, trans = quat_to_rot(q), offsets[..., 4:]
    if not getattr(args, "no_offsets", False):
        rot, trans = rigid_compose(rot0, trans0, rot, trans)          # rigids[:, 0:1].compose(...)
    torsions = torsions.reshape(B, T, L, 7, 2)
    if not getattr(args, "oracle", False):
        torsions = torsions / torch.linalg.norm(torsions, dim=-1, keepdim=True)
    aatype = seqres[:, None].expand(B, T, L)
    atom14 = frames_torsions_to_atom14(rot, trans, torsions, aatype, tables)
    aa_out = torch.argmax(logits, -1) if args.design else aatype
    return atom14, aa_out
