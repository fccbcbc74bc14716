"""state_dict layout of LatentMDGenModel and seeded random weights.

`param_shapes(args, latent_dim)` is the single source of truth for parameter/buffer names
and shapes; it mirrors what the reference constructor registers
(latent_model.py:50-126,354-376,413-454; mha.py:111-130; ipa.py:66-87; layers.py:23-27,64-69).
`random_state_dict` draws every entry from a seeded CPU generator with NO all-zero tensors
(the reference's adaLN-Zero init makes a fresh model output exactly 0, which
would make parity tests vacuous).
"""
import math
from collections import OrderedDict

import numpy as np
import torch


def _mha_shapes(p, C):
    o = OrderedDict()
    o[p + ".attn.bias_k"] = (1, 1, C)
    o[p + ".attn.bias_v"] = (1, 1, C)
    for n in ("k_proj", "v_proj", "q_proj", "out_proj"):
        o[f"{p}.attn.{n}.weight"] = (C, C)
        o[f"{p}.attn.{n}.bias"] = (C,)
    return o


def _ipa_shapes(p, C, H, c, Pq, Pv):
    o = OrderedDict()
    o[p + ".head_weights"] = (H,)
    for n, out in (("linear_q", H * c), ("linear_kv", 2 * H * c), ("linear_q_points", H * Pq * 3),
                   ("linear_kv_points", H * (Pq + Pv) * 3)):
        o[f"{p}.{n}.weight"] = (out, C)
        o[f"{p}.{n}.bias"] = (out,)
    o[p + ".linear_out.weight"] = (C, H * (c + Pv * 4))
    o[p + ".linear_out.bias"] = (C,)
    return o


def param_shapes(args, latent_dim):
    """Ordered {name: shape} in the reference's registration order; buffers included."""
    C = args.embed_dim
    F = 4 * C
    H, c, Pq, Pv = args.ipa_heads, args.ipa_head_dim, args.ipa_qk, args.ipa_v
    rope = not args.no_rope
    hd = C // args.mha_heads
    o = OrderedDict()
    if getattr(args, "abs_pos_emb", False):
        o["pos_embed"] = (1, args.crop, C)
    if getattr(args, "abs_time_emb", False):
        o["time_embed"] = (1, args.num_frames, C)
    o["latent_to_emb.weight"] = (C, latent_dim)
    o["latent_to_emb.bias"] = (C,)
    if args.tps_condition or args.inpainting or args.dynamic_mpnn:
        for n in ("latent_to_emb_f", "latent_to_emb_r"):
            o[n + ".weight"] = (C, 7)
            o[n + ".bias"] = (C,)
    cond_dim = latent_dim - 20 if args.design else latent_dim
    o["cond_to_emb.weight"] = (C, cond_dim)
    o["cond_to_emb.bias"] = (C,)
    o["mask_to_emb.weight"] = (2, C)
    if args.design:
        o["x_d_to_emb.weight"] = (C, 20)
        o["x_d_to_emb.bias"] = (C,)
    if args.prepend_ipa:
        if not args.no_aa_emb:
            o["aatype_to_emb.weight"] = (21, C)
        for i in range(args.num_layers):
            p = f"ipa_layers.{i}"
            o[p + ".adaLN_modulation.1.weight"] = (6 * C, C)
            o[p + ".adaLN_modulation.1.bias"] = (6 * C,)
            o[p + ".ipa_norm.weight"] = (C,)
            o[p + ".ipa_norm.bias"] = (C,)
            o.update(_ipa_shapes(p + ".ipa", C, H, c, Pq, Pv))
            o.update(_mha_shapes(p + ".mha_l", C))
            if rope:
                o[p + ".mha_l.attn.rot_emb.inv_freq"] = (hd // 2,)
            for n, (a, b) in (("fc1", (F, C)), ("fc2", (C, F))):
                o[f"{p}.{n}.weight"] = (a, b)
                o[f"{p}.{n}.bias"] = (a,)
    for i in range(args.num_layers):
        p = f"layers.{i}"
        o[p + ".adaLN_modulation.1.weight"] = (9 * C, C)
        o[p + ".adaLN_modulation.1.bias"] = (9 * C,)
        if args.interleave_ipa:
            o[p + ".ipa_norm.weight"] = (C,)
            o[p + ".ipa_norm.bias"] = (C,)
            o.update(_ipa_shapes(p + ".ipa", C, H, c, Pq, Pv))
        for m in ("mha_t", "mha_l"):
            o.update(_mha_shapes(f"{p}.{m}", C))
            if rope:
                o[f"{p}.{m}.attn.rot_emb.inv_freq"] = (hd // 2,)
        for n, (a, b) in (("fc1", (F, C)), ("fc2", (C, F))):
            o[f"{p}.{n}.weight"] = (a, b)
            o[f"{p}.{n}.bias"] = (a,)
    o["emb_to_latent.linear.weight"] = (latent_dim, C)
    o["emb_to_latent.linear.bias"] = (latent_dim,)
    o["emb_to_latent.adaLN_modulation.1.weight"] = (2 * C, C)
    o["emb_to_latent.adaLN_modulation.1.bias"] = (2 * C,)
    if args.design:
        for n in ("fc1", "fc2", "fc3"):
            o[n + ".weight"] = (C, C)
            o[n + ".bias"] = (C,)
        o["emb_to_logits.weight"] = (20, C)
        o["emb_to_logits.bias"] = (20,)
    o["t_embedder.mlp.0.weight"] = (C, 256)
    o["t_embedder.mlp.0.bias"] = (C,)
    o["t_embedder.mlp.2.weight"] = (C, C)
    o["t_embedder.mlp.2.bias"] = (C,)
    return o


def sincos_pos_embed(dim, n):
    """get_1d_sincos_pos_embed_from_grid (latent_model.py:22-40): float64 [sin | cos]."""
    omega = 1.0 / 10000 ** (np.arange(dim // 2, dtype=np.float64) / (dim / 2.0))
    out = np.einsum("m,d->md", np.arange(n, dtype=np.float64), omega)
    return torch.from_numpy(np.concatenate([np.sin(out), np.cos(out)], 1)).float().unsqueeze(0)


def rope_inv_freq(head_dim):
    """fair-esm RotaryEmbedding buffer: 1 / 10000^(arange(0,d,2)/d)."""
    return 1.0 / (10000 ** (torch.arange(0, head_dim, 2).float() / head_dim))


def random_state_dict(args, latent_dim, seed=0, mod_std=0.05):
    """Seeded, fully non-zero weights in the reference layout (CPU fp32).

    Matrices ~ U(-a, a) with the Xavier bound the reference uses (latent_model.py:136-142);
    the entries the reference zero-initialises (adaLN linears, final linear, IPA linear_out,
    all biases) are drawn N(0, mod_std or 0.02) instead (Q2)."""
    g = torch.Generator().manual_seed(seed)
    sd = OrderedDict()
    hd = args.embed_dim // args.mha_heads
    for name, shape in param_shapes(args, latent_dim).items():
        if name == "pos_embed":
            sd[name] = sincos_pos_embed(shape[-1], shape[1])
        elif name == "time_embed":
            sd[name] = sincos_pos_embed(shape[-1], shape[1])
        elif name.endswith("inv_freq"):
            sd[name] = rope_inv_freq(hd)
        elif name.endswith("head_weights"):
            sd[name] = 0.541324854612918 + 0.1 * torch.randn(shape, generator=g)
        elif name.endswith("ipa_norm.weight"):
            sd[name] = 1.0 + 0.1 * torch.randn(shape, generator=g)
        elif "adaLN_modulation" in name:
            sd[name] = mod_std * torch.randn(shape, generator=g) * (0.3 if name.endswith("weight") else 1.0)
        elif len(shape) == 2 and not name.startswith("mask_to_emb") and not name.startswith("aatype_to_emb"):
            fan_out, fan_in = shape
            a = math.sqrt(6.0 / (fan_in + fan_out))
            sd[name] = (torch.rand(shape, generator=g) * 2 - 1) * a
        elif len(shape) == 3:       # bias_k / bias_v : xavier_normal on [1,1,C] (mha.py:156-159)
            sd[name] = torch.randn(shape, generator=g) * math.sqrt(2.0 / (shape[-1] + 1))
        elif len(shape) == 2:       # embeddings ~ N(0,1) (nn.Embedding default)
            sd[name] = torch.randn(shape, generator=g)
        else:                       # biases
            sd[name] = 0.02 * torch.randn(shape, generator=g)
    return sd
