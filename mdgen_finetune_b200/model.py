"""Host-side mirror of the reference's LatentMDGenModel (src/mdgen/model/latent_model.py:43-326).

Same constructor (`LatentMDGenModel(args, latent_dim)`), same `forward` / `forward_inference`
signatures, same parameter / buffer names and shapes (strict `load_state_dict` of a reference
checkpoint works), real `nn.Linear` children under the reference names (PEFT/LoRA attaches by
name, gfn_diffusion.py:58-83; the merged weights reach the CUDA path, autograd.py).  The modules below are parameter CONTAINERS only: every call is
forwarded to the CUDA library through the C ABI (include/mdgen_b200.h); there is no PyTorch
compute path and no CPU fallback.
"""
import ctypes
import math

import numpy as np
import torch
import torch.nn as nn

from . import _lib
from .weights import param_shapes, sincos_pos_embed, rope_inv_freq


# ---- parameter containers (names = reference attribute names) ------------------------------
class _RotEmb(nn.Module):                       # esm.rotary_embedding.RotaryEmbedding buffer
    def __init__(self, dim):
        super().__init__()
        self.register_buffer("inv_freq", rope_inv_freq(dim))


class _MHA(nn.Module):                          # mha.py:69-130
    def __init__(self, C, heads, rope):
        super().__init__()
        self.k_proj = nn.Linear(C, C)
        self.v_proj = nn.Linear(C, C)
        self.q_proj = nn.Linear(C, C)
        self.out_proj = nn.Linear(C, C)
        self.bias_k = nn.Parameter(torch.empty(1, 1, C))
        self.bias_v = nn.Parameter(torch.empty(1, 1, C))
        nn.init.xavier_normal_(self.bias_k)
        nn.init.xavier_normal_(self.bias_v)
        if rope:
            self.rot_emb = _RotEmb(C // heads)


class _AttentionWithRoPE(nn.Module):            # latent_model.py:329-338
    def __init__(self, C, heads, rope):
        super().__init__()
        self.attn = _MHA(C, heads, rope)


class _IPA(nn.Module):                          # ipa.py:34-90
    def __init__(self, C, H, c, Pq, Pv):
        super().__init__()
        self.linear_q = nn.Linear(C, H * c)
        self.linear_kv = nn.Linear(C, 2 * H * c)
        self.linear_q_points = nn.Linear(C, H * Pq * 3)
        self.linear_kv_points = nn.Linear(C, H * (Pq + Pv) * 3)
        self.head_weights = nn.Parameter(torch.full((H,), 0.541324854612918))   # softplus^-1(1)
        self.linear_out = nn.Linear(H * (c + Pv * 4), C)


def _adaLN(C, n):
    return nn.Sequential(nn.SiLU(), nn.Linear(C, n * C, bias=True))


class _IPALayer(nn.Module):                     # latent_model.py:341-376
    def __init__(self, args):
        super().__init__()
        C = args.embed_dim
        self.adaLN_modulation = _adaLN(C, 6)
        self.ipa_norm = nn.LayerNorm(C)
        self.ipa = _IPA(C, args.ipa_heads, args.ipa_head_dim, args.ipa_qk, args.ipa_v)
        self.mha_l = _AttentionWithRoPE(C, args.mha_heads, not args.no_rope)
        self.fc1 = nn.Linear(C, 4 * C)
        self.fc2 = nn.Linear(4 * C, C)


class _MainLayer(nn.Module):                    # latent_model.py:397-454
    def __init__(self, args):
        super().__init__()
        C = args.embed_dim
        self.adaLN_modulation = _adaLN(C, 9)
        self.mha_t = _AttentionWithRoPE(C, args.mha_heads, not args.no_rope)
        self.mha_l = _AttentionWithRoPE(C, args.mha_heads, not args.no_rope)
        self.fc1 = nn.Linear(C, 4 * C)
        self.fc2 = nn.Linear(4 * C, C)


class _FinalLayer(nn.Module):                   # layers.py:58-69
    def __init__(self, C, D):
        super().__init__()
        self.linear = nn.Linear(C, D)
        self.adaLN_modulation = _adaLN(C, 2)


class _TimestepEmbedder(nn.Module):             # layers.py:17-27
    def __init__(self, C):
        super().__init__()
        self.mlp = nn.Sequential(nn.Linear(256, C), nn.SiLU(), nn.Linear(C, C))


def frames_to_tensors(fr):
    """Accepts a reference `Rigid` (rigid_utils.py:813: `_rots.get_rot_mats()`, `_trans`), any
    object with `.rot` / `.trans`, or a (rot [B,L,3,3], trans [B,L,3]) tuple."""
    if fr is None:
        return None, None
    if isinstance(fr, (tuple, list)):
        rot, trans = fr
    elif hasattr(fr, "_rots"):
        rot, trans = fr._rots.get_rot_mats(), fr._trans
    else:
        rot, trans = fr.rot, fr.trans
    return rot, trans


def _rel7_eigh(start, end):
    """to_tensor_7 of start^-1 o end and end^-1 o start exactly as the reference computes it
    (latent_model.py:198-199 -> rigid_utils.py:191-215): quaternion = top eigenvector from
    torch.linalg.eigh ON THE INPUT DEVICE, whose sign is implementation-defined.
    Step-invariant: runs once per conditioning, never inside the sampler loop."""
    def one(Ra, ta, Rb, tb):
        R = Ra.transpose(-1, -2) @ Rb
        t = (Ra.transpose(-1, -2) @ (tb - ta)[..., None])[..., 0]
        xx, xy, xz = R[..., 0, 0], R[..., 0, 1], R[..., 0, 2]
        yx, yy, yz = R[..., 1, 0], R[..., 1, 1], R[..., 1, 2]
        zx, zy, zz = R[..., 2, 0], R[..., 2, 1], R[..., 2, 2]
        k = [[xx + yy + zz, zy - yz, xz - zx, yx - xy], [zy - yz, xx - yy - zz, xy + yx, xz + zx],
             [xz - zx, xy + yx, yy - xx - zz, yz + zy], [yx - xy, xz + zx, yz + zy, zz - xx - yy]]
        k = (1.0 / 3.0) * torch.stack([torch.stack(r, -1) for r in k], -2)
        _, vec = torch.linalg.eigh(k)
        return torch.cat([vec[..., -1], t], -1).contiguous()
    return one(*start, *end), one(*end, *start)


class _DirichletFlowTables:
    """Host tables of mdgen.utils.DirichletConditionalFlow (utils.py:33-43) -- scipy, like the reference; the
    arithmetic that uses them (c_factor, utils.py:45-57) runs in dirichlet_flow_kernel.  One row (the alpha nearest
    to the requested one, utils.py:53) is uploaded per distinct alpha and cached."""

    def __init__(self, K=20, alpha_min=1, alpha_max=100, alpha_spacing=0.01):
        import numpy as np
        import scipy.special
        self.alphas = np.arange(alpha_min, alpha_max + alpha_spacing, alpha_spacing)
        self.bs = np.linspace(0, 1, 1000)
        cdfs = np.array([scipy.special.betainc(al, K - 1, self.bs) for al in self.alphas])
        self.beta_cdfs_derivative = np.diff(cdfs, axis=0) / alpha_spacing
        self.alpha_spacing = alpha_spacing
        self.K = K
        self._rows = {}

    def index(self, alpha):
        import numpy as np
        return int(np.argmin(np.abs(alpha - self.alphas)))

    def row(self, alpha, device):
        i = self.index(alpha)
        key = (i, str(device))
        if key not in self._rows:
            if len(self._rows) > 256:
                self._rows.clear()
            self._rows[key] = torch.from_numpy(self.beta_cdfs_derivative[i].copy()).to(device=device, dtype=torch.float64)
        return self._rows[key]

    def beta(self, alpha):
        import scipy.special
        return float(scipy.special.beta(alpha, self.K - 1))


class LatentMDGenModel(nn.Module):
    """Drop-in for mdgen.model.latent_model.LatentMDGenModel (latent_model.py:43)."""

    def __init__(self, args, latent_dim, precision="bf16", quat_sign="eigh"):
        super().__init__()
        self.args = args
        if self.args.design:
            assert self.args.prepend_ipa                                   # latent_model.py:47-48
        for flag in ("hyena", "interleave_ipa", "dynamic_mpnn", "mpnn"):
            if getattr(args, flag, False):
                raise NotImplementedError(f"--{flag} is outside the B200 hot path (DESIGN.md, out of scope)")
        assert precision in ("bf16", "fp32")
        assert quat_sign in ("eigh", "eigh_cpu", "closed_form")
        self.precision = precision
        self.quat_sign = quat_sign
        self._condflow = None
        self.latent_dim = latent_dim
        C = args.embed_dim
        self.latent_to_emb = nn.Linear(latent_dim, C)
        if args.tps_condition or args.inpainting:
            self.latent_to_emb_f = nn.Linear(7, C)
            self.latent_to_emb_r = nn.Linear(7, C)
        cond_dim = latent_dim - 20 if args.design else latent_dim
        self.cond_to_emb = nn.Linear(cond_dim, C)
        self.mask_to_emb = nn.Embedding(2, C)
        if args.design:
            self.x_d_to_emb = nn.Linear(20, C)
        if args.prepend_ipa:
            if not args.no_aa_emb:
                self.aatype_to_emb = nn.Embedding(21, C)
            self.ipa_layers = nn.ModuleList([_IPALayer(args) for _ in range(args.num_layers)])
        self.layers = nn.ModuleList([_MainLayer(args) for _ in range(args.num_layers)])
        self.emb_to_latent = _FinalLayer(C, latent_dim)
        if args.design:
            self.fc1 = nn.Linear(C, C)
            self.fc2 = nn.Linear(C, C)
            self.fc3 = nn.Linear(C, C)
            self.emb_to_logits = nn.Linear(C, 20)
        self.t_embedder = _TimestepEmbedder(C)
        if args.abs_pos_emb:
            self.register_buffer("pos_embed", torch.zeros(1, args.crop, C))
        if getattr(args, "abs_time_emb", False):
            self.register_buffer("time_embed", torch.zeros(1, args.num_frames, C))
        self.initialize_weights()
        self._handle = None
        self._weights_key = None
        self._cond_key = None
        self._keepalive = None
        expect = param_shapes(args, latent_dim)
        got = {k: tuple(v.shape) for k, v in self.state_dict().items()}
        assert got == {k: tuple(v) for k, v in expect.items()}, "state_dict layout drifted from SURVEY 8b"
        self._ref_names = tuple(expect.keys())        # reference state_dict names (stable under PEFT wrapping of children)

    @property
    def device(self):
        return next(self.parameters()).device

    def initialize_weights(self):
        """latent_model.py:134-177: Xavier on every Linear, zero biases, adaLN-Zero, zero final /
        IPA linear_out, sin-cos position tables, N(0, 0.02) timestep MLP."""
        def _basic(m):
            if isinstance(m, nn.Linear):
                nn.init.xavier_uniform_(m.weight)
                if m.bias is not None:
                    nn.init.constant_(m.bias, 0)
        self.apply(_basic)
        if self.args.prepend_ipa:
            for blk in self.ipa_layers:
                nn.init.constant_(blk.ipa.linear_out.weight, 0)
                nn.init.constant_(blk.ipa.linear_out.bias, 0)
        if self.args.abs_pos_emb:
            self.pos_embed.data.copy_(sincos_pos_embed(self.pos_embed.shape[-1], self.args.crop))
        if getattr(self.args, "abs_time_emb", False):
            self.time_embed.data.copy_(sincos_pos_embed(self.time_embed.shape[-1], self.args.num_frames))
        nn.init.normal_(self.t_embedder.mlp[0].weight, std=0.02)
        nn.init.normal_(self.t_embedder.mlp[2].weight, std=0.02)
        for blk in self.layers:
            nn.init.constant_(blk.adaLN_modulation[-1].weight, 0)
            nn.init.constant_(blk.adaLN_modulation[-1].bias, 0)
        nn.init.constant_(self.emb_to_latent.adaLN_modulation[-1].weight, 0)
        nn.init.constant_(self.emb_to_latent.adaLN_modulation[-1].bias, 0)
        nn.init.constant_(self.emb_to_latent.linear.weight, 0)
        nn.init.constant_(self.emb_to_latent.linear.bias, 0)

    # ---- C-ABI plumbing --------------------------------------------------------------------
    def _config(self):
        a = self.args
        flags = 0
        for cond, bit in ((a.sim_condition, _lib.F_SIM), (a.tps_condition, _lib.F_TPS),
                          (a.inpainting, _lib.F_INPAINT), (a.design, _lib.F_DESIGN),
                          (a.no_aa_emb, _lib.F_NO_AA_EMB), (a.abs_pos_emb, _lib.F_ABS_POS),
                          (getattr(a, "abs_time_emb", False), _lib.F_ABS_TIME), (a.no_rope, _lib.F_NO_ROPE),
                          (a.prepend_ipa, _lib.F_PREPEND_IPA)):
            if cond:
                flags |= bit
        return _lib.MdgenConfig(
            embed_dim=a.embed_dim, num_layers=a.num_layers, mha_heads=a.mha_heads, ipa_heads=a.ipa_heads,
            ipa_head_dim=a.ipa_head_dim, ipa_qk=a.ipa_qk, ipa_v=a.ipa_v, latent_dim=self.latent_dim,
            crop=a.crop, num_frames=a.num_frames, flags=flags,
            precision=_lib.PREC_BF16 if self.precision == "bf16" else _lib.PREC_FP32_CHECK,
            time_multiplier=float(a.time_multiplier))

    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _raw_key(self):
        """Identity of every tensor the effective weights are built from (parameters incl. LoRA adapters, buffers)."""
        items = list(self.named_parameters()) + list(self.named_buffers())
        return tuple((k, v.data_ptr(), v._version) for k, v in items)

    def _sync_weights(self):
        dev = self.device
        if dev.type != "cuda":
            raise RuntimeError("mdgen_finetune_b200.LatentMDGenModel runs on CUDA only (no CPU fallback); "
                               "call .cuda() first")
        key = (dev.index, self.precision) + self._raw_key()
        if key == self._weights_key:
            return
        from .autograd import effective_tensors
        with torch.no_grad():
            sd = effective_tensors(self, self._ref_names)      # LoRA-wrapped linears contribute W + scaling * B @ A
        L = _lib.lib()
        with torch.cuda.device(dev):
            if self._handle is None or self._weights_key is None or self._weights_key[:2] != key[:2]:
                self._destroy()
                h = ctypes.c_void_p()
                cfg = self._config()
                _lib.check(L.mdgen_create(ctypes.byref(cfg), ctypes.byref(h)), "mdgen_create")
                self._handle = h
            st = self._stream()
            for name, v in sd.items():
                t = v.detach().to(dev, torch.float32).contiguous()
                shape = (ctypes.c_int64 * t.dim())(*t.shape)
                _lib.check(L.mdgen_load_weight(self._handle, name.encode(), t.data_ptr(), shape, t.dim(), 1, st),
                           f"mdgen_load_weight({name})")
            _lib.check(L.mdgen_finalize_weights(self._handle, st), "mdgen_finalize_weights")
        self._weights_key = key
        self._cond_key = None

    def _destroy(self):
        if getattr(self, "_handle", None) is not None:
            try:
                _lib.lib().mdgen_destroy(self._handle)
            except Exception:
                pass
            try:
                self._handle = None
            except Exception:        # interpreter shutdown: torch's module machinery may already be torn down
                pass

    def __del__(self):
        try:
            self._destroy()
        except Exception:
            pass

    def __deepcopy__(self, memo):
        import copy
        new = LatentMDGenModel(copy.deepcopy(self.args, memo), self.latent_dim, self.precision, self.quat_sign)
        new.load_state_dict(self.state_dict())
        return new.to(self.device)

    # ---- conditioning (everything forward() gets besides x, t) -----------------------------
    @staticmethod
    def _tkey(t):
        return None if t is None else (t.data_ptr(), t._version, tuple(t.shape), tuple(t.stride()), t.dtype, str(t.device))

    def invalidate(self):
        """Drops the cached weights / conditioning of the CUDA handle.  Needed only after in-place edits that bypass
        autograd's version counter (`p.data.copy_()`, EMA-style updates): those change neither `_version` nor the
        address, so the caches cannot see them."""
        self._weights_key = None
        self._cond_key = None
        self._keepalive = None

    def _set_conditioning(self, x, mask, start_frames, end_frames, x_cond, x_cond_mask, aatype):
        dev = self.device
        B, T, L, D = x.shape
        if mask.dim() == 2:                                              # latent_model.py:233-234
            mask = mask.repeat(B, 1, 1)
        srot, strans = frames_to_tensors(start_frames)
        erot, etrans = frames_to_tensors(end_frames)
        originals = (mask, srot, strans, erot, etrans, x_cond, x_cond_mask, aatype)
        key = (B, T, L, self.quat_sign) + tuple(self._tkey(v) for v in originals)
        if key == self._cond_key:
            return
        f32 = lambda v: None if v is None else v.detach().to(dev, torch.float32).contiguous()
        i64 = lambda v: None if v is None else v.detach().to(dev, torch.int64).contiguous()
        m, sr, stt, er, et = f32(mask), f32(srot), f32(strans), f32(erot), f32(etrans)
        xc, cm, aa = f32(x_cond), i64(x_cond_mask), i64(aatype)
        f7 = r7 = None
        two_stream = self.args.tps_condition or self.args.inpainting
        if self.args.prepend_ipa and two_stream and self.quat_sign != "closed_form":
            if er is None:
                raise ValueError("tps_condition / inpainting need end_frames (latent_model.py:197-199)")
            if self.quat_sign == "eigh_cpu":
                f7, r7 = _rel7_eigh((sr.cpu(), stt.cpu()), (er.cpu(), et.cpu()))
                f7, r7 = f7.to(dev), r7.to(dev)
            else:
                f7, r7 = _rel7_eigh((sr, stt), (er, et))
        p = lambda v: None if v is None else ctypes.c_void_p(v.data_ptr())
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().mdgen_set_conditioning(
                self._handle, B, T, L, p(m), p(sr), p(stt), p(er), p(et), p(f7), p(r7), p(xc), p(cm), p(aa),
                self._stream()), "mdgen_set_conditioning")
        # The key holds data_ptr() / _version of the CALLER's tensors: keep those alive (not only the converted copies),
        # so that the address cannot be recycled by a different tensor while the key is current.
        self._keepalive = originals + (m, sr, stt, er, et, xc, cm, aa, f7, r7)
        self._cond_key = key

    # ---- the reference API ---------------------------------------------------------------------
    def forward(self, x, t, mask, start_frames=None, end_frames=None, x_cond=None, x_cond_mask=None,
                aatype=None, *args, **kwargs):
        """One velocity evaluation (latent_model.py:216-269).  Unknown kwargs (e.g. `peptide=`) are
        swallowed like the reference's **kwargs; `t` may be int64 (RTB DDPM timesteps).

        With gradients enabled and x or any parameter requiring them (RTB posterior: samplers.py:136-139, also under
        torch.utils.checkpoint) the call goes through autograd.VelocityFn: forward = mdgen_forward, backward =
        mdgen_backward (row N4).  Otherwise the output is a plain tensor."""
        if self.device.type != "cuda":
            raise RuntimeError("mdgen_finetune_b200.LatentMDGenModel runs on CUDA only (no CPU fallback); "
                               "call .cuda() first")
        if x.dim() != 4 or x.shape[-1] != self.latent_dim:
            raise ValueError(f"x must be [B,T,L,{self.latent_dim}], got {tuple(x.shape)}")
        if t.numel() != x.shape[0]:
            raise ValueError("t must have one entry per batch element")
        cond = (mask, start_frames, end_frames, x_cond, x_cond_mask, aatype)
        if torch.is_grad_enabled() and (x.requires_grad or any(p.requires_grad for p in self.parameters())):
            from .autograd import VelocityFn, effective_tensors
            eff = effective_tensors(self, self._ref_names)
            return VelocityFn.apply(self, cond, self._ref_names, x, t, *[eff[n] for n in self._ref_names])
        return self._forward_no_grad(x, t, cond)

    def _forward_no_grad(self, x, t, cond):
        self._sync_weights()
        dev = self.device
        self._set_conditioning(x, *cond)
        xx = x.detach().to(dev, torch.float32).contiguous()
        tt = t.detach().to(dev, torch.float32).contiguous()
        out = torch.empty_like(xx)
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().mdgen_forward(self._handle, xx.data_ptr(), tt.data_ptr(), out.data_ptr(),
                                                self._stream()), "mdgen_forward")
        return out

    def forward_inference(self, x, t, mask, start_frames=None, end_frames=None, x_cond=None,
                          x_cond_mask=None, aatype=None):
        """latent_model.py:272-326.  Non-design: pass-through to forward."""
        if not self.args.design:
            return self.forward(x, t, mask, start_frames, end_frames, x_cond, x_cond_mask, aatype)
        # ---- design: logits -> Dirichlet conditional-flow velocity of the 20 sequence channels (:279-326).
        # forward() leaves latent[..., -20:] = logits; the softmax, c_factor and the flow reduction run in one CUDA
        # kernel (mdgen_dirichlet_flow) instead of the reference's .cpu().numpy() round trip.
        latent = self.forward(x, t, mask, start_frames, end_frames, x_cond, x_cond_mask, aatype)
        dev = self.device
        xx = x.detach().to(dev, torch.float32).contiguous()
        a = self.args
        cf = self.condflow
        alpha = self._alpha_of(t.reshape(-1)[0])
        dalpha = float(a.alpha_max - 1)
        row = cf.row(alpha, dev)
        flags = (ctypes.c_int32 * 2)()
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().mdgen_dirichlet_flow(
                self._handle, xx.data_ptr(), latent.data_ptr(), row.data_ptr(), alpha, cf.beta(alpha),
                float(a.dirichlet_flow_temp), dalpha, int(bool(getattr(a, "allow_nan_cfactor", False))), flags,
                self._stream()), "mdgen_dirichlet_flow")
        self._report_flow_flags(flags)
        return latent

    def _alpha_of(self, t0):
        """t_to_alpha (transport.py:52-57) on t[0] in the tensor's own precision, then .item() and the clamp of
        latent_model.py:305-309."""
        t0 = torch.as_tensor(t0, dtype=torch.float32).detach().cpu()
        alpha = (1 * (1 - t0) + t0 * self.args.alpha_max).item()
        if alpha > self.args.alpha_max:
            alpha = self.args.alpha_max - self.condflow.alpha_spacing
        return alpha

    def _report_flow_flags(self, flags):
        if flags[0]:                                                      # latent_model.py:282-290
            print(f"WARNING: {flags[0]} positions of xt do not lie on the simplex (negative entries or sum != 1).")
        if flags[1]:                                                      # :312-319
            print(f"NAN cfactor: {flags[1]} entries")
            if not getattr(self.args, "allow_nan_cfactor", False):
                raise RuntimeError(f"NAN cfactor: {flags[1]} entries")

    @property
    def condflow(self):
        """DirichletConditionalFlow tables (utils.py:33-43; built by the reference constructor, latent_model.py:124-126,
        here on first use): d/dalpha of the regularised incomplete beta function on a 1000-point grid."""
        if not self.args.design:
            raise AttributeError("condflow exists only in design mode (latent_model.py:124)")
        if self._condflow is None:
            self._condflow = _DirichletFlowTables(K=20, alpha_spacing=0.001, alpha_max=self.args.alpha_max)
        return self._condflow

    @torch.no_grad()
    def euler_rollout(self, x0, num_steps=50, return_trajectory=False, **model_kwargs):
        """Whole fixed-grid Euler solve in one C-ABI call (what Sampler.sample_ode('euler') does,
        transport.py:407-450): returns x at t=1, or the [num_steps, ...] stack."""
        self._sync_weights()
        dev = self.device
        kw = dict(model_kwargs)
        self._set_conditioning(x0, kw.get("mask"), kw.get("start_frames"), kw.get("end_frames"),
                               kw.get("x_cond"), kw.get("x_cond_mask"), kw.get("aatype"))
        xx = x0.detach().to(dev, torch.float32).contiguous()
        out = torch.empty_like(xx)
        traj = torch.empty((num_steps,) + tuple(xx.shape), device=dev) if return_trajectory else None
        design = bool(self.args.design)
        if design:
            # per-step Dirichlet-flow tables (forward_inference, latent_model.py:305-311): alpha_i of the grid times,
            # the nearest beta-CDF-derivative row and B(alpha_i, 19); the flow itself runs inside the fused rollout
            cf, a = self.condflow, self.args
            grid = torch.linspace(0, 1, num_steps)                      # integrators.py:90
            alphas = np.array([self._alpha_of(grid[i]) for i in range(num_steps - 1)], dtype=np.float64)
            rows = np.ascontiguousarray(np.stack([cf.beta_cdfs_derivative[cf.index(al)] for al in alphas]), dtype=np.float64)
            betas = np.array([cf.beta(al) for al in alphas], dtype=np.float64)
            with torch.cuda.device(dev):
                _lib.check(_lib.lib().mdgen_set_dirichlet_tables(
                    self._handle, num_steps - 1, rows.ctypes.data, alphas.ctypes.data, betas.ctypes.data,
                    float(a.dirichlet_flow_temp), float(a.alpha_max - 1),
                    int(bool(getattr(a, "allow_nan_cfactor", False))), self._stream()), "mdgen_set_dirichlet_tables")
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().mdgen_euler_rollout(
                self._handle, xx.data_ptr(), int(num_steps), None if traj is None else traj.data_ptr(),
                out.data_ptr(), self._stream()), "mdgen_euler_rollout")
            if design:      # the reference warns / raises at the offending step; the fused rollout reports at its end
                flags = (ctypes.c_int32 * 2)()
                _lib.check(_lib.lib().mdgen_dirichlet_flags(self._handle, flags, self._stream()), "mdgen_dirichlet_flags")
                self._report_flow_flags(flags)
        return traj if return_trajectory else out

    def launches_per_forward(self):
        return int(_lib.lib().mdgen_launches_per_forward(self._handle)) if self._handle else 0
