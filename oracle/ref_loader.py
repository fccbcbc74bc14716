"""TEST INFRASTRUCTURE ONLY -- loads the UNMODIFIED reference from /root/reference/src.

Only `oracle/make_golden.py` (run in the build container, where /root/reference
exists) uses this module.  Nothing here travels to the GPU box at run time: the
golden vectors it produces are committed under tests/golden/.

What it does:
  * installs `sys.modules` stubs, each with a real ModuleSpec, for the third-party
    packages the reference imports but this image lacks: `esm.rotary_embedding`
    (fair-esm 1.0.2, environment.yml:21), `torchdiffeq` (0.2.5, environment.yml:266),
    `tree` (dm-tree), `Bio.PDB`;
  * hides `flash_attn` / `deepspeed` from importlib.util.find_spec only while
    mdgen/model/primitives.py:20-31 probes for them, then restores find_spec;
  * applies the Q1 fix: latent_model.py:255 passes
    `mask.reshape(-1, L)` to run_ipa, which only works for T == 1; upstream MDGen
    passes `mask[:, 0]`.  We monkey-patch `run_ipa` to take rows [::T] of that mask,
    which is identical to `mask[:, 0]` and leaves the reference source untouched.

The two arithmetic stubs restate the published algorithms of the pinned packages:
  RotaryEmbedding  -- fair-esm esm/rotary_embedding.py (call sites mha.py:13,129-130,356-357)
  odeint(euler)    -- torchdiffeq fixed-grid Euler (call site integrators.py:106-113)
Parity at those two boundaries is "unpinned" in the reference (it has no tests).
"""
import importlib
import importlib.machinery
import importlib.util
import sys
import types

import torch

import os

# The reference's sources: /root/reference/src in the build container; on the GPU box (no /root/reference) the staged
# copy under oracle/_ref/src made by oracle/stage_ref.py (git-ignored build output, shipped like the built .so files).
_HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = "/root/reference/src" if os.path.isdir("/root/reference/src") else os.path.join(_HERE, "_ref", "src")


def _stub(name, **attrs):
    m = types.ModuleType(name)
    m.__spec__ = importlib.machinery.ModuleSpec(name, loader=None)
    for k, v in attrs.items():
        setattr(m, k, v)
    sys.modules[name] = m
    return m


# ---- fair-esm RotaryEmbedding restatement -------------------------------------------
def _rotate_half(x):
    x1, x2 = x.chunk(2, dim=-1)
    return torch.cat((-x2, x1), dim=-1)


def _apply_rotary_pos_emb(x, cos, sin):
    cos = cos[:, : x.shape[-2], :]
    sin = sin[:, : x.shape[-2], :]
    return (x * cos) + (_rotate_half(x) * sin)


class RotaryEmbedding(torch.nn.Module):
    def __init__(self, dim, *_, **__):
        super().__init__()
        inv_freq = 1.0 / (10000 ** (torch.arange(0, dim, 2).float() / dim))
        self.register_buffer("inv_freq", inv_freq)
        self._seq_len_cached = None
        self._cos_cached = None
        self._sin_cached = None

    def _update_cos_sin_tables(self, x, seq_dimension=1):
        seq_len = x.shape[seq_dimension]
        if seq_len != self._seq_len_cached or self._cos_cached.device != x.device:
            self._seq_len_cached = seq_len
            t = torch.arange(x.shape[seq_dimension], device=x.device).type_as(self.inv_freq)
            freqs = torch.einsum("i,j->ij", t, self.inv_freq)
            emb = torch.cat((freqs, freqs), dim=-1).to(x.device)
            self._cos_cached = emb.cos()[None, :, :]
            self._sin_cached = emb.sin()[None, :, :]
        return self._cos_cached, self._sin_cached

    def forward(self, q, k):
        self._cos_cached, self._sin_cached = self._update_cos_sin_tables(k, seq_dimension=-2)
        return (_apply_rotary_pos_emb(q, self._cos_cached, self._sin_cached),
                _apply_rotary_pos_emb(k, self._cos_cached, self._sin_cached))


# ---- torchdiffeq fixed-grid Euler restatement -----------------------------------------
def odeint(func, y0, t, *, method="euler", atol=None, rtol=None, **_):
    if method != "euler":
        raise NotImplementedError("oracle stub restates fixed-grid Euler only (SURVEY Q3)")
    ys = [y0]
    y = y0
    for i in range(len(t) - 1):
        dt = t[i + 1] - t[i]
        y = y + dt * func(t[i], y)
        ys.append(y)
    return torch.stack(ys, 0)


_loaded = None


def load_reference():
    """Returns a namespace with the reference's LatentMDGenModel, Sampler, create_transport, Rigid..."""
    global _loaded
    if _loaded is not None:
        return _loaded
    _stub("esm")
    _stub("esm.rotary_embedding", RotaryEmbedding=RotaryEmbedding)
    _stub("torchdiffeq", odeint=odeint)
    def map_structure(f, s):          # dm-tree: recursive map over nested lists/dicts
        if isinstance(s, (list, tuple)):
            return type(s)(map_structure(f, e) for e in s)
        if isinstance(s, dict):
            return {k: map_structure(f, v) for k, v in s.items()}
        return f(s)

    _stub("tree", map_structure=map_structure)
    _stub("Bio")
    _stub("Bio.PDB", PDBParser=object)
    if not os.path.isdir(REF_SRC):
        raise RuntimeError(f"reference sources not found at {REF_SRC}: run `python oracle/stage_ref.py` in the build container")
    if REF_SRC not in sys.path:
        sys.path.insert(0, REF_SRC)

    real_find_spec = importlib.util.find_spec

    def hiding_find_spec(name, package=None):
        if name.split(".")[0] in ("flash_attn", "deepspeed"):
            return None
        return real_find_spec(name, package)

    importlib.util.find_spec = hiding_find_spec
    try:
        import mdgen.model.primitives  # noqa: F401
    finally:
        importlib.util.find_spec = real_find_spec

    from mdgen.model import latent_model as lm
    from mdgen.transport.transport import Sampler, create_transport
    from mdgen import rigid_utils
    from mdgen.model.mha import MultiheadAttention
    from mdgen.model.ipa import InvariantPointAttention

    # Q1: mask.reshape(-1, L) has B*T rows; rows [::T] are mask[:, 0].
    orig_run_ipa = lm.LatentMDGenModel.run_ipa

    def run_ipa_q1(self, t, mask, start_frames, end_frames, aatype, x_d=None):
        B = t.shape[0]
        if mask.shape[0] != B:
            T = mask.shape[0] // B
            mask = mask[::T]
        return orig_run_ipa(self, t, mask, start_frames, end_frames, aatype, x_d=x_d)

    lm.LatentMDGenModel.run_ipa = run_ipa_q1

    _loaded = types.SimpleNamespace(
        lm=lm, LatentMDGenModel=lm.LatentMDGenModel, IPALayer=lm.IPALayer,
        LatentMDGenLayer=lm.LatentMDGenLayer, Sampler=Sampler, create_transport=create_transport,
        Rigid=rigid_utils.Rigid, Rotation=rigid_utils.Rotation, rigid_utils=rigid_utils,
        MultiheadAttention=MultiheadAttention, InvariantPointAttention=InvariantPointAttention,
    )
    return _loaded


# ---- mdgen.wrapper (NewMDGenWrapper: prep_batch / inference, rows N1 / N2) ---------------------------------------
class _LightningModuleStub(torch.nn.Module):
    """pytorch_lightning.LightningModule as far as mdgen/wrapper.py uses it outside a Trainer: an nn.Module with
    `save_hyperparameters()` (wrapper.py:48) and a `device` property (wrapper.py:228,431-433,458-460)."""

    def save_hyperparameters(self, *a, **k):
        pass

    @property
    def device(self):
        try:
            return next(self.parameters()).device
        except StopIteration:
            return torch.device("cpu")


def load_wrapper():
    """The reference's NewMDGenWrapper (src/mdgen/wrapper.py:175) with `pytorch_lightning` and `wandb` stubbed
    (both absent from this image; neither does arithmetic on the prep_batch / inference path)."""
    ref = load_reference()
    if "pytorch_lightning" not in sys.modules:
        _stub("pytorch_lightning", LightningModule=_LightningModuleStub)
    if "wandb" not in sys.modules:
        _stub("wandb", log=lambda *a, **k: None)
    from mdgen import wrapper
    ref.wrapper = wrapper
    ref.NewMDGenWrapper = wrapper.NewMDGenWrapper
    return ref
