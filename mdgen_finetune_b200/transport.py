"""Sampler side of the hot path with the reference's interface
(src/mdgen/transport/transport.py:278-292,407-450,513-574; integrators.py:74-114).

`create_transport(args, ...) -> Transport`, `Sampler(transport).sample_ode(sampling_method="euler",
num_steps=50) -> fn(x, model, **model_kwargs) -> Tensor[num_steps, *x.shape]`.

Only the velocity-model / fixed-grid Euler branch is built (dopri5, SDE, likelihood and
score/noise parametrisations are out of scope).  When `model` is (a functools.partial of) a bound
method of our LatentMDGenModel, the whole solve runs inside ONE C-ABI call
(`mdgen_euler_rollout`: Euler AXPY fused into the final projection epilogue); for any other
callable the loop runs here, one `model(x, t, **kw)` per grid step, exactly like
torchdiffeq's fixed-grid Euler.
"""
import functools
import math

import torch


def expand_t_like_x(t, x):
    """path.py:5-13: t [B] -> [B, 1, ...] broadcastable against x."""
    return t.view(t.size(0), *([1] * (x.dim() - 1)))


class ICPlan:
    """Linear interpolant (path.py:18-135): x_t = t x1 + (1 - t) x0, u_t = x1 - x0."""

    def compute_alpha_t(self, t):
        return t, 1

    def compute_sigma_t(self, t):
        return 1 - t, -1

    def plan(self, t, x0, x1):          # path.py:112-135
        te = expand_t_like_x(t, x1)
        alpha_t, d_alpha_t = self.compute_alpha_t(te)
        sigma_t, d_sigma_t = self.compute_sigma_t(te)
        xt = alpha_t * x1 + sigma_t * x0
        ut = d_alpha_t * x1 + d_sigma_t * x0
        return t, xt, ut


class GVPCPlan(ICPlan):
    """Trigonometric interpolant (path.py:173-190): alpha = sin(pi t / 2), sigma = cos(pi t / 2)."""

    def compute_alpha_t(self, t):
        return torch.sin(t * math.pi / 2), math.pi / 2 * torch.cos(t * math.pi / 2)

    def compute_sigma_t(self, t):
        return torch.cos(t * math.pi / 2), -math.pi / 2 * torch.sin(t * math.pi / 2)


def mean_flat(x, mask):
    """transport.py:12-16: masked mean over all non-batch dimensions."""
    dims = list(range(1, x.dim()))
    return torch.sum(x * mask, dim=dims) / torch.sum(mask, dim=dims)


def t_to_alpha(t, args):
    """transport.py:52-57."""
    return 1 * (1 - t) + t * args.alpha_max, (args.alpha_max - 1)


class Transport:
    """transport.py:60-222 for velocity models on Linear / GVP paths: interpolant + training loss (training_losses) and the
    sampling interval (check_interval, transport.py:94-123).  The model call inside training_losses is the only heavy
    step; with mdgen_finetune_b200.LatentMDGenModel it runs on the CUDA path with gradients (autograd.VelocityFn)."""

    def __init__(self, *, args, model_type="velocity", path_type="Linear", loss_type=None,
                 train_eps=0, sample_eps=0):
        self.args = args
        self.model_type = model_type
        self.path_type = path_type
        self.loss_type = loss_type
        self.train_eps = train_eps
        self.sample_eps = sample_eps
        self.path_sampler = GVPCPlan() if path_type == "GVP" else ICPlan()

    def sample(self, x1):
        """transport.py:125-135: x0 ~ N(0, 1) like x1, t ~ U(t0, t1) per batch element (drawn on the CPU generator like the
        reference's th.rand, then moved)."""
        x0 = torch.randn_like(x1)
        t0, t1 = self.check_interval(self.train_eps, self.sample_eps)
        t = torch.rand((x1.shape[0],)) * (t1 - t0) + t0
        return t.to(x1), x0, x1

    def training_losses(self, model, x1, aatype1=None, mask=None, model_kwargs=None):
        """transport.py:137-222 (velocity branch; the mpnn / dynamic_mpnn variants are out of scope):
        masked MSE between the model output and u_t; design adds the Dirichlet-noised sequence channels to x_t and a
        cross-entropy on the logits."""
        if model_kwargs is None:
            model_kwargs = {}
        args = self.args
        t, x0, x1 = self.sample(x1)
        t, xt, ut = self.path_sampler.plan(t, x0, x1)
        design = bool(getattr(args, "design", False))
        if design:
            seq_one_hot = torch.nn.functional.one_hot(aatype1, num_classes=20)
            alphas, _ = t_to_alpha(t, args)
            alphas = 1 + seq_one_hot * (alphas[:, None, None] - 1)
            x_d = torch.distributions.Dirichlet(alphas).sample()
            x_d = x_d[:, None, :, :].expand(-1, xt.shape[1], -1, -1)
            xt = torch.cat([xt, x_d], dim=-1)
        model_output = model(xt, t, **model_kwargs)
        B, *_, C = xt.shape
        assert model_output.size() == (B, *xt.size()[1:-1], C)
        if design:
            logits = model_output[:, :, :, -20:]
            model_output = model_output[:, :, :, :-20]
        terms = {"t": t, "pred": model_output}
        terms["loss"] = mean_flat((model_output - ut) ** 2, mask)
        if design:
            terms["loss_continuous"] = terms["loss"]
            seq_expanded = aatype1[:, None, :].expand(-1, xt.shape[1], -1)
            loss_d = torch.nn.functional.cross_entropy(logits.reshape(-1, 20), seq_expanded.reshape(-1))
            w = args.discrete_loss_weight
            terms["loss"] = loss_d * w + (1 - w) * terms["loss"]
            terms["loss_discrete"] = loss_d
            terms["logits"] = logits
        return terms

    def check_interval(self, train_eps, sample_eps, *, diffusion_form="SBDM", sde=False, reverse=False,
                       eval=False, last_step_size=0.0):
        t0, t1 = 0, 1          # velocity + Linear/GVP: stable everywhere (transport.py:106-118)
        if reverse:
            t0, t1 = 1 - t0, 1 - t1
        return t0, t1

    def get_drift(self):
        def body_fn(x, t, model, **model_kwargs):      # transport.py:241-257 (velocity_ode)
            out = model(x, t, **model_kwargs)
            assert out.shape == x.shape, "Output shape from ODE solver must match input shape"
            return out
        return body_fn


def create_transport(args, path_type="Linear", prediction="velocity", loss_weight=None, train_eps=None,
                     sample_eps=None):
    """transport.py:513-574.  Only prediction='velocity' with a Linear/GVP path is on the hot path."""
    if prediction != "velocity" or path_type not in ("Linear", "GVP"):
        raise NotImplementedError("only velocity prediction on Linear/GVP paths is built (DESIGN.md scope)")
    return Transport(args=args, model_type="velocity", path_type=path_type, loss_type=loss_weight,
                     train_eps=0, sample_eps=0)


def _unwrap_native(model):
    """(LatentMDGenModel, bound kwargs) if `model` is its forward/forward_inference, possibly wrapped
    in functools.partial (wrapper.py:477); else (None, None)."""
    from .model import LatentMDGenModel
    kw = {}
    fn = model
    while isinstance(fn, functools.partial):
        if fn.args:
            return None, None
        kw = {**fn.keywords, **kw}
        fn = fn.func
    owner = getattr(fn, "__self__", None)
    if isinstance(fn, LatentMDGenModel):
        owner = fn
    elif not (isinstance(owner, LatentMDGenModel) and fn.__name__ in ("forward", "forward_inference", "__call__")):
        return None, None
    if owner.args.design and getattr(fn, "__name__", "") != "forward_inference":
        return None, None            # design: only forward_inference carries the Dirichlet flow the fused rollout applies
    return owner, kw


class Sampler:
    """transport.py:278-292."""

    def __init__(self, transport):
        self.transport = transport
        self.drift = transport.get_drift()

    def sample_ode(self, *, sampling_method="dopri5", num_steps=50, atol=1e-6, rtol=1e-3, reverse=False):
        if sampling_method != "euler":
            raise NotImplementedError("only the fixed-grid Euler sampler is built (SURVEY Q3); "
                                      f"got sampling_method={sampling_method!r}")
        t0, t1 = self.transport.check_interval(self.transport.train_eps, self.transport.sample_eps,
                                               sde=False, eval=True, reverse=reverse, last_step_size=0.0)
        grid = torch.linspace(0, 1, num_steps)            # integrators.py:90

        @torch.no_grad()
        def _sample(x, model, **model_kwargs):
            native, bound = _unwrap_native(model)
            if native is not None and not reverse:
                return native.euler_rollout(x, num_steps, return_trajectory=True, **{**bound, **model_kwargs})
            t = grid.to(x.device)
            ys = [x]
            y = x
            for i in range(num_steps - 1):               # torchdiffeq fixed-grid Euler
                tv = torch.ones(x.size(0), device=x.device) * t[i]      # integrators.py:98-99
                if reverse:
                    tv = torch.ones_like(tv) * (1 - tv)
                y = y + (t[i + 1] - t[i]) * self.drift(y, tv, model, **model_kwargs)
                ys.append(y)
            return torch.stack(ys, 0)
        return _sample
