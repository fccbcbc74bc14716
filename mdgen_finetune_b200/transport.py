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

import torch


class Transport:
    """transport.py:60-83 (state only) + check_interval for velocity models (transport.py:94-123)."""

    def __init__(self, *, args, model_type="velocity", path_type="Linear", loss_type=None,
                 train_eps=0, sample_eps=0):
        self.args = args
        self.model_type = model_type
        self.path_type = path_type
        self.loss_type = loss_type
        self.train_eps = train_eps
        self.sample_eps = sample_eps

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
