"""Row N4: gradients through the CUDA denoiser (what the reference gets from autograd when it trains the posterior through
`policy.unet(x, t_, **condition)` -- HGFNode.run_policy, src/models/samplers.py:120-146 -- and in
Transport.training_losses, src/mdgen/transport/transport.py:173-188).

`VelocityFn` is a recompute-based torch.autograd.Function around the C ABI: its forward is `mdgen_forward`, its backward
one `mdgen_backward` call (fp32 recompute + hand-written adjoint kernels, csrc/train*.cuh, csrc/backward*.cuh).  The
parameters enter as INPUTS of the Function under their reference state_dict names, so

* autograd accumulates into the nn.Parameters themselves: optimisers, `clip_grad_norm_`, DistributedDataParallel hooks
  (NCCL gradient all-reduce, gfn_diffusion.py:154-161,213-219) and `torch.utils.checkpoint` work unchanged;
* a LoRA-wrapped Linear (PEFT `LoraLayer` layout: `base_layer`, `lora_A[adapter]`, `lora_B[adapter]`, `scaling[adapter]`,
  gfn_diffusion.py:58-83) contributes its MERGED weight `W + scaling * B @ A` -- a small torch expression -- and autograd
  carries the weight gradient on to A and B.

No torch compute touches the activations: torch only merges / routes parameter tensors.
"""
import ctypes

import torch

from . import _lib


def effective_tensors(model, names):
    """{reference state_dict name: tensor} for the module tree as it is now: plain parameters / buffers as they are, LoRA-
    wrapped linears merged.  Differentiable w.r.t. the adapter matrices."""
    out = {}
    for name in names:
        mod = model
        parts = name.split(".")
        ok = True
        for p in parts[:-1]:
            # PEFT wraps the whole model as `base_model.model`; a wrapped child keeps its attribute name
            nxt = getattr(mod, p, None)
            if nxt is None and hasattr(mod, "base_layer"):
                nxt = getattr(mod.base_layer, p, None)
            if nxt is None:
                ok = False
                break
            mod = nxt
        if not ok:
            raise KeyError(f"parameter {name} not found in the module tree")
        leaf = parts[-1]
        if hasattr(mod, "base_layer") and hasattr(mod, "lora_A"):
            base = mod.base_layer
            t = getattr(base, leaf)
            if leaf == "weight" and not getattr(mod, "merged", False) and not getattr(mod, "disable_adapters", False):
                adapters = getattr(mod, "active_adapters", None) or list(mod.lora_A.keys())
                if isinstance(adapters, str):
                    adapters = [adapters]
                for a in adapters:
                    if a in mod.lora_A:
                        t = t + float(mod.scaling[a]) * (mod.lora_B[a].weight @ mod.lora_A[a].weight)
            out[name] = t
        else:
            t = getattr(mod, leaf, None)
            if t is None:
                raise KeyError(f"parameter {name} not found in the module tree")
            out[name] = t
    return out


class VelocityFn(torch.autograd.Function):
    """y = LatentMDGenModel.forward(x, t, **conditioning) with gradients for x and every listed parameter."""

    @staticmethod
    def forward(ctx, model, cond, names, x, t, *params):
        ctx.model, ctx.cond, ctx.names = model, cond, names
        ctx.save_for_backward(x, t)
        ctx.param_meta = [(tuple(p.shape), p.dtype) for p in params]
        with torch.no_grad():
            return model._forward_no_grad(x, t, cond)

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, dout):
        x, t = ctx.saved_tensors
        model = ctx.model
        dev = model.device
        with torch.no_grad():
            model._sync_weights()
            model._set_conditioning(x, *ctx.cond)
            xx = x.detach().to(dev, torch.float32).contiguous()
            tt = t.detach().to(dev, torch.float32).contiguous()
            dd = dout.detach().to(dev, torch.float32).contiguous()
            need_x = ctx.needs_input_grad[3]
            dx = torch.empty_like(xx) if need_x else None
            want = [i for i in range(len(ctx.names)) if ctx.needs_input_grad[5 + i]]
            sizes = [int(torch.Size(ctx.param_meta[i][0]).numel()) for i in want]
            flat = torch.zeros(sum(sizes), device=dev, dtype=torch.float32)          # one buffer, one memset
            n = len(want)
            c_names = (ctypes.c_char_p * max(n, 1))()
            c_ptrs = (ctypes.c_void_p * max(n, 1))()
            off = 0
            views = []
            for k, i in enumerate(want):
                v = flat[off:off + sizes[k]]
                off += sizes[k]
                views.append(v)
                c_names[k] = ctx.names[i].encode()
                c_ptrs[k] = v.data_ptr()
            with torch.cuda.device(dev):
                _lib.check(_lib.lib().mdgen_backward(
                    model._handle, xx.data_ptr(), tt.data_ptr(), dd.data_ptr(), None if dx is None else dx.data_ptr(), n,
                    c_names, c_ptrs, None, model._stream()), "mdgen_backward")
        grads = [None] * len(ctx.names)
        for k, i in enumerate(want):
            shape, dtype = ctx.param_meta[i]
            grads[i] = views[k].view(shape).to(dtype)
        return (None, None, None, dx.to(x.dtype) if need_x else None, None, *grads)
