"""Data-parallel gradient all-reduce for RTB posterior fine-tuning (BASELINE config 4; the one collective of the path).

Reference sites: HF Accelerate / DDP around the posterior (src/rtb_utils/gfn_diffusion.py:154-161, 192, 213-219) and
Lightning DDP (src/train.py:46-69): one process per GPU, the RTB batch split over the ranks, gradients averaged with an
NCCL all-reduce once per optimiser step (34.6 M fp32 parameters = 137 MB, or the LoRA adapters only).

Two ways to get it here:
  * `torch.nn.parallel.DistributedDataParallel(model, device_ids=[rank])` works unchanged: the parameters are real leaf
    nn.Parameters and autograd.VelocityFn returns their gradients through autograd, so DDP's bucketed hooks fire;
  * `GradAllReduce` below, for loops that do not wrap the module (the reference's `accelerator.backward(loss)` site): call
    `sync()` after `loss.backward()`.  It flattens the gradients that exist into few large buckets (NVSwitch: size buckets
    for launch latency, not link count), issues one `all_reduce(SUM)` per bucket asynchronously, and scales by 1 / world.
Backend-agnostic (NCCL on the GPU box, gloo in the CPU tests); 137 MB once per step needs no fused compute + collective
kernel.
"""
import torch
import torch.distributed as dist


class GradAllReduce:
    def __init__(self, params, group=None, bucket_mb=64.0, average=True):
        self.params = [p for p in params if p.requires_grad]
        self.group = group
        self.bucket_bytes = int(bucket_mb * (1 << 20))
        self.average = average

    def world(self):
        return dist.get_world_size(self.group) if dist.is_available() and dist.is_initialized() else 1

    def buckets(self):
        """Parameters that have a gradient, grouped by (device, dtype) into buckets of ~bucket_bytes, in reverse order
        (the order backward produces them)."""
        out, cur, size, key = [], [], 0, None
        for p in reversed(self.params):
            if p.grad is None:
                continue
            k = (p.grad.device, p.grad.dtype)
            nbytes = p.grad.numel() * p.grad.element_size()
            if cur and (k != key or size + nbytes > self.bucket_bytes):
                out.append(cur)
                cur, size = [], 0
            key = k
            cur.append(p)
            size += nbytes
        if cur:
            out.append(cur)
        return out

    @torch.no_grad()
    def sync(self):
        """All-reduce (mean over the ranks) of every existing gradient.  Returns the number of bytes reduced.
        Ranks must hold gradients for the same parameters (same trainable set), as under DDP."""
        w = self.world()
        if w == 1:
            return 0
        pending, total = [], 0
        for bucket in self.buckets():
            flat = torch.cat([p.grad.reshape(-1) for p in bucket])
            total += flat.numel() * flat.element_size()
            pending.append((dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group, async_op=True), flat, bucket))
        for work, flat, bucket in pending:
            work.wait()
            if self.average:
                flat.div_(w)
            off = 0
            for p in bucket:
                n = p.grad.numel()
                p.grad.copy_(flat[off:off + n].view_as(p.grad))
                off += n
        return total


def all_reduce_mean_scalar(value, group=None):
    """The `vargrad` logZ estimate of RTB is a batch mean (gfn_diffusion.py:446-455): with the batch split over the ranks
    it needs one scalar all-reduce to keep single-GPU semantics."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return value
    v = value.detach().clone()
    dist.all_reduce(v, op=dist.ReduceOp.SUM, group=group)
    return v / dist.get_world_size(group)
