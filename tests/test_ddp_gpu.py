"""RTB data-parallel fine-tuning across GPUs (BASELINE config 4: "NCCL grad allreduce"): two processes, one per GPU, the
batch split over the ranks, gradients averaged over NCCL -- (a) through torch's DistributedDataParallel around the module
(the parameters are real leaves, autograd.VelocityFn feeds DDP's hooks), (b) through mdgen_finetune_b200.GradAllReduce
(the reference's `accelerator.backward` site, gfn_diffusion.py:213-219).  Both must equal the gradient of the mean loss over
the WHOLE batch computed in one process.  Needs 2 GPUs: `gpurun --gpus 2 -- python -m pytest tests/test_ddp_gpu.py -m gpu`."""
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def _case():
    from mdgen_finetune_b200 import configs as cfgs
    from mdgen_finetune_b200.weights import random_state_dict
    c = cfgs.small_variant(4, B=4, T=3, L=4)
    c["id"] = 4
    args = cfgs.make_args(c)
    return c, args, random_state_dict(args, c["D"], seed=0), cfgs.synthetic_inputs(c, seed=0)


def _loss_and_grads(m, inp, sl, dev):
    g = lambda v: v[sl].to(dev)
    kw = dict(mask=g(inp["mask"]), start_frames=(g(inp["start_rot"]), g(inp["start_trans"])),
              end_frames=(g(inp["end_rot"]), g(inp["end_trans"])), x_cond=g(inp["x_cond"]),
              x_cond_mask=g(inp["x_cond_mask"]), aatype=g(inp["aatype"]))
    x = g(inp["x0"])
    t = torch.linspace(0.2, 0.8, inp["x0"].shape[0])[sl].to(dev)
    y = m(x, t, **kw)
    return (y ** 2).mean()


def _worker(rank, world, port, mode, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch.distributed as dist
    dev = torch.device("cuda", rank)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    from mdgen_finetune_b200 import GradAllReduce, LatentMDGenModel
    c, args, sd, inp = _case()
    m = LatentMDGenModel(args, c["D"], precision="fp32", quat_sign="eigh_cpu")
    m.load_state_dict(sd)
    m = m.to(dev).train()
    per = c["B"] // world
    sl = slice(rank * per, (rank + 1) * per)
    if mode == "ddp":
        ddp = torch.nn.parallel.DistributedDataParallel(m, device_ids=[rank])
        loss = _loss_and_grads(ddp, inp, sl, dev)
        loss.backward()
    else:
        loss = _loss_and_grads(m, inp, sl, dev)
        loss.backward()
        nbytes = GradAllReduce(m.parameters()).sync()
        assert nbytes > 100e6                         # 34.6 M fp32 parameters
    torch.cuda.synchronize(dev)
    probe = {n: (float(p.grad.double().norm()), float(p.grad.double().flatten()[::97].sum()))
             for n, p in m.named_parameters() if p.grad is not None}
    q.put((rank, float(loss.detach()), probe))
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.is_available() and torch.cuda.device_count() < 2, reason="needs two GPUs (gpurun --gpus 2)")
@pytest.mark.parametrize("mode", ["ddp", "grad_allreduce"])
def test_two_rank_gradients_equal_single_process_full_batch(mode):
    import torch.multiprocessing as mp
    from mdgen_finetune_b200 import LatentMDGenModel
    c, args, sd, inp = _case()
    dev = torch.device("cuda", 0)
    m = LatentMDGenModel(args, c["D"], precision="fp32", quat_sign="eigh_cpu")
    m.load_state_dict(sd)
    m = m.to(dev).train()
    loss = _loss_and_grads(m, inp, slice(0, c["B"]), dev)
    loss.backward()
    ref = {n: (float(p.grad.double().norm()), float(p.grad.double().flatten()[::97].sum()))
           for n, p in m.named_parameters() if p.grad is not None}
    del m
    torch.cuda.empty_cache()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29700 + os.getpid() % 200
    procs = [ctx.Process(target=_worker, args=(r, 2, port, mode, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted((q.get(timeout=600) for _ in range(2)), key=lambda r: r[0])
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    # mean of the per-rank mean losses = full-batch mean loss (equal shard sizes)
    assert abs((res[0][1] + res[1][1]) / 2 - float(loss.detach())) < 1e-5 * max(1.0, abs(float(loss.detach())))
    assert set(res[0][2]) == set(res[1][2]) == set(ref)
    for n, (norm, probe) in ref.items():
        for r in res:
            gn, gp = r[2][n]
            assert abs(gn - norm) <= 1e-4 * norm + 1e-12, (mode, n, gn, norm)
            assert abs(gp - probe) <= 1e-3 * norm + 1e-12, (mode, n, gp, probe)
        assert res[0][2][n] == res[1][2][n] or abs(res[0][2][n][0] - res[1][2][n][0]) <= 1e-6 * norm
