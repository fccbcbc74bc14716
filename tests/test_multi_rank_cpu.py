"""N > 1 host logic on CPU (gloo, world_size 2): the sampling path shards by independent rollouts -- one process per
GPU, different seeds per rank, replicated weights, NO data-path collective; the only exchange is the MAX reduction of
the step time, and only rank 0 prints the JSON line."""
import json
import os
import subprocess
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import bench
    from mdgen_finetune_b200 import configs as cfgs
    dev = torch.device("cpu")
    bench.dist_barrier(world, dev)
    slowest = bench.dist_max(10.0 + 5.0 * rank, world, dev)
    c = cfgs.small_variant(3, B=2, T=8, L=4)
    c["id"] = 3
    x0 = bench.rank_inputs(c, rank)["x0"]
    gathered = [torch.empty_like(x0) for _ in range(world)]
    dist.all_gather(gathered, x0)                       # test-only exchange: the product path never gathers
    q.put((rank, slowest, float((gathered[0] - gathered[1]).abs().max()), tuple(x0.shape)))
    dist.destroy_process_group()


def test_rank_sharding_and_time_reduction_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29000 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, slowest, diff, shape in res:
        assert slowest == 15.0                          # every rank reports the slowest rank's time
        assert diff > 0.1                               # the ranks sample different rollouts
        assert shape == (2, 8, 4, 21)                   # per-rank work is fixed: weak scaling


def test_reference_arm_under_torchrun_prints_once():
    """`bench.py --impl reference` launched like the driver launches it for N = 2: rank 0 alone runs the CPU arm and
    prints one JSON line, the other rank exits 0 without work."""
    env = dict(os.environ, OMP_NUM_THREADS="4")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(29500 + os.getpid() % 400), os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2",
           "--steps", "1", "--warmup", "0", "--config", "2"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = lines[0]
    assert d["impl"] == "reference" and d["n_gpus"] == 2 and d["value"] > 0
    # the reference's own modules (staged under oracle/_ref by __graft_entry__.build(), or /root/reference/src here)
    assert d["cpu_baseline"]["kind"] == "reference" and d["e2e"]["h2d_bytes_per_step"] == 0
    assert d["cpu_baseline"]["same_config"] in (True, False) and d["cpu_baseline"]["cores"] >= 1


def test_strong_scaling_shards_partition_the_batch():
    """--scaling strong: the config's ONE batch is cut contiguously over the ranks; the union of the shards is the batch."""
    from mdgen_finetune_b200 import configs as cfgs
    cf = dict(cfgs.CONFIGS[3]); cf["id"] = 3
    cf["T"] = 20
    full = cfgs.synthetic_inputs(cf, seed=0)
    for world in (2, 4, 8):
        per = cf["B"] // world
        shards = [{k: v[r * per:(r + 1) * per] for k, v in full.items()} for r in range(world)]
        for k in full:
            assert torch.equal(torch.cat([s[k] for s in shards]), full[k])
        assert all(s["x0"].shape[0] == per for s in shards)


def _grad_worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from mdgen_finetune_b200 import GradAllReduce, LatentMDGenModel, all_reduce_mean_scalar, configs as cfgs
    args = cfgs.make_args(cfgs.small_variant(4, B=2, T=1, L=4))
    torch.manual_seed(0)
    m = LatentMDGenModel(args, 48)
    # LoRA-style: only part of the model is trainable; some trainable parameters get no gradient at all
    m.requires_grad_(False)
    train = [p for n, p in m.named_parameters() if ".fc1." in n or "bias_k" in n or "head_weights" in n]
    for p in train:
        p.requires_grad_(True)
    g = torch.Generator().manual_seed(100 + rank)
    for i, p in enumerate(train):
        if i % 7 != 3:                                   # the same subset on every rank, as under DDP
            p.grad = torch.randn(p.shape, generator=g)
    mine = {i: p.grad.clone() for i, p in enumerate(train) if p.grad is not None}
    red = GradAllReduce(m.parameters(), bucket_mb=1.0)
    nb = len(red.buckets())
    nbytes = red.sync()
    logz = all_reduce_mean_scalar(torch.tensor(float(rank + 1)))
    # plain numpy copies: tensors would travel as shared-memory handles that die with this process
    q.put((rank, nb, nbytes, {i: p.grad.numpy().copy() for i, p in enumerate(train) if p.grad is not None},
           {i: v.numpy().copy() for i, v in mine.items()}, float(logz)))
    dist.destroy_process_group()


def test_gradient_allreduce_gloo_world2():
    """RTB data parallelism (gfn_diffusion.py:154-161, 213-219): after sync() every rank holds the MEAN of the ranks'
    gradients, bucketed; parameters without a gradient are skipped; the vargrad scalar is averaged too."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31000 + os.getpid() % 2000
    procs = [ctx.Process(target=_grad_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted((q.get(timeout=180) for _ in range(2)), key=lambda r: r[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    (_, nb0, bytes0, red0, mine0, z0), (_, nb1, bytes1, red1, mine1, z1) = res
    assert nb0 == nb1 and nb0 > 1 and bytes0 == bytes1 > 0          # several buckets at bucket_mb = 1
    assert set(red0) == set(mine0) == set(mine1)
    import numpy as np
    for i in mine0:
        mean = (mine0[i] + mine1[i]) / 2
        assert np.allclose(red0[i], mean, atol=1e-6) and np.array_equal(red0[i], red1[i])
    assert z0 == z1 == 1.5
