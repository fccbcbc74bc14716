#!/usr/bin/env python
"""bench.py -- sampled frames/sec of 4AA 1000-frame Euler rollouts (BASELINE.json metric).

A "step" = one full 49-evaluation Euler sampler call (Sampler.sample_ode('euler', num_steps=50),
SURVEY Q3) over one batch of config 3 (B=8 upsampling windows x T=1000 frames x L=4) per GPU.
  value : frames/s with x0 + conditioning already resident in HBM (whole job, all ranks)
  e2e   : same metric through the public Sampler API with pinned-HOST inputs, H2D/D2H in the timed region
  --impl reference : the CPU oracle port (oracle/mdgen_oracle.py = restated reference) on the host cores
"""
import argparse
import functools
import json
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
from mdgen_finetune_b200 import configs as cfgs  # noqa: E402
from mdgen_finetune_b200.weights import random_state_dict  # noqa: E402

NUM_STEPS = 50          # linspace(0,1,50) -> 49 velocity evaluations (wrapper.py:472, integrators.py:90)
METRIC = "sampled_frames_per_sec"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        j = json.load(open(p))
        return dict(tflops=j["bf16_tflops_sustained"], tflops_burst=j["bf16_tflops"], hbm=j["hbm_gbs"], src="measured")
    return dict(tflops=1590.0, tflops_burst=1590.0, hbm=6650.0, src="fallback")


class ClockSampler(threading.Thread):
    def __init__(self, idx):
        super().__init__(daemon=True)
        self.idx, self.stop_flag, self.rows = idx, False, []

    def run(self):
        # NVML in-process: forking nvidia-smi out of a process with a CUDA context every 200 ms stalled the launching
        # thread for milliseconds at a time (one device-arm measurement came out 10 % slow); nvidia-smi is the fallback
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.idx)
            mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            bits = (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20), ("sw_power_cap", 0x4))
            while not self.stop_flag:
                sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                try:
                    r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
                except Exception:
                    r = 0
                self.rows.append([str(sm), str(mx)] + ["Active" if r & bit else "Not Active" for _, bit in bits])
                time.sleep(0.1)
            return
        except Exception:
            pass
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        while not self.stop_flag:
            try:
                o = subprocess.run(["nvidia-smi", "-i", str(self.idx), f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                                   capture_output=True, text=True, timeout=5).stdout.strip()
                if o:
                    self.rows.append([c.strip() for c in o.split(",")])
            except Exception:
                pass
            time.sleep(0.5)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(int(r[0]) for r in self.rows if r[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[2 + i] == "Active" for r in self.rows if len(r) > 2 + i)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": int(self.rows[0][1]) if self.rows[0][1].isdigit() else None,
                "reasons": reasons, "samples": len(self.rows)}


_REF_CACHE = {}


def reference_forward_fn(cfg_id, B):
    """The REFERENCE'S OWN LatentMDGenModel (unmodified sources: /root/reference/src in the build container, the copy
    staged by oracle/stage_ref.py under oracle/_ref/src on the GPU box) on the host cores, fp32, all threads: returns
    (fn(t_scalar) -> velocity, config dict, threads).  Falls back to the oracle port only if the sources are missing."""
    key = (cfg_id, B)
    if key in _REF_CACHE:
        return _REF_CACHE[key]
    c = dict(cfgs.CONFIGS[cfg_id]); c["B"] = B; c["id"] = cfg_id
    args = cfgs.make_args(c)
    sd = random_state_dict(args, c["D"], seed=0)
    inp = cfgs.synthetic_inputs(c, seed=0)
    torch.set_num_threads(os.cpu_count() or 1)
    try:
        from oracle.ref_loader import load_reference
        ref = load_reference()
        model = ref.LatentMDGenModel(args, c["D"])
        model.load_state_dict(sd, strict=True)
        model.eval()
        rigid = lambda r, t: ref.Rigid(ref.Rotation(rot_mats=r), t)
        kw = dict(mask=inp["mask"], start_frames=rigid(inp["start_rot"], inp["start_trans"]), x_cond=inp["x_cond"],
                  x_cond_mask=inp["x_cond_mask"], aatype=inp["aatype"])
        if "end_rot" in inp:
            kw["end_frames"] = rigid(inp["end_rot"], inp["end_trans"])
        f = lambda tv: model.forward_inference(inp["x0"], torch.full((B,), tv), **kw)
        kind = "reference"
    except Exception as e:      # sources not staged: the restated port (oracle/mdgen_oracle.py), and say so
        sys.stderr.write(f"bench.py: reference sources unavailable ({e}); timing the oracle port instead\n")
        from oracle import mdgen_oracle as O
        start = (inp["start_rot"], inp["start_trans"])
        end = (inp["end_rot"], inp["end_trans"]) if "end_rot" in inp else None
        f = lambda tv: O.forward_inference(sd, args, inp["x0"], torch.full((B,), tv), inp["mask"], start, end, inp["x_cond"],
                                           inp["x_cond_mask"], inp["aatype"])
        kind = "port"
    _REF_CACHE[key] = (f, c, torch.get_num_threads(), kind)
    return _REF_CACHE[key]


def reference_time_forward(cfg_id, B, nfwd=3, warm=1):
    """Median seconds per velocity evaluation of the reference on the host cores at batch B."""
    f, c, cores, kind = reference_forward_fn(cfg_id, B)
    ts = []
    with torch.no_grad():
        for _ in range(warm):
            f(0.0)
        for i in range(nfwd):
            t0 = time.perf_counter()
            f(i / 49.0)
            ts.append(time.perf_counter() - t0)
    ts.sort()
    return ts[len(ts) // 2], c, cores, kind


def ncu_traffic(category, cfg_id):
    """DRAM bytes (read + write) per launch of the dominant kernel family from the committed `ncu --set full` capture
    (profiles/r02/traffic_cfg<id>.json, else r01; written by tools/ncu_traffic.py); None when no capture exists for this config."""
    names = {"gemm_tcgen05": ("gemm_tc_kernel", "gemm_tc2_kernel", "gemm_ln_kernel", "gemm_mlp_ln_kernel"), "attention_tcgen05": ("attention_tc_kernel",),
             "attention_short": ("attention_seq4_bulk_kernel", "attention_seq_kernel", "attention_short_kernel"),
             "layernorm_modulate": ("ln_modulate_kernel",), "embed": ("embed_step_kernel", "embed_kernel"),
             "qkv_attention_fused": ("gemm_qkv_attn4_kernel",)}
    t = None
    for rnd in ("r02", "r01"):          # the latest committed capture
        try:
            t = json.load(open(os.path.join(ROOT, "profiles", rnd, f"traffic_cfg{cfg_id}.json")))
            break
        except (OSError, ValueError):
            continue
    if t is None:
        return None
    tot = n = 0
    for k in names.get(category, ()):
        if k in t:
            tot += t[k]["dram_bytes_per_launch"] * t[k]["launches"]
            n += t[k]["launches"]
    return tot / n if n else None


def algorithmic_flops(c):
    """Per-category algorithmic FLOPs of ONE velocity evaluation (BASELINE.md section 3 formula)."""
    B, T, L, D = c["B"], c["T"], c["L"], c["D"]
    C, F = 384, 1536
    N = B * T * L
    M = B * L * (2 if c["task"] in ("tps", "design") else 1)
    gemm = 5 * 32 * N * C * C + 2 * N * C * D + 5 * (8 * M * C * C + 4 * M * C * F)
    if c["task"] == "design":
        gemm += 2 * N * C * C
    qkv_l = 5 * 6 * N * C * C                      # the residue-axis q/k/v projections of the 5 main layers
    res_main = 5 * 4 * B * T * L * (L + 1) * C     # residue-axis attention of the main layers
    return {"gemm_tcgen05": float(gemm), "attention_frame": float(5 * 4 * B * L * T * (T + 1) * C),
            "attention_residue": float(res_main + 5 * 4 * M * (L + 1) * C),
            # gemm_qkv_attn4_kernel (L = 4): projection + 4 x 5 attention in one launch; its FLOPs leave the GEMM family
            "qkv_attention_fused": float(qkv_l + res_main), "qkv_l": float(qkv_l)}


def run_reference(a, rank, world):
    """--impl reference: the reference's own CPU implementation of the path (its unmodified LatentMDGenModel.forward_
    inference, staged under oracle/_ref) on all host threads.  Rank 0 only.  A step = ONE velocity evaluation of the
    workload's full batch when the whole run then fits in a few minutes, else of a B = 1 slice; the metric extrapolates
    the 49 evaluations of a rollout from it (a full reference rollout of config 3 takes minutes per step)."""
    if rank != 0:
        return
    cfg = cfgs.CONFIGS[a.config]
    Bfull = cfg["B"]
    nev = NUM_STEPS - 1
    budget_s = 200.0
    t1, _, cores, kind = reference_time_forward(a.config, 1, nfwd=1, warm=1)
    full_s = None
    B = 1
    if Bfull > 1 and t1 * Bfull * 1.5 * (a.steps + 2) < budget_s:
        full_s, _, _, _ = reference_time_forward(a.config, Bfull, nfwd=1, warm=0)
        if full_s * (a.steps + 1) < budget_s:
            B = Bfull
    elif Bfull == 1:
        B = 1
    secs = []
    for _ in range(max(0, min(a.warmup, 2) - 1)):
        reference_time_forward(a.config, B, nfwd=1, warm=0)
    for _ in range(a.steps):
        s, c, cores, kind = reference_time_forward(a.config, B, nfwd=1, warm=0)
        secs.append(s)
    s = sum(secs) / len(secs)
    value = B * cfg["T"] / (nev * s)
    sample = (f"1 velocity evaluation of {'the full batch' if B == Bfull else 'a B=1 slice'} (B={B} x {cfg['T']} frames x {cfg['L']} "
              f"residues) per step on the reference's own LatentMDGenModel; x{nev} extrapolated to a full Euler call")
    cb = {"value": value, "unit": "frames/s", "cores": cores, "kind": kind, "sample": sample, "batch": B,
          "same_config": B == Bfull}
    if full_s is not None:
        cb["full_batch_forward_s"] = full_s
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": s * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": {"workload": workload_name(a.config)},
        "cpu_baseline": cb,
        "e2e": {"value": value, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def workload_name(cfg_id):
    c = cfgs.CONFIGS[cfg_id]
    return (f"config {cfg_id}: {c['name']}, B={c['B']} x T={c['T']} frames x L={c['L']} residues, D={c['D']}, "
            f"{NUM_STEPS - 1}-evaluation fixed-grid Euler rollout per step")


def dist_barrier(world, dev):
    """Barrier over the ranks (one process per GPU; NCCL on the GPU box, gloo in the CPU tests) + device sync."""
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    if dev.type == "cuda":
        torch.cuda.synchronize(dev)


def dist_max(ms, world, dev):
    """Step time of the job = the slowest rank's (the only collective of the sampling path: rollouts are independent)."""
    if world > 1:
        import torch.distributed as dist
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
    return ms


def rank_inputs(c, rank):
    """Weak scaling: every rank samples its own batch of independent rollouts (seed = rank), weights are replicated."""
    return cfgs.synthetic_inputs(c, seed=rank)


def build_model(cfg_id, dev, B=None, precision="bf16"):
    from mdgen_finetune_b200 import LatentMDGenModel
    c = dict(cfgs.CONFIGS[cfg_id]); c["id"] = cfg_id
    if B is not None:
        c["B"] = B
    args = cfgs.make_args(c)
    sd = random_state_dict(args, c["D"], seed=0)
    model = LatentMDGenModel(args, c["D"], precision=precision)
    model.load_state_dict(sd)
    return model.to(dev).eval().requires_grad_(False), c, args


def kwargs_from(inp, f):
    kw = dict(mask=f(inp["mask"]), start_frames=(f(inp["start_rot"]), f(inp["start_trans"])),
              x_cond=f(inp["x_cond"]), x_cond_mask=f(inp["x_cond_mask"]), aatype=f(inp["aatype"]))
    if "end_rot" in inp:
        kw["end_frames"] = (f(inp["end_rot"]), f(inp["end_trans"]))
    return kw


def time_device_rollouts(model, x0, kw, steps, warm, world, dev):
    """ms per rollout (device-resident inputs), CUDA events on the launching stream, barrier + sync on both sides, MAX over ranks."""
    for _ in range(warm):
        model.euler_rollout(x0, NUM_STEPS, **kw)
    dist_barrier(world, dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        model.euler_rollout(x0, NUM_STEPS, **kw)
    e1.record()
    dist_barrier(world, dev)
    return dist_max(e0.elapsed_time(e1) / steps, world, dev)


def other_config_lines(a, rank, world, dev, pk):
    """north_star: throughput on tetrapeptide AND ATLAS-shaped trajectories at every GPU count.  Configs 1, 2, 4, 5 at
    their full BASELINE shapes, measured in the same run with the same timing rules (device-resident arm, weak scaling:
    every rank its own rollouts), fewer steps."""
    out = {}
    for cid in (1, 2, 4, 5):
        if cid == a.config:
            continue
        model, c, _ = build_model(cid, dev)
        inp = cfgs.synthetic_inputs(c, seed=rank)
        kw = kwargs_from(inp, lambda v: v.to(dev))
        x0 = inp["x0"].to(dev)
        ms = time_device_rollouts(model, x0, kw, max(3, a.steps // 4), 3, world, dev)
        fps = world * c["B"] * c["T"] / (ms * 1e-3)
        out[f"config{cid}"] = {"workload": workload_name(cid), "value": fps, "unit": "frames/s", "ms_per_step": ms,
                               "tensor_core_frac": cfgs.flops_per_forward(c) * (NUM_STEPS - 1) / (ms * 1e-3) / 1e12 / pk["tflops"],
                               "launches_per_evaluation": model.launches_per_forward()}
        del model
        torch.cuda.empty_cache()
    return out


def run_gpu(a, rank, world, local_rank):
    from mdgen_finetune_b200 import Sampler, create_transport, _lib
    import torch.distributed as dist
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    strong = a.scaling == "strong"
    Bfull = cfgs.CONFIGS[a.config]["B"]
    if strong and Bfull % world != 0:
        raise SystemExit(f"--scaling strong shards the batch of {Bfull} over the ranks: --gpus must divide it")
    model, c, args = build_model(a.config, dev, B=Bfull // world if strong else None)
    # weak: every rank samples its own full batch (seed = rank).  strong: the ONE batch of the config (seed 0) is sharded
    # contiguously over the ranks.  Either way rollouts are independent: no data-path collective.
    if strong:
        cf = dict(cfgs.CONFIGS[a.config]); cf["id"] = a.config
        full = cfgs.synthetic_inputs(cf, seed=0)
        per = Bfull // world
        inp = {k: v[rank * per:(rank + 1) * per].contiguous() for k, v in full.items()}
    else:
        inp = rank_inputs(c, rank)
    kwargs_on = lambda f: kwargs_from(inp, f)

    barrier = lambda: dist_barrier(world, dev)
    max_over_ranks = lambda ms: dist_max(ms, world, dev)

    frames_per_step = c["B"] * c["T"]
    # ---------------- device-resident arm ----------------
    kw_dev = kwargs_on(lambda v: v.to(dev))
    x0_dev = inp["x0"].to(dev)
    rollout = lambda: model.euler_rollout(x0_dev, NUM_STEPS, **kw_dev)   # design mode: Dirichlet flow applied per step inside
    clocks = ClockSampler(local_rank)
    clocks.start()                  # NVML initialises during the warm-up, not inside the timed region
    # warm-up: the W requested steps, and at least ~1.5 s of load -- on a fresh box the first process measured 3 % slow
    # with only W = 3 rollouts (0.34 s) of warm-up (clock / power state ramp, first-touch of the workspace)
    t_warm = time.perf_counter()
    n_warm = 0
    while n_warm < a.warmup or time.perf_counter() - t_warm < 1.5:
        rollout()
        torch.cuda.synchronize(dev)
        n_warm += 1
    clocks.rows.clear()             # keep only the samples taken under the timed load
    # parity of the timed dispatch, once, after warm-up: finite, and one bf16 velocity against the fp32 check mode of the
    # same library (itself pinned to the reference at 1e-4 by tests/test_gpu_parity.py) within north_star's 2e-2
    t_chk = torch.full((c["B"],), 0.3, device=dev)
    v_bf16 = model.forward_inference(x0_dev, t_chk, **kw_dev)
    m32, _, _ = build_model(a.config, dev, B=c["B"], precision="fp32")
    v_f32 = m32.forward_inference(x0_dev, t_chk, **kw_dev)
    parity = float((v_bf16.double() - v_f32.double()).norm() / v_f32.double().norm())
    del m32, v_f32
    torch.cuda.empty_cache()
    if not a.no_parity_check and not (torch.isfinite(v_bf16).all() and torch.isfinite(rollout()).all() and parity < 2e-2):
        raise SystemExit(f"bench.py: parity check failed (rel-L2 bf16 vs fp32 check mode = {parity})")
    rollout()
    torch.cuda.synchronize(dev)
    l0 = _lib.lib().mdgen_launch_count()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        out = rollout()
    e1.record()
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1) / a.steps)
    launches = int(_lib.lib().mdgen_launch_count() - l0)
    # ---------------- end-to-end arm: public Sampler API, pinned host inputs ----------------
    pin = lambda v: v.pin_memory()
    host = {k: pin(v) for k, v in inp.items()}
    out_host = torch.empty_like(inp["x0"]).pin_memory()
    sample_fn = Sampler(create_transport(args)).sample_ode(sampling_method="euler", num_steps=NUM_STEPS)
    h2d = sum(v.numel() * v.element_size() for v in host.values())
    d2h = out_host.numel() * out_host.element_size()

    def e2e_step():
        f = lambda v: v.to(dev, non_blocking=True)
        kw = dict(mask=f(host["mask"]), start_frames=(f(host["start_rot"]), f(host["start_trans"])),
                  x_cond=f(host["x_cond"]), x_cond_mask=f(host["x_cond_mask"]), aatype=f(host["aatype"]))
        if "end_rot" in host:
            kw["end_frames"] = (f(host["end_rot"]), f(host["end_trans"]))
        traj = sample_fn(f(host["x0"]), functools.partial(model.forward_inference, **kw))
        out_host.copy_(traj[-1], non_blocking=True)
        torch.cuda.current_stream(dev).synchronize()

    for _ in range(max(3, a.warmup // 2)):
        e2e_step()
    barrier()
    e0.record()
    for _ in range(a.steps):
        e2e_step()
    e1.record()
    barrier()
    ms_e2e = max_over_ranks(e0.elapsed_time(e1) / a.steps)
    clocks.stop_flag = True
    clocks.join(timeout=2)
    # ---------------- per-kernel-family timing (one extra profiled step) ----------------
    _lib.profile_begin()
    rollout()
    prof = _lib.profile_end()
    pk = peaks()
    others = other_config_lines(a, rank, world, dev, pk) if not a.no_other_configs else None
    strong_block = None
    if world > 1 and not strong and Bfull % world == 0 and not a.no_other_configs:
        # BASELINE config 3 read literally ("batch 8 ... data-parallel across 1/2/4/8"): the ONE batch sharded over the ranks
        ms_model, cs, _ = build_model(a.config, dev, B=Bfull // world)
        cf = dict(cfgs.CONFIGS[a.config]); cf["id"] = a.config
        full = cfgs.synthetic_inputs(cf, seed=0)
        per = Bfull // world
        sh = {k: v[rank * per:(rank + 1) * per].contiguous() for k, v in full.items()}
        ms_s = time_device_rollouts(ms_model, sh["x0"].to(dev), kwargs_from(sh, lambda v: v.to(dev)), max(3, a.steps // 2), 3, world, dev)
        strong_block = {"value": Bfull * c["T"] / (ms_s * 1e-3), "unit": "frames/s", "ms_per_step": ms_s,
                        "batch_per_gpu": per, "note": "fixed total batch sharded over the ranks (strong scaling), same timing rules"}
        del ms_model
    if rank != 0:
        return
    total_prof = sum(v[0] for v in prof.values()) or 1.0
    fl = algorithmic_flops(c)
    nev = NUM_STEPS - 1
    kernels = {k: {"ms_per_step": round(v[0], 3), "launches": v[1], "share": round(v[0] / total_prof, 4)} for k, v in prof.items()}
    dom = max(prof, key=lambda k: prof[k][0])
    fused = "qkv_attention_fused" in prof
    flop_of = {"gemm_tcgen05": fl["gemm_tcgen05"] - (fl["qkv_l"] if fused else 0.0), "attention_tcgen05": fl["attention_frame"],
               "attention_simt": fl["attention_frame"], "attention_short": fl["attention_residue"],
               "qkv_attention_fused": fl["qkv_attention_fused"]}
    roof = {"bound": "tensor", "kernel": dom, "unit": "TFLOP/s", "peak": pk["tflops"], "peak_source": pk["src"] + " (sustained cuBLAS bf16)",
            "traffic": ncu_traffic(dom, a.config)}
    if dom.startswith("attention"):
        roof["note"] = ("head_dim 24: 96 FLOP per (query, key) pair but one ex2 -- the kernel is bounded by the MUFU unit (16 ex2/clk/SM, "
                        "1536 clk per 4 x 128 x 48 block round; measured ~2450), not by the tensor pipe; DESIGN.md section 4")
    if dom in flop_of:
        t_ms, n = prof[dom]
        ach = flop_of[dom] * nev / (t_ms * 1e-3) / 1e12
        roof.update(achieved=ach, frac=ach / pk["tflops"], avg_launch_us=1e3 * t_ms / n,
                    algorithmic_gflop_per_launch=flop_of[dom] * nev / n / 1e9)
    else:
        roof.update(achieved=None, frac=None)
    for other in ("gemm_tcgen05", "attention_tcgen05", "qkv_attention_fused"):      # the other tensor-core families, same definition
        if other in prof and dom != other:
            t_ms, n = prof[other]
            g = flop_of[other] * nev / (t_ms * 1e-3) / 1e12
            roof[other] = {"achieved": g, "frac": g / pk["tflops"], "avg_launch_us": 1e3 * t_ms / n}
    if "attention_tcgen05" in prof:
        # the attention kernel's own roofline is the MUFU unit: one ex2 per (query, key) pair, 16 ex2 / clk / SM
        t_ms, n = prof["attention_tcgen05"]
        pairs = 5 * 16 * c["B"] * c["L"] * c["T"] * (c["T"] + 1)            # 5 layers x 16 heads, frame axis
        if c["L"] >= 64:
            pairs += 5 * 16 * c["B"] * c["T"] * c["L"] * (c["L"] + 1)       # ATLAS: the residue axis runs on this kernel too
        mufu_peak = 16 * 148 * 1.965e9
        roof["attention_mufu"] = {"achieved_gex2_per_s": pairs * nev / (t_ms * 1e-3) / 1e9, "peak_gex2_per_s": mufu_peak / 1e9,
                                  "frac": pairs * nev / (t_ms * 1e-3) / mufu_peak,
                                  "note": "peak = 16 ex2/clk/SM x 148 SMs x 1.965 GHz (nominal boost clock)"}
    total_flops = cfgs.flops_per_forward(c) * nev
    value = world * frames_per_step / (ms * 1e-3)
    line = {
        "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": ms, "higher_is_better": True, "scaling": "strong" if strong else "weak", "vs_baseline": None, "dtype": "bf16",
        "data": "synthetic", "gpu_launches": launches,
        "config": {"workload": workload_name(a.config), "frames_per_step_per_gpu": frames_per_step,
                   "l2": "no explicit flush: the per-evaluation activation working set (>0.4 GB) exceeds the 126 MB L2",
                   "weights": "random-init (seeded), adaLN-Zero entries re-drawn non-zero"},
        "tensor_core_frac_whole_step": total_flops / (ms * 1e-3) / 1e12 / pk["tflops"],
        "clocks": clocks.summary(),
        "e2e": {"value": world * frames_per_step / (ms_e2e * 1e-3), "unit": "frames/s", "ms_per_step": ms_e2e,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "api": "Sampler.sample_ode('euler')(x0, partial(model.forward_inference, **kw))"},
        "roofline": roof, "kernels": kernels,
        "parity_check": {"rel_l2_velocity_bf16_vs_fp32_check": parity, "tolerance": 2e-2, "finite": True},
        "rollout": "one CUDA graph per Euler call (all 49 evaluations), replayed; launches counted per graph kernel node",
    }
    if others is not None:
        line["other_configs"] = others
    if strong_block is not None:
        line["strong_scaling"] = strong_block
    if world == 1 and not a.no_cpu_baseline:
        s, cc, cores, kind = reference_time_forward(a.config, 1, nfwd=3, warm=1)
        cb = {"value": cc["T"] / (nev * s), "unit": "frames/s", "cores": cores, "kind": kind, "batch": 1,
              "sample": f"median of 3 velocity evaluations of a B=1 slice (1 of {Bfull} rollouts) on the reference's own "
                        f"LatentMDGenModel, x{nev} extrapolated"}
        if Bfull > 1 and s * Bfull < 20.0:
            sf, _, _, _ = reference_time_forward(a.config, Bfull, nfwd=1, warm=0)
            cb["full_batch_forward_s"] = sf
            cb["full_batch_value"] = Bfull * cc["T"] / (nev * sf)
        line["cpu_baseline"] = cb
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--config", type=int, default=3)
    ap.add_argument("--no-cpu-baseline", dest="no_cpu_baseline", action="store_true")
    ap.add_argument("--no-other-configs", dest="no_other_configs", action="store_true")
    ap.add_argument("--no-parity-check", dest="no_parity_check", action="store_true",
                    help="kernel timing experiments that break the results on purpose (MDGEN_ATTN_VAR); never for a reported number")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: every rank its own full batch (default); strong: the config's one batch sharded over the ranks")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    if a.impl == "reference":
        run_reference(a, rank, world)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (the product path has no CPU fallback); use --impl reference for the CPU arm")
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local_rank))
    try:
        run_gpu(a, rank, world, local_rank)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
