#!/bin/bash
# GPU box: attention kernel tests, probe timing, then (optionally) the whole suite + bench.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_kernels.py -q -m gpu --timeout 60 -x -k attention > gpurun_out/t_attn.log 2>&1; rc=$?; echo "attn tests rc=$rc"; tail -n 12 gpurun_out/t_attn.log
[ $rc -ne 0 ] && exit 1
timeout 40 python tools/attn_probe.py > gpurun_out/attn_probe.log 2>&1; rc=$?; echo "probe rc=$rc"; head -c 1500 gpurun_out/attn_probe.log
[ $rc -ne 0 ] && exit 1
SKIP_NCU=${SKIP_NCU:-1} bash tools/gpu_round.sh
