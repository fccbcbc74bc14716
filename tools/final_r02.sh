#!/bin/bash
# final-state evidence of the round: full GPU suite, the default bench line (all configs), the ncu launch list and full capture
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/t_gpu.log 2>&1; tail -3 gpurun_out/t_gpu.log
python bench.py --steps 10 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench.err; tail -c 600 gpurun_out/bench.log; echo
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.log 2> gpurun_out/bench_ref.err; tail -c 400 gpurun_out/bench_ref.log; echo
bash tools/gpu_profile_r02.sh > gpurun_out/profile.log 2>&1; tail -30 gpurun_out/profile.log
