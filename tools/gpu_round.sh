#!/bin/bash
# GPU box: full gpu test suite, a short bench, then the ncu launch list of ~4 velocity evaluations.
mkdir -p gpurun_out
if [ "${SKIP_TESTS:-0}" != "1" ]; then
timeout 900 python -m pytest tests -q -m gpu --timeout 300 -x > gpurun_out/t_gpu.log 2>&1; echo "tests rc=$?"; tail -n 6 gpurun_out/t_gpu.log
fi
timeout 400 python bench.py --steps 3 --warmup 3 ${BENCH_ARGS:-} > gpurun_out/bench.log 2>&1; echo "bench rc=$?"; tail -c 3000 gpurun_out/bench.log
if [ "${SKIP_NCU:-0}" != "1" ]; then
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
ncu --metrics gpu__time_duration.sum --clock-control none -s ${SKIP:-9000} -c ${COUNT:-240} --csv \
    --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
fi
