#!/bin/bash
# GPU box: kernel + parity tests, then a short bench.  Outputs under gpurun_out/.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_kernels.py -q -m gpu --timeout 120 -x ${PYTEST_K:+-k "$PYTEST_K"} > gpurun_out/t_kernels.log 2>&1; echo "kernels rc=$?"; tail -n 15 gpurun_out/t_kernels.log
if [ "${SKIP_PARITY:-0}" != "1" ]; then
timeout 400 python -m pytest tests/test_gpu_parity.py -q -m gpu --timeout 200 > gpurun_out/t_parity.log 2>&1; echo "parity rc=$?"; tail -n 8 gpurun_out/t_parity.log
fi
if [ "${SKIP_BENCH:-0}" != "1" ]; then
timeout 300 python bench.py --steps 3 --warmup 3 ${BENCH_ARGS:-} > gpurun_out/bench.log 2>&1; echo "bench rc=$?"; tail -c 2500 gpurun_out/bench.log
fi
