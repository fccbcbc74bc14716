#!/bin/bash
# kernel tests + per-family times of the configs given in CONFIGS (default 3)
if [ "$1" != "quick" ]; then
  timeout 1500 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_parity.py tests/test_graph.py -m gpu -x -q 2>&1 | tail -3
fi
for c in ${CONFIGS:-3}; do
for i in 1 2; do
  env $ENVV timeout 600 python bench.py --config $c --steps 5 --warmup 3 --no-cpu-baseline --no-other-configs 2>&1 | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('config $c', round(d['value']), round(d['ms_per_step'], 2), {k: v['ms_per_step'] for k, v in d['kernels'].items() if v['ms_per_step'] > 0.5})"
done
done
