#!/bin/bash
# A/B of programmatic dependent launch (MDGEN_NO_PDL=1 launches every kernel as an ordinary full dependency)
timeout 1200 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_parity.py tests/test_graph.py -m gpu -x -q 2>&1 | tail -3
for c in 3 1 4; do
for v in "A=1" MDGEN_NO_PDL=1 "A=1"; do
  env $v timeout 600 python bench.py --config $c --steps 5 --warmup 3 --no-cpu-baseline --no-other-configs 2>&1 | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('config $c', '$v', round(d['value']), round(d['ms_per_step'], 2))"
done
done
