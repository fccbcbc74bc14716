#!/bin/bash
# fused projection + residual + LayerNorm kernel on the small configs: threshold on the token count
for c in 1 2; do
for v in "A=1" MDGEN_GEMM_LN_MIN_M=2048 "A=1" MDGEN_GEMM_LN_MIN_M=2048; do
  env $v timeout 600 python bench.py --config $c --steps 5 --warmup 3 --no-cpu-baseline --no-other-configs 2>&1 | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('config $c', '$v', round(d['value']), round(d['ms_per_step'], 2), d.get('parity_check'))"
done
done
