#!/bin/bash
# A/B of the one-kernel MLP block (MDGEN_NO_MLP_FUSED=1: fc1 + GELU on the pair GEMM, fc2 in gemm_ln)
if [ "$1" != "quick" ]; then
  timeout 1500 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_parity.py tests/test_graph.py -m gpu -x -q 2>&1 | tail -3
fi
for c in ${CONFIGS:-3 4 5}; do
for v in "A=1" MDGEN_NO_MLP_FUSED=1 "A=1"; do
  env $v timeout 600 python bench.py --config $c --steps 5 --warmup 3 --no-cpu-baseline --no-other-configs 2>&1 | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('config $c', '$v', round(d['value']), round(d['ms_per_step'], 2), {k: v['ms_per_step'] for k, v in d['kernels'].items() if 'gemm_tc' in k}, d.get('parity_check', {}).get('rel_l2_velocity_bf16_vs_fp32_check'))"
done
done
