#!/bin/bash
# A/B of the cluster-of-4 projection + residual + LayerNorm kernel on the small configs (MDGEN_NO_GEMM_LN4=1: GEMM + stand-alone LayerNorm)
timeout 1500 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_parity.py tests/test_graph.py -m gpu -x -q 2>&1 | tail -3
for c in ${CONFIGS:-1 2}; do
for v in "A=1" MDGEN_NO_GEMM_LN4=1 "A=1" MDGEN_NO_GEMM_LN4=1; do
  env $v timeout 600 python bench.py --config $c --steps 5 --warmup 3 --no-cpu-baseline --no-other-configs 2>&1 | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('config $c', '$v', round(d['value']), round(d['ms_per_step'], 2), d['parity_check']['rel_l2_velocity_bf16_vs_fp32_check'])"
done
done
