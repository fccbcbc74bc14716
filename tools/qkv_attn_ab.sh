#!/bin/bash
# A/B of the fused q/k/v projection + residue-axis attention kernel (gemm_qkv_attn.cuh) on config 3
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "full_size_velocity or variants_agree" 2>&1 | tail -3
for v in "" MDGEN_NO_QKV_ATTN=1; do
  env $v python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-other-configs 2>&1 | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('$v', round(d['value']), round(d['ms_per_step'], 2), {k: v['ms_per_step'] for k, v in d['kernels'].items()})"
done
