#!/bin/bash
# A/B of the pair kernel's 16-epilogue-warp variant (default for GELU; MDGEN_NO_PAIR16=1 turns it off; MDGEN_PAIR16_STORE=1
# also uses it for the plain bf16 stores)
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -2
for v in "A=1" MDGEN_NO_PAIR16=1 MDGEN_PAIR16_STORE=1 "A=1"; do
  env $v python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-other-configs 2>&1 | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print('$v', round(d['value']), round(d['ms_per_step'], 2), {k: v['ms_per_step'] for k, v in d['kernels'].items() if 'gemm' in k})"
done
