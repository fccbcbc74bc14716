for c in 1 2; do
for e in "X=1" "MDGEN_NO_GEMM_LN=1" "MDGEN_NO_GEMM_2CTA=1" "MDGEN_NO_GEMM_LN=1 MDGEN_NO_GEMM_2CTA=1"; do
  echo -n "cfg $c [$e] "; env $e timeout 200 python bench.py --config $c --steps 3 --warmup 3 --no-cpu-baseline 2>&1 | grep '^{' | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(round(d['value']), round(d['ms_per_step'],2), {k:v['ms_per_step'] for k,v in d['kernels'].items() if v['ms_per_step']>0.4})"
done; done
