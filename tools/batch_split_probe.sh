#!/bin/bash
# does config 3 run faster per frame at a smaller batch (activations closer to the 126 MB L2)?
for b in 8 4 2; do
  echo -n "B=$b: "; timeout 200 python -c "
import sys, json, io, contextlib
import mdgen_finetune_b200.configs as c
c.CONFIGS[3]['B'] = $b
import bench
sys.argv = ['bench.py', '--steps', '3', '--warmup', '3', '--no-cpu-baseline']
buf = io.StringIO()
with contextlib.redirect_stdout(buf):
    bench.main()
d = json.loads([l for l in buf.getvalue().splitlines() if l.startswith('{')][-1])
print(round(d['value']), round(d['ms_per_step'], 2), {k: v['ms_per_step'] for k, v in d['kernels'].items() if v['ms_per_step'] > 0.4})
" 2>&1 | tail -1
done
