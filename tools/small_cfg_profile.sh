#!/bin/bash
# per-kernel-family breakdown of the small configs (profile mode = eager launches bracketed by events) next to the graph time
for c in 1 2 4 5; do
  python bench.py --config $c --steps 5 --warmup 3 --no-cpu-baseline --no-other-configs 2>&1 | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read())
k = {n: (v['ms_per_step'], v['launches']) for n, v in d['kernels'].items()}
print('config $c:', round(d['value']), 'frames/s', round(d['ms_per_step'], 2), 'ms; eager kernel sum', round(sum(v[0] for v in k.values()), 2), 'ms', k)"
done
