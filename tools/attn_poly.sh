#!/bin/bash
# attention kernel alone (config-3 frame axis shape) for each exp2 polynomial share; then kernel tests with the default
for v in ${POLYS:-0 4 3 2}; do
  echo -n "poly=$v: "; MDGEN_ATTN_POLY=$v NOSTAMP=1 timeout 60 python tools/attn_probe.py 2>&1 | head -1
done
timeout 200 python -m pytest tests/test_gpu_kernels.py -q -m gpu --timeout 60 -x -k attention 2>&1 | tail -3
