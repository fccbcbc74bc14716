#!/bin/bash
# A/B of the half-block pipelined softmax (MDGEN_ATTN_PIPE=1) against v7
for P in 0 1; do
  echo "== MDGEN_ATTN_PIPE=$P"
  MDGEN_ATTN_PIPE=$P timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k attention 2>&1 | tail -2
  for shape in "32 1000 4" "256 250 1" "250 256 1" "64 100 4" "128 100 4"; do
    MDGEN_ATTN_PIPE=$P NOSTAMP=1 timeout 120 python tools/attn_probe.py $shape 2>&1 | tail -1 | sed "s/^/  shape [$shape]: /"
  done
done
