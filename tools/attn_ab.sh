#!/bin/bash
# A/B of the tcgen05 attention variants: correctness (kernel tests) then timing on the config-3 / config-5 / config-2 shapes
for SB in 48 32; do
  echo "== MDGEN_ATTN_SB=$SB"
  MDGEN_ATTN_SB=$SB timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k attention 2>&1 | tail -2
  for shape in "32 1000 4" "256 250 1" "250 256 1" "64 100 4"; do
    MDGEN_ATTN_SB=$SB NOSTAMP=1 timeout 120 python tools/attn_probe.py $shape 2>&1 | tail -1 | sed "s/^/  shape [$shape]: /"
  done
done
