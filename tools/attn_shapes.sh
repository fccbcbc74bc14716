#!/bin/bash
# correctness + stand-alone timing of the tcgen05 attention kernel on the config-3 / ATLAS (both axes) / config-2,4 shapes
timeout 300 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k attention 2>&1 | tail -2
for shape in "32 1000 4" "256 250 1" "250 256 1" "64 100 4" "128 100 4"; do
  NOSTAMP=1 timeout 120 python tools/attn_probe.py $shape 2>&1 | tail -1 | sed "s/^/  shape [$shape]: /"
done
