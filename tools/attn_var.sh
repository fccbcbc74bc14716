#!/bin/bash
# timing experiments on the attention kernel: MDGEN_ATTN_VAR bits switch parts of the work off (results wrong)
for v in ${VARS:-0 1 2 4 8 16 3 7 31}; do
  echo -n "var=$v: "; MDGEN_ATTN_VAR=$v NOSTAMP=1 timeout 120 python tools/attn_probe.py 2>&1 | head -1
done
