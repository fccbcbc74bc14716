#!/bin/bash
# one `ncu --set full` capture of the attention kernel alone (tools/attn_probe.py shape = config 3 frame axis)
mkdir -p gpurun_out
NOSTAMP=1 ncu --set full --clock-control none --import-source on -k regex:attention_tc_kernel -s 4 -c 1 -o gpurun_out/attn -f \
  python tools/attn_probe.py > gpurun_out/attn_ncu.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/attn.ncu-rep --page raw --csv > gpurun_out/attn_raw.csv 2>/dev/null
ncu -i gpurun_out/attn.ncu-rep --page source --csv > gpurun_out/attn_source.csv 2>/dev/null
ncu -i gpurun_out/attn.ncu-rep --page details > gpurun_out/attn_details.txt 2>/dev/null
rm -f gpurun_out/attn.ncu-rep; ls -la gpurun_out | head -30
