#!/bin/bash
# Round-2 profile evidence (run on the GPU box via gpurun): plain bench line first, then
#   1. the ncu launch list (gpu__time_duration.sum, --clock-control none) of ~10 velocity evaluations of the graph rollout,
#   2. one `ncu --set full` capture of one layer's worth of the tensor-core kernels, raw page -> traffic JSON + summary.
set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-other-configs"
$CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/plain.log; exit 1; }
tail -c 400 gpurun_out/plain.log; echo
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
python tools/launch_summary.py gpurun_out/launches.csv 380 > gpurun_out/launches_summary.txt; cat gpurun_out/launches_summary.txt
ncu --set full --clock-control none --import-source on \
    -k regex:"attention_tc_kernel|gemm_tc_kernel|gemm_tc2_kernel|gemm_ln_kernel|gemm_mlp_ln_kernel|gemm_qkv_attn4_kernel|embed_step_kernel|ln_modulate_kernel" \
    -s 500 -c 34 -o gpurun_out/prof -f $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
ncu -i gpurun_out/prof.ncu-rep --page raw --csv > gpurun_out/prof_raw.csv 2>/dev/null
python tools/ncu_traffic.py gpurun_out/prof_raw.csv gpurun_out/traffic_cfg3.json > gpurun_out/ncu_full_summary.txt; cat gpurun_out/ncu_full_summary.txt
rm -f gpurun_out/prof.ncu-rep
ls -la gpurun_out | head -30
