#!/bin/bash
# Run on the GPU box (via gpurun): plain bench first, then the ncu launch list of one velocity
# evaluation and one `--set full` capture of the two top kernels.  Outputs under gpurun_out/.
set -u
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
SKIP=${SKIP:-9000}     # launches of conditioning + the first warm-up rollout
$CMD > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -20 gpurun_out/plain.log; exit 1; }
tail -c 600 gpurun_out/plain.log
ncu --metrics gpu__time_duration.sum --clock-control none -s $SKIP -c ${COUNT:-200} --csv \
    --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:"${KREGEX:-attention_tc_kernel|gemm_tc_kernel}" -s ${KSKIP:-60} -c ${KCOUNT:-8} \
    -o gpurun_out/prof -f $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la gpurun_out
