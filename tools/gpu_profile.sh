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
# the raw page as CSV travels back (gpurun_out is capped at 64 MiB); the report itself only when small or KEEP_REP=1
ncu -i gpurun_out/prof.ncu-rep --page raw --csv > gpurun_out/prof_raw.csv 2>/dev/null
if [ "${KEEP_REP:-0}" != "1" ]; then rm -f gpurun_out/prof.ncu-rep; fi
ls -la gpurun_out
