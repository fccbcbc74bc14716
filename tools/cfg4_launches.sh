#!/bin/bash
# ncu launch list of the design config (4): which kernels make up its 107 launches per evaluation
CMD="python bench.py --config 4 --steps 1 --warmup 3 --no-cpu-baseline --no-other-configs"
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches_cfg4.csv $CMD > gpurun_out/ncu_cfg4.log 2>&1
python tools/launch_summary.py gpurun_out/launches_cfg4.csv 1070
