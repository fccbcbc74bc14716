#!/bin/bash
# one `ncu --set full` source-level capture of one GEMM kernel instance inside a config-3 evaluation
mkdir -p gpurun_out
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"${KREGEX:-gemm_tc_kernel<128, 1}" -s ${KSKIP:-12} -c 1 -o gpurun_out/gemm -f \
  python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/gemm_ncu.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/gemm.ncu-rep --page raw --csv > gpurun_out/gemm_raw.csv 2>/dev/null
ncu -i gpurun_out/gemm.ncu-rep --page source --csv > gpurun_out/gemm_source.csv 2>/dev/null
rm -f gpurun_out/gemm.ncu-rep; ls -la gpurun_out | grep gemm
