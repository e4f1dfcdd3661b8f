#!/bin/bash
# ncu --set full of the chunked prefetch kernel at config 4's size (D3Q27 KBC fp32 512^3), one launch
set -x
timeout 600 ncu --set full --clock-control none --import-source on -k regex:step_pf_kernel -s 12 -c 1 -o gpurun_out/r2zd_ncu_pf_512 python scripts/run_case.py D3Q27 kbc f32 512 20 > /dev/null 2>&1
ncu -i gpurun_out/r2zd_ncu_pf_512.ncu-rep --page raw --csv > gpurun_out/r2zd_ncu_pf_512_raw.csv 2>/dev/null
rm -f gpurun_out/r2zd_ncu_pf_512.ncu-rep; ls -la gpurun_out | grep r2zd
