#!/bin/bash
# (an x-marching enstrophy kernel was timed here too: no gain, not kept) ncu --set full of the chunked prefetch kernel
# (D3Q27 KBC fp32 256^3), of the classic kernel beside it, and of the tile-resident kernel (config 1)
set -x
timeout 900 python -m pytest tests -m gpu -x -q -p no:cacheprovider -k "enstrophy or moments or slab or observ or reporter" 2>&1 | tail -4
timeout 300 python scripts/bench_output.py 256 --no-vtk-runs 2>&1 | grep "^{" | grep -i enstrophy | tee gpurun_out/r2u_enstrophy.jsonl
LB_ENSTROPHY_NO_MARCH=1 timeout 300 python scripts/bench_output.py 256 --no-vtk-runs 2>&1 | grep "^{" | grep -i enstrophy | sed 's/^{/{"LB_ENSTROPHY_NO_MARCH": 1, /' | tee -a gpurun_out/r2u_enstrophy.jsonl
timeout 300 python scripts/bench_output.py 128 --no-vtk-runs 2>&1 | grep "^{" | grep -i enstrophy | tee -a gpurun_out/r2u_enstrophy.jsonl
timeout 300 ncu --set full --clock-control none --import-source on -k regex:step_pf_kernel -s 20 -c 1 -o gpurun_out/r2u_ncu_pf python scripts/run_case.py D3Q27 kbc f32 256 30 > /dev/null 2>&1
ncu -i gpurun_out/r2u_ncu_pf.ncu-rep --page raw --csv > gpurun_out/r2u_ncu_pf_raw.csv 2>/dev/null
LB_PF=0 timeout 300 ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 20 -c 1 -o gpurun_out/r2u_ncu_classic python scripts/run_case.py D3Q27 kbc f32 256 30 > /dev/null 2>&1
ncu -i gpurun_out/r2u_ncu_classic.ncu-rep --page raw --csv > gpurun_out/r2u_ncu_classic_raw.csv 2>/dev/null
timeout 300 ncu --set full --clock-control none --import-source on -k regex:step_resident -c 1 -o gpurun_out/r2u_ncu_resident python scripts/run_case.py D2Q9 bgk f64 256 1000 > /dev/null 2>&1
ncu -i gpurun_out/r2u_ncu_resident.ncu-rep --page raw --csv > gpurun_out/r2u_ncu_resident_raw.csv 2>/dev/null
rm -f gpurun_out/r2u_ncu_*.ncu-rep.tmp; ls -la gpurun_out/ | grep r2u
