#!/bin/bash
# the parity tests of the suite with each kernel variant forced wherever it is compiled and applies (environment
# switches of lb_capi.cu): every variant must reproduce the goldens / the oracle exactly as the default choice does
set -x
T="tests/test_gpu_parity.py tests/test_gpu_baseline_sizes.py tests/test_gpu_round2.py tests/test_gpu_reference_suite.py tests/test_gpu_stock_flows.py tests/test_gpu_run_host.py"
for env in "LB_PF=1" "LB_PF=0" "LB_TMA=1" "LB_TMA=0" "LB_NO_LOOP=1" "LB_NO_RESIDENT=1" "LB_NO_PDL=1" "LB_VEC=1"; do
  echo "== $env"
  env $env timeout 900 python -m pytest $T -m gpu -x -q -p no:cacheprovider 2>&1 | tail -2
done
