#!/bin/bash
# moments_kernel<OBS> (lb_observables) with per-block partials + finish kernel instead of the per-block ticket
set -x
timeout 1200 python -m pytest tests -m gpu -x -q -p no:cacheprovider 2>&1 | tail -4
timeout 300 python scripts/bench_output.py 256 --no-vtk-runs 2>&1 | grep "^{" | tee gpurun_out/r2y_small_kernels.jsonl | cut -c1-250
