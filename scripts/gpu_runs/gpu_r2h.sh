#!/bin/bash
# round 2, call h (1 GPU): chain kernel (config 1), TMA kernel with lane-parallel issue, the reference's own GPU paths
# (torch-CUDA and generated cuda_native kernel) beside lettuce_b200, compute-sanitizer
set -x
timeout 300 python -m pytest tests/test_gpu_round2.py -m gpu -x -q -p no:cacheprovider 2>&1 | tail -8
timeout 900 python -m pytest tests -m gpu -x -q -p no:cacheprovider 2>&1 | tail -4
timeout 600 python bench.py --no-cpu-baseline --no-e2e 2>/dev/null | tail -1 > gpurun_out/r2h_bench_n1_extras.json
python - <<'PY'
import json
d = json.load(open("gpurun_out/r2h_bench_n1_extras.json"))
print("value", d["value"], "frac", d["roofline"]["frac"])
for e in d["extra"]:
    print(e)
PY
timeout 900 python scripts/bench_tma.py --quick 2>&1 | grep "^{" | tee gpurun_out/r2h_bench_tma.jsonl
timeout 900 python scripts/bench_reference_gpu.py 2>&1 | grep "^{" | tee gpurun_out/r2h_reference_gpu.jsonl
bash scripts/gpu_runs/gpu_r2g_sanitize.sh
