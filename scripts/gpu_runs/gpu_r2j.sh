#!/bin/bash
# round 2, call j (1 GPU): TMA kernel with masks (config 3), write-coverage tests, full suite, the full bench line
set -x
timeout 300 python -m pytest tests/test_gpu_round2.py -m gpu -x -q -p no:cacheprovider 2>&1 | tail -8
timeout 900 python -m pytest tests -m gpu -x -q -p no:cacheprovider 2>&1 | tail -4
timeout 900 python scripts/bench_tma.py --quick 2>&1 | grep "^{" | tee gpurun_out/r2j_bench_tma.jsonl
timeout 300 python scripts/bench_small.py 2>&1 | grep "^{" | tee gpurun_out/r2j_bench_small.jsonl
timeout 600 python bench.py 2>/dev/null | tail -1 > gpurun_out/r2j_bench_n1.json
python - <<'PY'
import json
d = json.load(open("gpurun_out/r2j_bench_n1.json"))
print("value", d["value"], "frac", d["roofline"]["frac"], "parity", d["parity"], "e2e", d["e2e"]["value"])
for e in d["extra"]:
    print(e)
PY
