#!/bin/bash
# round 2, call m (1 GPU): after the packed-rounding fix -- vector width independence, whole suite, smoke, bench line
set -x
timeout 600 python scripts/check_vec_width.py 2>&1 | grep "vector\|check_vec" | tee gpurun_out/r2m_vec_width.log
timeout 900 python -m pytest tests -m gpu -x -q -p no:cacheprovider 2>&1 | tail -5
timeout 60 python __graft_entry__.py smoke 2>&1 | tail -1
timeout 600 python bench.py 2>/dev/null | tail -1 > gpurun_out/r2m_bench_n1.json
python - <<'PY'
import json
d = json.load(open("gpurun_out/r2m_bench_n1.json"))
print("value", d["value"], "frac", d["roofline"]["frac"], "parity", d["parity"], "e2e", d["e2e"]["value"])
for e in d["extra"]:
    print(e["case"][:70], e.get("mlups"), e.get("us_per_step"), e.get("frac"), e.get("error"))
PY
timeout 300 python scripts/bench_tma.py --quick 2>&1 | grep "^{" | head -4 | tee gpurun_out/r2m_bench_tma_head.jsonl
