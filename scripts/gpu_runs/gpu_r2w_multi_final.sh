#!/bin/bash
# final library of round 2 on N GPUs: slab correctness (all transports, stock obstacle in slabs, enstrophy across
# slabs), the multi-GPU tests of the suite, bench.py --gpus N
N=${1:-2}
set -x
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29511 scripts/slab_check.py --res 48 --steps 21 > gpurun_out/r2w_slab_check_n${N}_full.log 2>&1
grep "slab_check" gpurun_out/r2w_slab_check_n${N}_full.log | tee gpurun_out/r2w_slab_check_n${N}.log | grep -v "bit-exact=True" | tail -6
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -x -q -p no:cacheprovider 2>&1 | tail -3
timeout 900 $TR --master-port 29512 bench.py --gpus $N 2>gpurun_out/r2w_bench_n${N}.err | tail -1 > gpurun_out/r2w_bench_n${N}.json
python - <<PY
import json
d = json.load(open("gpurun_out/r2w_bench_n${N}.json"))
print("N=${N} value", d["value"], "us/step", d["ms_per_step"], "slab_bit_exact", d.get("slab_bit_exact"), "e2e", d["e2e"]["value"], d["config"]["parallelism"])
for e in d.get("extra", []):
    print(e)
PY
