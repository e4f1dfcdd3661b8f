#!/bin/bash
# round 2, multi-GPU call (N = number of GPUs of the box): slab correctness (all transports, stock obstacle in slabs,
# enstrophy across slabs), bench.py --gpus N with the default transport (+ extras: config 5), and round 1's 'p2p' for A/B
N=${1:-2}
set -x
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29511 scripts/slab_check.py --res 48 --steps 21 > gpurun_out/r2k_slab_check_n${N}_full.log 2>&1
grep "slab_check" gpurun_out/r2k_slab_check_n${N}_full.log | tee gpurun_out/r2k_slab_check_n${N}.log | tail -4
timeout 900 $TR --master-port 29512 bench.py --gpus $N 2>gpurun_out/r2k_bench_n${N}.err | tail -1 > gpurun_out/r2k_bench_n${N}.json
python - <<PY
import json
d = json.load(open("gpurun_out/r2k_bench_n${N}.json"))
print("N=${N} value", d["value"], "us/step", d["ms_per_step"], "slab_bit_exact", d.get("slab_bit_exact"), "e2e", d["e2e"]["value"], d["config"]["parallelism"])
for e in d.get("extra", []):
    print(e)
PY
LB_SLAB_TRANSPORT=p2p timeout 300 $TR --master-port 29513 bench.py --gpus $N --no-extra --no-e2e 2>/dev/null | tail -1 > gpurun_out/r2k_bench_n${N}_p2p.json
python -c "
import json; d=json.load(open('gpurun_out/r2k_bench_n${N}_p2p.json')); print('p2p (round-1 transport) N=${N}', d['value'], d['ms_per_step'])"
