#!/bin/bash
# 8 GPUs: does binding every rank to the NUMA node of its GPU help the end-to-end leg (pinned host copies)?  A/B
N=${1:-8}
set -x
nvidia-smi topo -m 2>&1 | head -24
for d in /sys/devices/system/node/node*; do echo "$d: $(cat $d/cpulist)"; done
nproc
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29531 bench.py --gpus $N --steps 3 --warmup 3 --inner 500 --no-extra --no-cpu-baseline 2>gpurun_out/r2zk_bind.err | tail -1 > gpurun_out/r2zk_bench_n${N}_numa_bound.json
LB_NO_NUMA_BIND=1 timeout 600 $TR --master-port 29532 bench.py --gpus $N --steps 3 --warmup 3 --inner 500 --no-extra --no-cpu-baseline 2>gpurun_out/r2zk_nobind.err | tail -1 > gpurun_out/r2zk_bench_n${N}_not_bound.json
python - <<PY
import json
for f in ("numa_bound","not_bound"):
    d=json.load(open(f"gpurun_out/r2zk_bench_n${N}_{f}.json"))
    print(f, "value", round(d["value"]), "e2e", round(d["e2e"]["value"]), "one step per call", round(d["e2e"]["one_timestep_per_call"]["value"]), d["e2e"].get("host_numa_binding"))
PY
