# round 1, session d: observables reduced before the stores, programmatic dependent launch
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/pytest_gpu.log
cat gpurun_out/pytest_gpu.log
echo "=== default (PDL)"
timeout 900 python scripts/bench_configs.py > gpurun_out/configs_d.jsonl 2> gpurun_out/configs_d.err
cut -c1-220 gpurun_out/configs_d.jsonl; tail -3 gpurun_out/configs_d.err
echo "=== LB_NO_PDL"
LB_NO_PDL=1 timeout 300 python scripts/bench_configs.py --only=cfg2 --only=cfg3 --only="2048" | cut -c1-220
