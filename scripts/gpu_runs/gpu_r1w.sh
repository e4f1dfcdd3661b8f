set -x
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 600 python bench.py > gpurun_out/r1w_bench.json 2> gpurun_out/r1w_bench.err; tail -c 600 gpurun_out/r1w_bench.json
timeout 900 python scripts/bench_configs.py > gpurun_out/r1w_configs.jsonl 2> gpurun_out/r1w_configs.err; tail -3 gpurun_out/r1w_configs.err
CMD="python bench.py --steps 2 --warmup 3 --inner 10 --no-cpu-baseline --no-e2e"
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r1w_launches.csv $CMD > gpurun_out/r1w_ncu_list.log 2>&1
tail -1 gpurun_out/r1w_ncu_list.log | cut -c1-200
