#!/bin/bash
# final library: whole GPU suite, smoke, then the ncu launch list (gpu__time_duration.sum, --clock-control none) of a
# short bench command -- first run without ncu, then the same command under ncu (its numbers are not bench values)
set -x
timeout 1200 python -m pytest tests -m gpu -x -q -p no:cacheprovider 2>&1 | tail -3
timeout 60 python __graft_entry__.py smoke 2>&1 | tail -1
timeout 600 python bench.py --steps 2 --warmup 3 --inner 100 --no-extra --no-cpu-baseline > gpurun_out/r2zc_bench_short.json 2> gpurun_out/r2zc_bench_short.err && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r2zc_launches.csv \
  python bench.py --steps 2 --warmup 3 --inner 100 --no-extra --no-cpu-baseline > gpurun_out/r2zc_ncu.log 2>&1
tail -c 300 gpurun_out/r2zc_bench_short.json; wc -l gpurun_out/r2zc_launches.csv
