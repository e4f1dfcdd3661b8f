# round 1 final single-GPU evidence: tests, bench (both arms), config sweep, ncu launch list + full captures
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -4 | tee gpurun_out/r1_pytest_gpu.log
timeout 900 python bench.py > gpurun_out/r1_bench_n1.json 2> gpurun_out/r1_bench_n1.err; cat gpurun_out/r1_bench_n1.json | cut -c1-1500
timeout 600 python bench.py --impl reference > gpurun_out/r1_bench_reference.json 2>/dev/null; cut -c1-400 gpurun_out/r1_bench_reference.json
timeout 900 python scripts/bench_configs.py > gpurun_out/r1_bench_configs.jsonl 2>/dev/null; cut -c1-200 gpurun_out/r1_bench_configs.jsonl
CMD="python bench.py --steps 2 --warmup 3 --inner 10 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain_ll.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r1_launches_bench.csv $CMD > gpurun_out/ncu_list.log 2>&1
tail -1 gpurun_out/ncu_list.log | cut -c1-200
for c in "D3Q19 bgk f32" "D3Q27 kbc f32"; do
  set -- $c; tag=$(echo $1_$2_$3 | tr A-Z a-z)
  CMD="python scripts/run_case.py $1 $2 $3 256 12"
  $CMD > gpurun_out/plain_$tag.log 2>&1 && \
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 6 -c 2 -o gpurun_out/r1_ncu_$tag $CMD > gpurun_out/ncu_$tag.log 2>&1
  tail -1 gpurun_out/ncu_$tag.log
done
