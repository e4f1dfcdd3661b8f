# round 1, session b: packed-fp32 / opposite-pair physics.  Parity first, then timings, then ncu.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > gpurun_out/gpu.txt
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/pytest_gpu.log
cat gpurun_out/pytest_gpu.log
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -s -k "test_collisions_vs_oracle" 2>&1 | grep -E "rel_err|passed|failed" | sort | uniq > gpurun_out/rel_errs.log
tail -3 gpurun_out/rel_errs.log
echo "=== default config sweep"
timeout 900 python scripts/bench_configs.py > gpurun_out/configs_default.jsonl 2> gpurun_out/configs_default.err
cat gpurun_out/configs_default.jsonl | cut -c1-200
for v in 1 2 4; do
  echo "=== LB_VEC=$v D3Q27"
  LB_VEC=$v timeout 300 python scripts/bench_configs.py --only="D3Q27" --only="cfg3" --only="cfg1" 2>/dev/null | grep -v "512" > gpurun_out/configs_vec$v.jsonl
  cat gpurun_out/configs_vec$v.jsonl | cut -c1-200
done
echo "=== bench.py"
timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err
cat gpurun_out/bench_default.json
# ncu: KBC step kernel, full set
CMD="python scripts/run_case.py D3Q27 kbc f32 256 12"
$CMD > gpurun_out/plain_kbc.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 6 -c 2 -o gpurun_out/prof_r1b_d3q27_kbc_f32 $CMD > gpurun_out/ncu_kbc.log 2>&1
tail -2 gpurun_out/ncu_kbc.log
CMD="python scripts/run_case.py D3Q19 bgk f32 256 12"
$CMD > gpurun_out/plain_bgk.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 6 -c 2 -o gpurun_out/prof_r1b_d3q19_bgk_f32 $CMD > gpurun_out/ncu_bgk.log 2>&1
tail -2 gpurun_out/ncu_bgk.log
