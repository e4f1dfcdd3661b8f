# round 1, session c: cooperative multi-step loop for small lattices, two-level reductions, TRT tuning
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/pytest_gpu.log
cat gpurun_out/pytest_gpu.log
echo "=== default"
timeout 900 python scripts/bench_configs.py --only=cfg1 --only=cfg3 --only=cfg4 --only="trt" --only="kbc f64" --only="2048" > gpurun_out/configs_c.jsonl 2> gpurun_out/configs_c.err
cut -c1-220 gpurun_out/configs_c.jsonl; tail -3 gpurun_out/configs_c.err
echo "=== LB_NO_LOOP"
LB_NO_LOOP=1 timeout 300 python scripts/bench_configs.py --only=cfg1 | cut -c1-220
echo "=== LB_VEC=2"
LB_VEC=2 timeout 300 python scripts/bench_configs.py --only="D3Q19 trt" --only="D3Q19 bgk f32 256" --only="cfg3" --only="2048" | cut -c1-220
