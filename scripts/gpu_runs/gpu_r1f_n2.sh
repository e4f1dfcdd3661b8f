mkdir -p gpurun_out
nvidia-smi -L
timeout 900 python -m pytest tests/test_multi_gpu.py tests/test_gpu_slab_emulated.py -m gpu -x -q 2>&1 | tail -15
N=2 bash scripts/gpu_n2_experiments.sh 2>&1 | tail -20
python bench.py --steps 3 --warmup 3 --inner 300 --no-e2e --no-cpu-baseline | python -c 'import sys,json; d=json.loads(sys.stdin.read()); print("N=1", d["value"], d["roofline"]["avg_launch_us"])'
