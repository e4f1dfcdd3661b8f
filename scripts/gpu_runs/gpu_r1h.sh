timeout 900 python -m pytest tests/test_gpu_initialize.py -m gpu -q -s 2>&1 | grep -E "initialise|passed|failed|Error" | head -40
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
