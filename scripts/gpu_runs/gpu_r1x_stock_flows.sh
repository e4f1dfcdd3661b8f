#!/bin/bash
# stock flows + output side on the GPU (tests/test_gpu_stock_flows.py, flow_* golden files through test_gpu_parity),
# cost of one VTK report, smoke()
timeout 100 python -m pytest tests/test_gpu_stock_flows.py tests/test_gpu_parity.py -m gpu -q -k "stock_flows or flow_" -p no:cacheprovider 2>&1 | tail -15
timeout 60 python scripts/bench_output.py 256 3 2>&1 | tail -3
timeout 60 python __graft_entry__.py smoke 2>&1 | tail -2
