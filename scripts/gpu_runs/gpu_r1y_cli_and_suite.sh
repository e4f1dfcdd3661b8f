#!/bin/bash
# console driver on the GPU (benchmark on five stock flows, the reference's convergence check), then as much
# of the whole GPU suite as fits in the remaining budget
timeout 60 python -m pytest tests/test_cli.py -m gpu -q -p no:cacheprovider 2>&1 | tail -12
timeout 30 python -m lettuce_b200 -p double convergence 2>&1 | tail -7
timeout 58 python -m pytest tests -m gpu -x -q -p no:cacheprovider --deselect tests/test_cli.py 2>&1 | tail -12
