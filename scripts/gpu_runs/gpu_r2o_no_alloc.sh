#!/bin/bash
# lb_observables_static: its test alone in a fresh process, then behind the whole GPU suite (pytest -x order)
set -x
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider 2>&1 | tail -5
timeout 900 python -m pytest tests -m gpu -x -q -p no:cacheprovider 2>&1 | tail -5
