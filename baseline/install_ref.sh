#!/bin/bash
# Installs the UNMODIFIED reference (PhiSpel/lettuce, /root/reference) into the git-ignored baseline/_ref/ so
# that it travels to the GPU box with `gpurun` and `bench.py --impl reference` can time the reference's own
# CPU PyTorch path there (cpu_baseline.kind = "reference").
#
# The contract's `pip install --no-index --no-build-isolation --find-links /opt/wheelhouse --target
# baseline/_ref /root/reference` (also with --no-deps, from a writable copy) fails in this image: the reference's
# setup.py has `setup_requires=['pytest-runner']`, which is in neither site-packages nor /opt/wheelhouse, and
# there is no network.  lettuce is a pure-Python package (setup.py: packages=find_packages(), no extension
# modules), so what pip would put under --target is exactly the `lettuce/` directory: this script copies it
# verbatim.  Nothing under baseline/_ref is tracked by git; nothing in the product imports it.
set -e
here="$(cd "$(dirname "$0")" && pwd)"
src=/root/reference/lettuce
[ -d "$src" ] || { echo "install_ref: $src not found (GPU box?) -- keeping what is there"; exit 0; }
if [ -f "$here/_ref/lettuce/_simulation.py" ] && ! diff -rq "$src" "$here/_ref/lettuce" -x __pycache__ >/dev/null 2>&1; then
  rm -rf "$here/_ref/lettuce"
fi
if [ ! -d "$here/_ref/lettuce" ]; then
  mkdir -p "$here/_ref"
  cp -r "$src" "$here/_ref/lettuce"
  find "$here/_ref" -name __pycache__ -type d -prune -exec rm -rf {} +
fi
echo "install_ref: baseline/_ref/lettuce is the reference's lettuce/ ($(find "$here/_ref/lettuce" -name '*.py' | wc -l) python files)"
