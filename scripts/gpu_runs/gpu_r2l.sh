#!/bin/bash
set -x
scripts/dev/svp.bin | tee gpurun_out/r2l_scalar_vs_packed.log
