#!/bin/bash
# round 2, call g (1 GPU): compute-sanitizer memcheck / racecheck / synccheck over scripts/sanitize_cases.py
set -x
timeout 120 python scripts/sanitize_cases.py 2>&1 | tail -15
for tool in memcheck racecheck synccheck; do
  timeout 1500 compute-sanitizer --tool $tool --error-exitcode 9 python scripts/sanitize_cases.py > gpurun_out/r2g_sanitize_$tool.log 2>&1
  echo "$tool rc=$?"
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|sanitize_cases OK|MISMATCH" gpurun_out/r2g_sanitize_$tool.log | tail -5
done
