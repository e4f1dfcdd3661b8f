cd lettuce_b200/csrc
for variant in "-DLB_KBC_SUMMED_FEQ" ""; do
  echo "=== EXTRA='$variant'"
  touch lb_physics.cuh
  make -j16 EXTRA="$variant" > /tmp/mk.log 2>&1 || tail -5 /tmp/mk.log
  (cd ../.. && python -m pytest tests/test_gpu_parity.py -m gpu -q -s -k "test_collisions_vs_oracle and kbc and f32" 2>&1 | grep -E "rel_err|passed|failed" | sort | uniq | head -40)
done
