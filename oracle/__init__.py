"""oracle/ -- TEST INFRASTRUCTURE, not product code.

CPU restatements of the reference's stream-collide algorithm used only as the checker:
  * lbm_oracle.py      torch-CPU restatement (pinned against tests/golden/*.npz)
  * lbm_oracle.c       plain-C/OpenMP fused restatement (second, independent checker and
                       the multi-threaded `cpu_baseline` of bench.py), built by oracle/Makefile
  * reference_shim.py  imports the unmodified reference from /root/reference (build
                       container only) to generate the golden vectors

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import anything from here.  The product package (lettuce_b200/) never does and raises when
its CUDA library is missing.
"""
