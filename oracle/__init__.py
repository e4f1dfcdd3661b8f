"""oracle/ -- TEST INFRASTRUCTURE, not product code.

CPU restatements of the reference's stream-collide algorithm used only as the checker:
  * lbm_oracle.py      torch-CPU restatement (pinned against tests/golden/*.npz)
  * reference_shim.py  imports the UNMODIFIED reference -- from baseline/_ref (git-ignored install,
                       travels to the GPU box: bench.py's reference arm) or /root/reference (build
                       container: golden-vector generation)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import anything from here.  The product package (lettuce_b200/) never does and raises when
its CUDA library is missing.
"""
