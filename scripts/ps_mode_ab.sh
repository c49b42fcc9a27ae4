#!/bin/bash
# A/B of the persistent PCG kernel's ring prefetch (SBA_PS_MODE bit 0) and interleaved L2 share (bit 1), panels and dense tiles.
mkdir -p gpurun_out
out=gpurun_out/ps_mode_ab.log; : > $out
timeout 300 python -m pytest tests -m gpu -x -q > gpurun_out/ps_mode_pytest.log 2>&1; echo "pytest rc=$? $(tail -1 gpurun_out/ps_mode_pytest.log)" >> $out
for sp in 1 0; do for mode in 0 1 2 3; do
  echo "== SBA_PS_SPARSE=$sp SBA_PS_MODE=$mode" >> $out
  SBA_PS_SPARSE=$sp SBA_PS_MODE=$mode SBA_PS_CLOCK=1 timeout 100 python scripts/prof_target.py venice 2>&1 | tail -3 >> $out
done; done
cat $out
