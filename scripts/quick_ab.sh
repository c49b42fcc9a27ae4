#!/bin/bash
# quick check of a kernel change: block-sparse / persistent-kernel parity tests, then CTA-0 phase clocks on the Venice shape
mkdir -p gpurun_out
out=gpurun_out/quick_ab.log; : > $out
timeout 300 python -m pytest tests -m gpu -x -q -k "sparse or persistent or venice or golden" > gpurun_out/quick_pytest.log 2>&1; echo "pytest rc=$? $(tail -1 gpurun_out/quick_pytest.log)" >> $out
for sp in 1 0; do
  echo "== SBA_PS_SPARSE=$sp (default SBA_PS_MODE)" >> $out
  SBA_PS_SPARSE=$sp SBA_PS_CLOCK=1 timeout 100 python scripts/prof_target.py venice 2>&1 | tail -2 >> $out
done
cat $out
