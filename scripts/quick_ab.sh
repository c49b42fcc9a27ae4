#!/bin/bash
# quick check of a kernel change: block-sparse / persistent-kernel parity tests, then CTA-0 phase clocks on the Venice shape
mkdir -p gpurun_out
out=gpurun_out/quick_ab.log; : > $out
timeout 300 python -m pytest tests -m gpu -x -q -k "sparse or persistent or venice or golden" > gpurun_out/quick_pytest.log 2>&1; echo "pytest rc=$? $(tail -1 gpurun_out/quick_pytest.log)" >> $out
for mode in 0 3; do
  echo "== SBA_PS_MODE=$mode" >> $out
  SBA_PS_MODE=$mode SBA_PS_CLOCK=2 timeout 100 python scripts/prof_target.py venice 2>&1 | grep -v "per-CTA" | tail -4 >> $out
done
cat $out
