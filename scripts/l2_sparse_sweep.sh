#!/bin/bash
# L2-resident share of the panel stream (SBA_PS_L2_MB, read once per process) vs time per CG iteration; per-CTA mat-vec times.
mkdir -p gpurun_out
for mb in 64 80 96 104 112; do
  echo "== SBA_PS_L2_MB=$mb"; SBA_PS_L2_MB=$mb SBA_PS_CLOCK=1 timeout 100 python scripts/prof_target.py venice 2>&1 | tail -3
done > gpurun_out/l2_sparse_sweep.log 2>&1
SBA_PS_CLOCK=2 timeout 100 python scripts/prof_target.py venice > gpurun_out/ps_clock2.log 2>&1
cat gpurun_out/l2_sparse_sweep.log
