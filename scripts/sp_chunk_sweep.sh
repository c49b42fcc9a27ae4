#!/bin/bash
# Chunk size of the block-sparse panel stream (SBA_SP_CHUNK_MAX) x ring prefetch mode: time per CG iteration (CTA 0 clock).
mkdir -p gpurun_out
out=gpurun_out/sp_chunk_sweep.log; : > $out
for ch in 40960 49152 57344 65536; do for mode in 0 3; do
  echo "== SBA_SP_CHUNK_MAX=$ch SBA_PS_MODE=$mode" >> $out
  SBA_SP_CHUNK_MAX=$ch SBA_PS_MODE=$mode SBA_PS_CLOCK=1 timeout 100 python scripts/prof_target.py venice 2>&1 | tail -2 >> $out
done; done
cat $out
