#!/bin/bash
# One gpurun call: GPU parity tests, bench line, ncu launch lists (profiling target and bench.py) + full capture of the dominant kernel.
# Usage (from the repo root on the GPU box): bash scripts/round_gpu.sh <tag>
tag=${1:-r1i}
mkdir -p gpurun_out
timeout 420 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/${tag}_pytest_gpu.log
timeout 300 python bench.py > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"
timeout 120 python scripts/prof_target.py venice > gpurun_out/${tag}_prof_plain.log 2>&1 && {
  timeout 240 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv \
      python scripts/prof_target.py venice > gpurun_out/${tag}_ncu_list.log 2>&1
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:'k_pcg_solve|k_explicit_compact' -c 3 -f -o gpurun_out/${tag}_full \
      python scripts/prof_target.py venice > gpurun_out/${tag}_ncu_full.log 2>&1
}
timeout 60 env SBA_PS_CLOCK=1 python scripts/prof_target.py venice > gpurun_out/${tag}_ps_clock.log 2>&1
# launch list of bench.py itself (the judged command), short form; its printed numbers are not bench values
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_bench_launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/${tag}_bench_ncu.log 2>&1
tail -3 gpurun_out/${tag}_pytest_gpu.log; cat gpurun_out/${tag}_bench.json | cut -c1-600
