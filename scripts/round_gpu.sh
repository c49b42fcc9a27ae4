#!/bin/bash
# One gpurun call: GPU parity tests, bench line, ncu launch lists (profiling target and bench.py) + full captures of the main kernels.
# Usage (from the repo root on the GPU box): bash scripts/round_gpu.sh <tag>
tag=${1:-r2h}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q > gpurun_out/${tag}_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/${tag}_pytest_gpu.log
timeout 400 python bench.py --steps 20 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; echo "bench rc=$?"
timeout 400 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/${tag}_reference_arm.json 2> gpurun_out/${tag}_reference_arm.err; echo "reference arm rc=$?"
timeout 120 python scripts/prof_target.py venice > gpurun_out/${tag}_prof_plain.log 2>&1 && {
  timeout 240 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv \
      python scripts/prof_target.py venice > gpurun_out/${tag}_ncu_list.log 2>&1
  timeout 400 ncu --set full --clock-control none --import-source on \
      -k regex:'k_pcg_solve|k_residual_jacobian|k_linearize_q|k_camera_blocks|k_schur_diag$|k_explicit_contrib|k_explicit_assemble|k_explicit_compact|k_camera_pass_f|k_model_tiles' -c 16 -f -o gpurun_out/${tag}_full \
      python scripts/prof_target.py venice > gpurun_out/${tag}_ncu_full.log 2>&1
}
timeout 120 python scripts/prof_dlt.py > gpurun_out/${tag}_dlt_plain.log 2>&1 && \
  timeout 240 ncu --set full --clock-control none --import-source on -k regex:k_triangulate -c 1 -f -o gpurun_out/${tag}_dlt_full \
      python scripts/prof_dlt.py > gpurun_out/${tag}_ncu_dlt.log 2>&1
timeout 60 env SBA_PS_CLOCK=1 python scripts/prof_target.py venice > gpurun_out/${tag}_ps_clock.log 2>&1
# launch list of bench.py itself (the judged command), short form; its printed numbers are not bench values
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_bench_launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --eta 0 > gpurun_out/${tag}_bench_ncu.log 2>&1
python -c "import sys; sys.path.insert(0, '.'); import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; echo "smoke rc=$?"
tail -3 gpurun_out/${tag}_pytest_gpu.log; cat gpurun_out/${tag}_bench.json | cut -c1-400; tail -2 gpurun_out/${tag}_smoke.log
