#!/bin/bash
# 2-GPU check: the two-GPU parity tests and the bench line at N=2
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q -k "two_gpu" > gpurun_out/two_gpu_pytest.log 2>&1; echo "pytest rc=$? $(tail -1 gpurun_out/two_gpu_pytest.log)"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 6 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench rc=$?"
cut -c1-400 gpurun_out/bench_n2.json; tail -3 gpurun_out/bench_n2.err
