#!/bin/bash
# 2-GPU check: all GPU tests (incl. new push kernel and flag barrier default), pull vs push bench
mkdir -p gpurun_out/c2
nvidia-smi -L > gpurun_out/c2/env.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/c2/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c2/pytest.log
tail -5 gpurun_out/c2/pytest.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29711 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e > gpurun_out/c2/bench_pull_n2.json 2> gpurun_out/c2/bench_pull_n2.err; echo "pull rc=$?"
tail -1 gpurun_out/c2/bench_pull_n2.json | cut -c1-300
timeout 300 $TR --master-port 29712 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e --exchange push > gpurun_out/c2/bench_push_n2.json 2> gpurun_out/c2/bench_push_n2.err; echo "push rc=$?"
tail -1 gpurun_out/c2/bench_push_n2.json | cut -c1-300
tail -5 gpurun_out/c2/bench_push_n2.err
