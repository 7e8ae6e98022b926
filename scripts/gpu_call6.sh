#!/bin/bash
mkdir -p gpurun_out/c6
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_flags_gpu.py tests/test_loader_gpu.py -m gpu -x -q > gpurun_out/c6/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c6/pytest.log
tail -4 gpurun_out/c6/pytest.log
timeout 200 python benchmarks/kernel_bench.py --only perm --out gpurun_out/c6/perm.jsonl > gpurun_out/c6/perm.log 2>&1; cat gpurun_out/c6/perm.jsonl
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 150 $TR --master-port 29751 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e --exchange push --timeout-s 10 --watchdog 120 > gpurun_out/c6/push_633.json 2> gpurun_out/c6/push_633.err; echo "push 6/3/3 rc=$?"
tail -1 gpurun_out/c6/push_633.json | cut -c1-250
grep -E "LoaderTimeout" gpurun_out/c6/push_633.err | tail -2 | cut -c1-1500
timeout 150 $TR --master-port 29752 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e > gpurun_out/c6/pull_n2.json 2> gpurun_out/c6/pull_n2.err; echo "pull rc=$?"
tail -1 gpurun_out/c6/pull_n2.json | cut -c1-250
