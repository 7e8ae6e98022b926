#!/bin/bash
# 2-GPU: full GPU tests with new waits/perm_gen(512)/canary; push bench default config; pull bench; perm_gen timings
mkdir -p gpurun_out/c4
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/c4/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c4/pytest.log
tail -4 gpurun_out/c4/pytest.log
timeout 200 python benchmarks/kernel_bench.py --only perm --out gpurun_out/c4/perm.jsonl > gpurun_out/c4/perm.log 2>&1; cat gpurun_out/c4/perm.jsonl
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 150 $TR --master-port 29731 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e --exchange push --timeout-s 15 --watchdog 100 > gpurun_out/c4/push_633.json 2> gpurun_out/c4/push_633.err; echo "push 6/3/3 rc=$?"
tail -1 gpurun_out/c4/push_633.json | cut -c1-250
grep -E "LoaderTimeout|Error" gpurun_out/c4/push_633.err | tail -5 | cut -c1-1500
timeout 150 $TR --master-port 29732 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e > gpurun_out/c4/pull_n2.json 2> gpurun_out/c4/pull_n2.err; echo "pull rc=$?"
tail -1 gpurun_out/c4/pull_n2.json | cut -c1-250
timeout 150 python bench.py --steps 20 --warmup 5 --no-e2e > gpurun_out/c4/pull_n1.json 2> gpurun_out/c4/pull_n1.err; echo "pull1 rc=$?"
tail -1 gpurun_out/c4/pull_n1.json | cut -c1-250
