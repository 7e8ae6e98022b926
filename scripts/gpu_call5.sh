#!/bin/bash
mkdir -p gpurun_out/c5
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/c5/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c5/pytest.log
tail -4 gpurun_out/c5/pytest.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 150 $TR --master-port 29741 bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e --exchange push --timeout-s 10 --watchdog 120 > gpurun_out/c5/push_633.json 2> gpurun_out/c5/push_633.err; echo "push 6/3/3 rc=$?"
tail -1 gpurun_out/c5/push_633.json | cut -c1-250
grep -E "LoaderTimeout" gpurun_out/c5/push_633.err | tail -4 | cut -c1-2500
