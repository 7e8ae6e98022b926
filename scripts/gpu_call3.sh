#!/bin/bash
# 2-GPU: diagnose the push bench hang; 1-GPU kernels tests for new perm_gen + convert loop
mkdir -p gpurun_out/c3
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_loader_gpu.py tests/test_varlen.py -m gpu -x -q > gpurun_out/c3/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c3/pytest.log
tail -4 gpurun_out/c3/pytest.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
i=0
for cfg in "--ring 4 --prefetch-depth 2 --fetch-streams 2" "--ring 6 --prefetch-depth 3 --fetch-streams 3" "--ring 6 --prefetch-depth 3 --fetch-streams 1" "--ring 3 --prefetch-depth 1 --fetch-streams 1"; do
  i=$((i+1))
  timeout 100 $TR --master-port $((29720+i)) bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e --exchange push --min-seconds 0.2 --timeout-s 15 --watchdog 60 $cfg > gpurun_out/c3/push_$i.json 2> gpurun_out/c3/push_$i.err; echo "push[$cfg] rc=$?"
  tail -1 gpurun_out/c3/push_$i.json | cut -c1-250
  grep -E "LoaderTimeout|Error|File \"/.*(bench|ddl_b200)" gpurun_out/c3/push_$i.err | tail -12 | cut -c1-400
done
