#!/bin/bash
# 4 GPUs: diagnose the 8-rank failures (push test exit 1, replicated config hang) with tracebacks and short timeouts
mkdir -p gpurun_out/c10
T=tests/test_loader_gpu.py
timeout 150 python -m pytest -m gpu -q "$T::test_multi_gpu_push_exchange[4]" "$T::test_multi_gpu_global_shuffle_over_nvlink[4-flags]" > gpurun_out/c10/pytest.log 2>&1; echo "pytest rc=$?"
grep -E "passed|failed" gpurun_out/c10/pytest.log | tail -2
grep -A25 "error_rank" gpurun_out/c10/pytest.log | head -70 | cut -c1-300
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
p=29770
for mc in off auto; do
  p=$((p+1))
  DDL_B200_EXCHANGE=push DDL_B200_MULTICAST=$mc timeout 70 $TR --master-port $p benchmarks/configs.py --config imagenet --replicate --steps 20 --timeout-s 10 --min-seconds 0.2 --out gpurun_out/c10/configs.jsonl > gpurun_out/c10/rep_$mc.log 2>&1; echo "replicate mc=$mc rc=$?"
  grep -E "LoaderTimeout|Error" gpurun_out/c10/rep_$mc.log | head -3 | cut -c1-1500
done
cat gpurun_out/c10/configs.jsonl 2>/dev/null | cut -c1-500
