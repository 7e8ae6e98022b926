#!/bin/bash
mkdir -p gpurun_out/c11
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
DDL_B200_EXCHANGE=push DDL_B200_MULTICAST=auto timeout 60 $TR --master-port 29781 benchmarks/configs.py --config imagenet --replicate --steps 20 --timeout-s 8 --min-seconds 0.2 --out gpurun_out/c11/configs.jsonl > gpurun_out/c11/rep_auto.log 2>&1; echo "replicate mc=auto rc=$?"
grep -E "LoaderTimeoutError:" gpurun_out/c11/rep_auto.log | sort | uniq | head -8 | cut -c1-900
cat gpurun_out/c11/configs.jsonl 2>/dev/null | cut -c1-400
