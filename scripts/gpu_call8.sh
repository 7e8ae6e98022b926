#!/bin/bash
# 2 GPUs: NVLink efficiency knobs of the pull form (stage-ring geometry, streams in flight)
mkdir -p gpurun_out/c8
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
i=0
for cfg in "--tune 0" "--tune 1" "--tune 2" "--tune 3" "--tune 0 --ring 8 --prefetch-depth 4 --fetch-streams 4" "--tune 2 --ring 8 --prefetch-depth 4 --fetch-streams 4"; do
  i=$((i+1))
  timeout 100 $TR --master-port $((29760+i)) bench.py --gpus 2 --steps 20 --warmup 5 --no-e2e --min-seconds 0.4 $cfg > gpurun_out/c8/pull_$i.json 2> gpurun_out/c8/pull_$i.err
  echo "[$cfg] rc=$? $(python -c "import json,sys; d=json.loads(open('gpurun_out/c8/pull_$i.json').read().strip().splitlines()[-1]); print(d['ms_per_step'], d['exchange']['nvlink_in_gbs_per_gpu'], d['exchange'].get('nvlink_rx_bytes_counted'), d['exchange'].get('nvlink_rx_bytes_algorithmic'))" 2>&1 | tail -1)"
done
