#!/bin/bash
# One 8-GPU session: headline bench (pull and push), exchange sweep vs NCCL, the other configs,
# ResNet-50 stall, replicated delivery.  Every command has its own timeout.
cd "$(dirname "$0")/.."
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
O=gpurun_out
mkdir -p $O
run() { local t=$1; shift; echo "== $*" ; timeout $t "$@" 2>&1 | grep -E '^\{' | cut -c1-1600; }
run 240 $TR --master-port 29601 bench.py --gpus $N --steps 300 --warmup 10 | tee $O/bench_pull_n$N.json
run 200 $TR --master-port 29602 bench.py --gpus $N --steps 300 --warmup 10 --exchange push --no-e2e | tee $O/bench_push_n$N.json
run 200 $TR --master-port 29603 benchmarks/exchange_sweep.py --sizes 4096,65536,1048576,16777216,67108864 --out $O/exchange_sweep_n$N.jsonl
run 150 $TR --master-port 29604 benchmarks/configs.py --config gpt2 --out $O/configs_n$N.jsonl
run 150 $TR --master-port 29605 benchmarks/configs.py --config volume --out $O/configs_n$N.jsonl
run 150 $TR --master-port 29606 benchmarks/configs.py --config volume_bf16 --out $O/configs_n$N.jsonl
if [ $N -ge 4 ]; then
  run 150 $TR --master-port 29607 benchmarks/configs.py --config volume --storage $((N/4)) --out $O/configs_n$N.jsonl
fi
run 240 $TR --master-port 29608 benchmarks/train_stall.py --steps 20 --warmup 6 --out $O/train_stall_n$N.jsonl
run 120 $TR --master-port 29609 benchmarks/multicast_bench.py | tee $O/multicast_n$N.json
