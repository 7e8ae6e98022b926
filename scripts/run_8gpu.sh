#!/bin/bash
# One 8-GPU session (round 2).  Most important first; every command has its own timeout.
cd "$(dirname "$0")/.."
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
O=gpurun_out/n$N
mkdir -p $O
t0=$(date +%s)
stamp() { echo "-- $(( $(date +%s) - t0 ))s: $*"; }
last() { tail -1 "$1" | cut -c1-${2:-600}; }

nvidia-smi nvlink -gt d -i 0 2>&1 | head -8
stamp "bench pull (value + e2e)"
timeout 200 $TR --master-port 29601 bench.py --gpus $N --steps 20 --warmup 5 > $O/bench_pull.json 2> $O/bench_pull.err; echo "rc=$?"; last $O/bench_pull.json 900
stamp "reference arm"
BDDL_REF_WIRING_FIX=0 BDDL_REF_KEEP_LOGS=$O/ref_logs timeout 300 $TR --master-port 29602 bench.py --impl reference --gpus $N --steps 20 --warmup 5 > $O/reference.json 2> $O/reference.err; echo "rc=$?"; last $O/reference.json 400
python - <<PY
import json
try:
    d=json.loads(open("$O/reference.json").read().strip().splitlines()[-1]); print("reference value", d.get("value"), "e2e", d.get("e2e"), d.get("unavailable"))
except Exception as e: print("reference parse", e)
PY
stamp "bench push"
timeout 150 $TR --master-port 29603 bench.py --gpus $N --steps 20 --warmup 5 --no-e2e --exchange push --timeout-s 30 > $O/bench_push.json 2> $O/bench_push.err; echo "rc=$?"; last $O/bench_push.json 400
grep -E "LoaderTimeout" $O/bench_push.err | tail -1 | cut -c1-800
stamp "multi-GPU tests at $N ranks"
T=tests/test_loader_gpu.py
timeout 240 python -m pytest -m gpu -q -x "$T::test_multi_gpu_push_exchange[$N]" "$T::test_multi_gpu_u8_to_bf16_with_integrity_mode[$N-pull]" "$T::test_replicated_window_delivered_to_every_consumer[$N-auto]" > $O/pytest.log 2>&1; echo "rc=$?"; tail -3 $O/pytest.log
stamp "h2d probe"
timeout 90 $TR --master-port 29604 benchmarks/h2d_probe.py > $O/h2d_probe.json 2> $O/h2d_probe.err; echo "rc=$?"; last $O/h2d_probe.json 900
stamp "exchange sweep copy / u8bf16"
timeout 120 $TR --master-port 29605 benchmarks/exchange_sweep.py --sizes 4096,65536,1048576,16777216,67108864 --out $O/exchange_sweep_copy.jsonl > $O/sweep_copy.log 2>&1; echo "rc=$?"; cut -c1-330 $O/exchange_sweep_copy.jsonl
timeout 120 $TR --master-port 29606 benchmarks/exchange_sweep.py --mode u8bf16 --sizes 4096,150528,1048576,16777216 --out $O/exchange_sweep_u8bf16.jsonl > $O/sweep_u8.log 2>&1; echo "rc=$?"; cut -c1-330 $O/exchange_sweep_u8bf16.jsonl
stamp "replicated window: multicast vs unicast"
DDL_B200_EXCHANGE=push DDL_B200_MULTICAST=auto timeout 100 $TR --master-port 29607 benchmarks/configs.py --config imagenet --replicate --steps 20 --out $O/configs.jsonl > $O/rep_mc.log 2>&1; echo "rc=$?"
DDL_B200_EXCHANGE=push DDL_B200_MULTICAST=off timeout 100 $TR --master-port 29608 benchmarks/configs.py --config imagenet --replicate --steps 20 --out $O/configs.jsonl > $O/rep_uc.log 2>&1; echo "rc=$?"
stamp "configs gpt2 / volume / 1:3 storage"
timeout 100 $TR --master-port 29609 benchmarks/configs.py --config gpt2 --out $O/configs.jsonl > $O/gpt2.log 2>&1; echo "rc=$?"
timeout 100 $TR --master-port 29610 benchmarks/configs.py --config volume --out $O/configs.jsonl > $O/volume.log 2>&1; echo "rc=$?"
if [ $N -ge 4 ]; then
  timeout 100 $TR --master-port 29611 benchmarks/configs.py --config volume --storage $((N/4)) --out $O/configs.jsonl > $O/volume13.log 2>&1; echo "rc=$?"
fi
cut -c1-420 $O/configs.jsonl
stamp "ResNet-50 stall"
timeout 200 $TR --master-port 29612 benchmarks/train_stall.py --steps 20 --warmup 6 --out $O/train_stall.jsonl > $O/train_stall.log 2>&1; echo "rc=$?"; cat $O/train_stall.jsonl
stamp "done"
