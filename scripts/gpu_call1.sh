#!/bin/bash
# 1-GPU check: GPU tests, the truthful bench, the reworked reference arm (also with 4 instances on one GPU)
mkdir -p gpurun_out/c1
df -h /dev/shm > gpurun_out/c1/env.txt; nproc >> gpurun_out/c1/env.txt; free -g >> gpurun_out/c1/env.txt
ls /sys/devices/system/node >> gpurun_out/c1/env.txt 2>&1
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/c1/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c1/pytest.log
tail -5 gpurun_out/c1/pytest.log
timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/c1/bench.json 2> gpurun_out/c1/bench.err; echo "bench rc=$?"
cat gpurun_out/c1/bench.json | cut -c1-1500
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/c1/ref1.json 2> gpurun_out/c1/ref1.err; echo "ref rc=$?"
cut -c1-400 gpurun_out/c1/ref1.json
BDDL_REF_INSTANCES=4 BDDL_REF_KEEP_LOGS=gpurun_out/c1/ref_logs timeout 400 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/c1/ref4.json 2> gpurun_out/c1/ref4.err; echo "ref4 rc=$?"
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/c1/ref4.json')); print('ref4', d.get('value'), d.get('e2e'), d.get('unavailable'))
except Exception as e: print('ref4 parse', e)
PY
