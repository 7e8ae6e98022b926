#!/bin/bash
# 2 GPUs: the whole GPU suite on the final code, then the 1-GPU bench lines (both arms) kept as raw JSON
mkdir -p gpurun_out/c13
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/c13/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c13/pytest.log
tail -4 gpurun_out/c13/pytest.log
timeout 200 python bench.py --steps 20 --warmup 5 > gpurun_out/c13/bench_n1.json 2> gpurun_out/c13/bench_n1.err; echo "bench rc=$?"; tail -1 gpurun_out/c13/bench_n1.json | cut -c1-300
timeout 200 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/c13/reference_n1.json 2> gpurun_out/c13/reference_n1.err; echo "ref rc=$?"; tail -1 gpurun_out/c13/reference_n1.json | cut -c1-200
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c13/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/c13/smoke.log
