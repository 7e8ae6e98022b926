#!/bin/bash
# 8 GPUs, short: the two oracle tests that matter most at 8 ranks (push order; replicated window through multicast)
mkdir -p gpurun_out/c14
T=tests/test_loader_gpu.py
timeout 100 python -m pytest -m gpu -q "$T::test_multi_gpu_push_exchange[8]" "$T::test_replicated_window_delivered_to_every_consumer[8-auto]" > gpurun_out/c14/pytest.log 2>&1; echo "pytest rc=$?"
grep -E "passed|failed" gpurun_out/c14/pytest.log | tail -2
grep -A22 "error_rank" gpurun_out/c14/pytest.log | head -60 | cut -c1-400
