#!/bin/bash
# 8 GPUs, short: replicated window through the multicast push kernel against the oracle
mkdir -p gpurun_out/c14
T=tests/test_loader_gpu.py
timeout 60 python -m pytest -m gpu -q "$T::test_replicated_window_delivered_to_every_consumer[8-auto]" > gpurun_out/c14/pytest.log 2>&1; echo "pytest rc=$?"
grep -E "passed|failed" gpurun_out/c14/pytest.log | tail -2
grep -A14 "error_rank" gpurun_out/c14/pytest.log | head -30 | cut -c1-300
