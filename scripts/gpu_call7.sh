#!/bin/bash
# 1 GPU: kernel bench of the batch assembly (new convert loop), then ncu captures of every kernel
mkdir -p gpurun_out/c7
timeout 300 python benchmarks/kernel_bench.py --only batch_sweep --out gpurun_out/c7/kernel_batch_sweep.jsonl > gpurun_out/c7/kb.log 2>&1; echo "kb rc=$?"
cat gpurun_out/c7/kernel_batch_sweep.jsonl | cut -c1-400
python benchmarks/ncu_targets.py > gpurun_out/c7/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'gather_|push_tma|index_exchange|slot_iota|perm_gen|flag_|checksum' -o gpurun_out/c7/all_kernels python benchmarks/ncu_targets.py > gpurun_out/c7/ncu_all.log 2>&1
echo "ncu all rc=$?"; tail -3 gpurun_out/c7/plain.log
python benchmarks/ncu_targets.py --steady 12 > gpurun_out/c7/plain2.log 2>&1 && \
ncu --set full --clock-control none --cache-control none -k regex:gather_tma_cvt -s 4 -c 6 -o gpurun_out/c7/cvt_steady python benchmarks/ncu_targets.py --steady 12 > gpurun_out/c7/ncu_steady.log 2>&1
echo "ncu steady rc=$?"
ls -la gpurun_out/c7
