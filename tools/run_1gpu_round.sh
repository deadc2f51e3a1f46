#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py -q -x -k "v1_attention_kv_arena" > gpurun_out/arena_tests.log 2>&1
echo "arena tests rc=$?"; tail -3 gpurun_out/arena_tests.log | cut -c1-300
for w in decode prefill; do
timeout 600 ncu --set full --clock-control none --import-source on -k regex:paged_flash_kernel -c 1 -f -o gpurun_out/ncu_paged_${w}_v3 python tools/paged_flash_bench.py --profile $w > gpurun_out/ncu_paged_${w}_v3.log 2>&1
echo "ncu $w rc=$?"
done
