#!/bin/bash
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:paged_decode_kernel -c 1 -f -o gpurun_out/ncu_paged_decode_v4 python tools/paged_flash_bench.py --profile decode > gpurun_out/ncu_paged_decode_v4.log 2>&1
echo "ncu decode rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:paged_flash_kernel -c 1 -f -o gpurun_out/ncu_paged_prefill_v4 python tools/paged_flash_bench.py --profile prefill > gpurun_out/ncu_paged_prefill_v4.log 2>&1
echo "ncu prefill rc=$?"
timeout 1800 python -m pytest tests -m gpu -q -x > gpurun_out/gpu_tests.log 2>&1
echo "pytest rc=$?"; tail -5 gpurun_out/gpu_tests.log | cut -c1-300
