#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kernels_gpu.py -q -x -k "fp8_block_scaled or rowwise_quantize or block_scaled_fp8 or w8a8" > gpurun_out/kernel_tests.log 2>&1
echo "kernel tests rc=$?"; tail -4 gpurun_out/kernel_tests.log | cut -c1-400
timeout 900 python -m pytest tests/test_peer_one_gpu.py -q -x > gpurun_out/peer_one_gpu.log 2>&1
echo "one-gpu peer rc=$?"; tail -4 gpurun_out/peer_one_gpu.log | cut -c1-600
timeout 900 python -m pytest tests/test_multigpu.py -q -s -k "moe_peer or ulysses" > gpurun_out/mgpu_tests.log 2>&1
echo "pytest rc=$?"; grep -E "MOE_PEER|ULYSSES|passed|failed|skipped|Error|error|assert" gpurun_out/mgpu_tests.log | cut -c1-400 | head -40
timeout 600 python tools/paged_flash_bench.py --out gpurun_out/paged_flash_bench.json > gpurun_out/paged_flash_bench.log 2>&1
echo "paged bench rc=$?"; grep -E "paged_flash_ms|tflops|kv_fraction|speedup|cuda_core|sm_mhz|\"(prefill|decode)" gpurun_out/paged_flash_bench.json | tr -d '\n' | cut -c1-2000
