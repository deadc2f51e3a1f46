#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py -q -x -k "mxfp8 or transformer_layer_bf16" > gpurun_out/mx_tests.log 2>&1
echo "mx tests rc=$?"; grep -E "passed|failed|Error|error|assert|rel" gpurun_out/mx_tests.log | cut -c1-400 | head -20
timeout 300 python tools/gemm_fp8_bench.py --out gpurun_out/gemm_fp8_bench_r2.json > gpurun_out/gemm_fp8_bench.log 2>&1
echo "fp8 bench rc=$?"; cut -c1-700 gpurun_out/gemm_fp8_bench.log
