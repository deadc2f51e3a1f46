#!/bin/bash
# one 1-GPU box visit: new-kernel tests first, then the full GPU suite, headline bench, attention head-to-head
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_paged_flash.py tests/test_engine_gpu.py -q -s -k "paged or mixed or small_blocks or split_kv or dispatch or gradients" > gpurun_out/new_tests.log 2>&1
echo "new tests rc=$?"; grep -E "GRADCHECK|passed|failed|Error|error|assert" gpurun_out/new_tests.log | cut -c1-400 | head -30
timeout 1800 python -m pytest tests -m gpu -q -x --deselect tests/test_paged_flash.py > gpurun_out/gpu_tests.log 2>&1
echo "pytest rc=$?"; tail -5 gpurun_out/gpu_tests.log | cut -c1-300
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_own_1gpu.json 2> gpurun_out/bench_own_1gpu.err
echo "bench rc=$?"; tail -1 gpurun_out/bench_own_1gpu.json | cut -c1-1800
timeout 600 python tools/attn_time.py --out gpurun_out/attn_head_to_head.json > gpurun_out/attn_time.log 2>&1
echo "attn rc=$?"; grep -E "fwd_ms|bwd_ms|tflops|unavailable|\"S\"|sm_mhz" gpurun_out/attn_head_to_head.json | tr -d '\n' | cut -c1-2500
