#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_paged_flash.py -q -x > gpurun_out/pf_tests.log 2>&1
echo "pf tests rc=$?"; tail -8 gpurun_out/pf_tests.log | cut -c1-400
timeout 300 python tools/paged_flash_bench.py --out gpurun_out/paged_flash_bench2.json > gpurun_out/pf_bench.log 2>&1
echo "pf bench rc=$?"; grep -E "paged_flash_ms|kv_fraction|tflops|\"(prefill|decode)" gpurun_out/pf_bench.log | cut -c1-200
DSB200_PF_TRACE=1 timeout 300 python tools/paged_flash_bench.py --profile decode > gpurun_out/pf_trace_decode.log 2>&1
echo "trace decode rc=$?"; tail -8 gpurun_out/pf_trace_decode.log | cut -c1-60
