#!/bin/bash
# one 1-GPU box visit: the GPU test suite, smoke(), the headline bench (own and reference arm)
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x > gpurun_out/gpu_tests.log 2>&1
echo "pytest rc=$?"; tail -3 gpurun_out/gpu_tests.log | cut -c1-300
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1
echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log | cut -c1-200
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_own_1gpu.json 2> gpurun_out/bench_own_1gpu.err
echo "bench rc=$?"; tail -1 gpurun_out/bench_own_1gpu.json | cut -c1-700
[ -n "$WITH_REFERENCE" ] && { timeout 1500 python bench.py --impl reference --steps 5 --warmup 3 > gpurun_out/bench_ref_1gpu.json 2> gpurun_out/bench_ref_1gpu.err; echo "ref rc=$?"; tail -1 gpurun_out/bench_ref_1gpu.json | cut -c1-400; }
