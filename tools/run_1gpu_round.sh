#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py tests/test_model_ops.py -q -x -k "evoformer" > gpurun_out/evo_tests.log 2>&1
echo "evo rc=$?"; tail -12 gpurun_out/evo_tests.log | cut -c1-300
