#!/bin/bash
# one 2-GPU box visit: multi-GPU tests, own bench (+ZeRO++), step breakdown (logs under gpurun_out/)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_multigpu.py -x -q -s -k "nvls or equals or shard_write or zeropp or two_gpus" > gpurun_out/mgpu_tests.log 2>&1
echo "pytest rc=$?"; grep -E "SYMM|RS |AG |ZERO3|ZEROPP|passed|failed|skipped|Error|error|assert" gpurun_out/mgpu_tests.log | cut -c1-400 | head -40
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29601 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_own_2gpu.json 2> gpurun_out/bench_own_2gpu.err
echo "own rc=$?"; tail -1 gpurun_out/bench_own_2gpu.json | cut -c1-2500
timeout 600 $TR --master-port 29603 bench.py --gpus 2 --steps 5 --warmup 3 --zeropp > gpurun_out/bench_own_2gpu_zeropp.json 2> gpurun_out/bench_own_2gpu_zeropp.err
echo "own zeropp rc=$?"; tail -1 gpurun_out/bench_own_2gpu_zeropp.json | cut -c1-2500; tail -5 gpurun_out/bench_own_2gpu_zeropp.err | cut -c1-300
timeout 600 $TR --master-port 29604 tools/step_breakdown.py --out gpurun_out/step_breakdown_7b_2gpu.txt > gpurun_out/step_breakdown.log 2>&1
echo "breakdown rc=$?"; head -25 gpurun_out/step_breakdown_7b_2gpu.txt | cut -c1-160
