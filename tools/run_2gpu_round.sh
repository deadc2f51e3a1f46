#!/bin/bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
$TR --master-port 29611 tools/determinism_probe.py --reps 2 2>&1 | grep -E "instance|rep |Error" | cut -c1-400
$TR --master-port 29612 tools/determinism_probe.py --reps 2 --no-p2p 2>&1 | grep -E "instance|rep |Error" | cut -c1-400
timeout 1500 python -m pytest tests/test_multigpu.py -q -s -k "equals or shard_write or zeropp or fused_reduce or fastgen" > gpurun_out/mgpu_tests.log 2>&1
echo "pytest rc=$?"; grep -E "SYMM|ZERO3|ZEROPP|FASTGEN|  unit |passed|failed|skipped|Error|error|assert" gpurun_out/mgpu_tests.log | cut -c1-400 | head -60
DSB200_PROBE_TRACE=1 timeout 600 $TR --master-port 29620 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_own_2gpu.json 2> gpurun_out/bench_own_2gpu.err
tail -1 gpurun_out/bench_own_2gpu.json | cut -c1-3000
timeout 600 $TR --master-port 29621 bench.py --gpus 2 --steps 5 --warmup 3 --zeropp > gpurun_out/bench_own_2gpu_zeropp.json 2> gpurun_out/bench_own_2gpu_zeropp.err
tail -1 gpurun_out/bench_own_2gpu_zeropp.json | cut -c1-2200
timeout 600 $TR --master-port 29604 tools/step_breakdown.py --out gpurun_out/step_breakdown_7b_2gpu.txt > gpurun_out/step_breakdown.log 2>&1
head -14 gpurun_out/step_breakdown_7b_2gpu.txt | cut -c1-160
