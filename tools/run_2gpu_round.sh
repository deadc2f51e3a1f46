#!/bin/bash
# one 2-GPU box visit: reference repro, multi-GPU tests, own + reference bench (logs under gpurun_out/)
mkdir -p gpurun_out
bash tools/ref_repro.sh 2 base nockpt > gpurun_out/ref_repro_summary.txt 2>&1
cat gpurun_out/ref_repro_summary.txt
timeout 900 python -m pytest tests/test_multigpu.py -x -q -s -k "nvls or equals or shard_write or symmetric or two_gpus" > gpurun_out/mgpu_tests.log 2>&1
echo "pytest rc=$?"; grep -E "SYMM|RS |AG |ZERO3|passed|failed|skipped|Error|error" gpurun_out/mgpu_tests.log | head -30
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29601 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_own_2gpu.json 2> gpurun_out/bench_own_2gpu.err
echo "own rc=$?"; tail -1 gpurun_out/bench_own_2gpu.json | cut -c1-1500
timeout 900 $TR --master-port 29602 bench.py --impl reference --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_ref_2gpu.json 2> gpurun_out/bench_ref_2gpu.err
echo "ref rc=$?"; tail -1 gpurun_out/bench_ref_2gpu.json | cut -c1-1200
