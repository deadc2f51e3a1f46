#!/bin/bash
# one 2-GPU box visit: multi-GPU tests, headline bench, shake-down of the config #3 - #5 workloads at reduced depth
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_multigpu.py -q -x -s > gpurun_out/mgpu_tests.log 2>&1
echo "2gpu tests rc=$?"; grep -E "FASTGEN|ULYSSES|passed|failed|Error|assert|timed out" gpurun_out/mgpu_tests.log | cut -c1-300 | head -12
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
PORT=29841
run() { name=$1; shift; timeout 400 $TR --master-port $PORT bench.py --gpus 2 "$@" > gpurun_out/wl2_$name.json 2> gpurun_out/wl2_$name.err; echo "$name rc=$?"; tail -1 gpurun_out/wl2_$name.json | cut -c1-900; PORT=$((PORT+1)); }
run headline --steps 5 --warmup 3
run zeropp --steps 5 --warmup 3 --zeropp
run fastgen_bf16 --workload fastgen70b --layers 8 --sequences 16 --steps 3 --warmup 2 --gen-tokens 8
run fastgen_mx --workload fastgen70b --layers 8 --sequences 16 --steps 3 --warmup 2 --gen-tokens 8 --quant wf8af8_mx
run mixtral --workload mixtral --layers 4 --micro-batch 1 --steps 2 --warmup 1
run ulysses --workload ulysses13b --layers 4 --seq-len 32768 --steps 2 --warmup 1
