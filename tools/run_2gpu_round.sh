#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_kernels_gpu.py -q -x -k "fastgen or skinny or gemm_bf16 or linear_autograd" > gpurun_out/graph_tests.log 2>&1
echo "1gpu graph tests rc=$?"; tail -6 gpurun_out/graph_tests.log | cut -c1-400
timeout 600 python -m pytest tests/test_multigpu.py -q -x -s -k "fastgen_tp2 or gemm_allreduce or fused" > gpurun_out/mgpu_tests.log 2>&1
echo "2gpu tests rc=$?"; grep -E "FASTGEN|passed|failed|Error|assert|timed out" gpurun_out/mgpu_tests.log | cut -c1-300 | head -12
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
run() { name=$1; shift; timeout 400 $TR --master-port $PORT bench.py --gpus 2 "$@" > gpurun_out/wl_$name.json 2> gpurun_out/wl_$name.err; echo "$name rc=$?"; tail -1 gpurun_out/wl_$name.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print({k:d[k] for k in ('value','ms_per_step','gpu_launches','prefill','decode') if k in d})"; grep -E "Error|error|Traceback|assert|could not be captured" gpurun_out/wl_$name.err | sort | uniq -c | head -8; PORT=$((PORT+1)); }
PORT=29841
run fastgen_bf16 --workload fastgen70b --layers 8 --sequences 16 --steps 3 --warmup 2 --gen-tokens 8
run fastgen_mx --workload fastgen70b --layers 8 --sequences 16 --steps 3 --warmup 2 --gen-tokens 8 --quant wf8af8_mx
