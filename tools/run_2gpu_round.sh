#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_multigpu.py -q -x -s -k "fastgen_tp2" > gpurun_out/mgpu_tests.log 2>&1
echo "2gpu tests rc=$?"; grep -E "FASTGEN|passed|failed|Error|assert|timed out" gpurun_out/mgpu_tests.log | cut -c1-300 | head -8
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29851 bench.py --gpus 2 --workload fastgen70b --layers 8 --sequences 64 --steps 3 --warmup 2 --gen-tokens 8 > gpurun_out/wl_fastgen64.json 2> gpurun_out/wl_fastgen64.err
echo "fastgen64 rc=$?"; tail -1 gpurun_out/wl_fastgen64.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print({k:d[k] for k in ('value','ms_per_step','gpu_launches','prefill','decode','tp_collectives') if k in d})"
