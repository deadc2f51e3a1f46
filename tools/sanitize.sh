#!/bin/bash
# compute-sanitizer passes over the peer-memory / flag-protocol kernels (csrc/comm/symm.cu, csrc/gemm/gemm_ar.cu, gemm_ag.cu,
# gemm_rs) and the paged flash kernel, driven through their 1-GPU tests.  One log per tool under gpurun_out/sanitizer/;
# tools/sanitize_summary.py condenses them into profiles/sanitizer_r2.md.
mkdir -p gpurun_out/sanitizer
SEL="p2p_collectives_emulated_ranks or ag_gemm_emulated_ranks or fused_gemm_allreduce_single_rank or fused_gemm_reduce_scatter_rows_single_rank"
for tool in memcheck racecheck synccheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 --log-file gpurun_out/sanitizer/kernels_$tool.log \
    python -m pytest tests/test_kernels_gpu.py -q -x -k "$SEL" > gpurun_out/sanitizer/kernels_$tool.pytest 2>&1
  echo "$tool kernels rc=$? $(tail -1 gpurun_out/sanitizer/kernels_$tool.pytest) | $(grep -c 'ERROR SUMMARY' gpurun_out/sanitizer/kernels_$tool.log) summaries: $(grep 'ERROR SUMMARY' gpurun_out/sanitizer/kernels_$tool.log | tail -1)"
done
for tool in memcheck racecheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 --log-file gpurun_out/sanitizer/paged_$tool.log \
    python -m pytest tests/test_paged_flash.py -q -x > gpurun_out/sanitizer/paged_$tool.pytest 2>&1
  echo "$tool paged rc=$? $(tail -1 gpurun_out/sanitizer/paged_$tool.pytest) | $(grep 'ERROR SUMMARY' gpurun_out/sanitizer/paged_$tool.log | tail -1)"
done
