#!/bin/bash
# first GPU bring-up: non-GEMM kernels, SIMT GEMM, then each tcgen05 GEMM test in its own process
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/smi.txt 2>&1
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -q -x -k "not gemm" > gpurun_out/t_elem.log 2>&1; echo "elem rc=$?"
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -q -k "gemm and (1- or fp32)" > gpurun_out/t_simt.log 2>&1; echo "simt rc=$?"
for t in "test_gemm_kmajor_kmajor and 2-" "test_gemm_b_mnmajor_epilogues and 2-" "test_gemm_wgrad_mnmajor_accumulate and 2-" "test_gemm_grouped_heads and 2" "test_gemm_token_mix_forward_and_dx and 2-" "test_gemm_token_mix_weight_grad and 2" "test_gemm_tc_matches_simt"; do
  n=$(echo "$t" | tr ' ' '_')
  timeout 300 python -m pytest tests/test_kernels_gpu.py -m gpu -q -k "$t" > "gpurun_out/t_tc_$n.log" 2>&1; echo "tc [$t] rc=$?"
done
timeout 600 python benchmarks/kernel_probe.py > gpurun_out/probe.log 2>&1; echo "probe rc=$?"
tail -5 gpurun_out/t_elem.log gpurun_out/t_simt.log
tail -4 gpurun_out/t_tc_*.log
cat gpurun_out/probe.log | tail -30
