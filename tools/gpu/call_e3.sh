#!/bin/bash
O=gpurun_out/r02e3; mkdir -p $O
for G in 12 16 24 32; do BCG_GEMM_2SM_GROUP=$G timeout 60 python tools/prefill_gemm_time.py 12390 > $O/gemm_time_g$G.log 2>&1; done
BCG_GEMM_2SM_STAGES=7 timeout 60 python tools/prefill_gemm_time.py 12390 > $O/gemm_time_s7.log 2>&1
BCG_GEMM_2SM_STAGES=7 timeout 100 python -m pytest tests/test_gpu_kernels.py -q -k "gemm" > $O/tests_s7.log 2>&1
for f in $O/gemm_time_*.log; do echo $f; tail -n 1 $f; done; tail -n 2 $O/tests_s7.log
