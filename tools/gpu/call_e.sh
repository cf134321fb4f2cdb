#!/bin/bash
# GPU call E (1 GPU): prefill GEMM on CTA pairs (tcgen05.mma.cta_group::2): correctness, then time vs the single-CTA kernel.
O=gpurun_out/r02e; mkdir -p $O
timeout 200 python -m pytest tests/test_gpu_kernels.py -q -k "gemm or qknorm_rope" > $O/gemm_tests.log 2>&1; RC=$?; echo "rc=$RC" >> $O/gemm_tests.log
BCG_GEMM_2SM=0 timeout 100 python tools/prefill_gemm_time.py 12390 > $O/gemm_time_1sm.log 2>&1
if [ $RC -eq 0 ]; then timeout 100 python tools/prefill_gemm_time.py 12390 > $O/gemm_time_2sm.log 2>&1; fi
grep -E "passed|failed|^E  |Error" $O/gemm_tests.log | head -n 12; cat $O/gemm_time_1sm.log $O/gemm_time_2sm.log 2>/dev/null
