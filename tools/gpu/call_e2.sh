#!/bin/bash
O=gpurun_out/r02e2; mkdir -p $O
timeout 200 python -m pytest tests/test_gpu_kernels.py -q -k "gemm or qknorm_rope" > $O/tests.log 2>&1; echo "rc=$?" >> $O/tests.log
for G in 2 4 16 49; do BCG_GEMM_2SM_GROUP=$G timeout 60 python tools/prefill_gemm_time.py 12390 > $O/gemm_time_g$G.log 2>&1; done
BCG_ROPE_TOKEN_KERNEL=0 timeout 100 python tools/ncu_targets.py > /dev/null 2>&1
timeout 100 python tools/prefill_norm_time.py > $O/norm_time.log 2>&1
grep -E "passed|failed|^E  " $O/tests.log | head; for G in 2 4 16 49; do echo "group $G"; cat $O/gemm_time_g$G.log; done; cat $O/norm_time.log
