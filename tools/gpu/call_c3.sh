#!/bin/bash
# GPU call C3 (1 GPU): prefill attention with P in TMEM vs shared memory, BN 64 / 128; JSON validity test.
O=gpurun_out/r02c3; mkdir -p $O
for PT in 1 0; do for BN in 64 128; do
  BCG_PREFILL_ATTN_PTMEM=$PT BCG_PREFILL_ATTN_BN=$BN timeout 200 python -m pytest tests/test_gpu_kernels.py -q -k "prefill_attn" > $O/prefill_pt${PT}_bn${BN}.log 2>&1; echo "rc=$?" >> $O/prefill_pt${PT}_bn${BN}.log
  BCG_PREFILL_ATTN_PTMEM=$PT BCG_PREFILL_ATTN_BN=$BN timeout 200 python tools/prefill_attn_time.py > $O/time_pt${PT}_bn${BN}.log 2>&1
done; done
timeout 600 python -m pytest tests/test_gpu_json_validity.py tests/test_gpu_games.py tests/test_local_model_dir.py -m gpu -q -s > $O/newtests.log 2>&1; echo "rc=$?" >> $O/newtests.log
for f in $O/prefill_pt*.log; do echo $f; tail -n 2 $f; done; for f in $O/time_pt*.log; do echo $f; head -n 3 $f; done; tail -n 5 $O/newtests.log
