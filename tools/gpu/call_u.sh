#!/bin/bash
# GPU call U (1 GPU): ncu --set full of the decode step's kernels in the final build (10 rows, 1,600 tokens of context), plain run first.
O=gpurun_out/r02u; mkdir -p $O
timeout 200 python tools/ncu_targets.py 10 1600 > $O/plain.log 2>&1; echo "rc=$?" >> $O/plain.log; tail -2 $O/plain.log
if grep -q "rc=0" $O/plain.log; then
  timeout 500 ncu --set full --clock-control none --import-source on -k "regex:gemm_tcgen05_kernel|decode_attn|add_norm_stats|qknorm_rope_kvwrite_kernel" --launch-skip 7 -c 14 -o $O/prof_decode python tools/ncu_targets.py 10 1600 > $O/ncu.log 2>&1; echo "rc=$?" >> $O/ncu.log; tail -3 $O/ncu.log
fi
