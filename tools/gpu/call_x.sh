#!/bin/bash
# GPU call X (1 GPU): raster group of the CTA-pair GEMM under the final L2 hints; DRAM traffic of the final kernel (ncu after a plain run).
O=gpurun_out/r02x; mkdir -p $O
for G in 12 8 16 24 12; do echo "== group $G" >> $O/gemm_group.log; BCG_GEMM_2SM_GROUP=$G timeout 60 python tools/prefill_gemm_time.py 12390 >> $O/gemm_group.log 2>&1; done
grep "==\|layer total" $O/gemm_group.log
timeout 200 python tools/ncu_targets.py > $O/plain.log 2>&1; echo "rc=$?" >> $O/plain.log
if grep -q "rc=0" $O/plain.log; then
  timeout 300 ncu --set full --clock-control none -k "regex:gemm2sm" -c 4 -o $O/prof_gemm2sm python tools/ncu_targets.py > $O/ncu.log 2>&1; echo "rc=$?" >> $O/ncu.log; tail -2 $O/ncu.log
fi
