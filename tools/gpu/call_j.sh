#!/bin/bash
# 2-SM prefill GEMM: L2 eviction hints / raster group; decode step period vs context (7 vs 8 attention splits)
O=gpurun_out/r02j; mkdir -p $O
run() { echo "== $1" >> $O/gemm_variants.log; env $1 timeout 120 python tools/prefill_gemm_time.py 12390 >> $O/gemm_variants.log 2>&1; }
run "BCG_X=0"
run "BCG_GEMM_2SM_HINT_A=2 BCG_GEMM_2SM_HINT_B=0"
run "BCG_GEMM_2SM_HINT_A=2 BCG_GEMM_2SM_HINT_B=1"
run "BCG_GEMM_2SM_HINT_A=2 BCG_GEMM_2SM_HINT_B=1 BCG_GEMM_2SM_GROUP=16"
run "BCG_GEMM_2SM_HINT_A=2 BCG_GEMM_2SM_HINT_B=1 BCG_GEMM_2SM_GROUP=8"
run "BCG_GEMM_2SM_HINT_A=0 BCG_GEMM_2SM_HINT_B=0"
run "BCG_X=0"
grep "==\|layer total" $O/gemm_variants.log
for c in 1600 1780 1800 1900; do
  timeout 200 python tools/step_ablation.py --period-only --ctx $c --sustain-steps 240 2>&1 | grep "full step\|1 token\|replays" >> $O/period_vs_ctx.txt
done
cat $O/period_vs_ctx.txt
