#!/bin/bash
# decode attention with the chunk index as the slowest grid dimension
O=gpurun_out/r02m; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_zz_no_alloc.py -m gpu -x -q -k "attn or no_alloc or static" > $O/tests.log 2>&1; tail -3 $O/tests.log
for spec in "1700 7" "1700 8" "1500 8" "1800 8" "1900 8"; do set -- $spec
  echo "ctx $1 max_splits $2" >> $O/period_vs_splits.txt
  timeout 200 python tools/step_ablation.py --period-only --ctx $1 --max-splits $2 --sustain-steps 240 2>&1 | grep "full step\|replays" >> $O/period_vs_splits.txt
done
cat $O/period_vs_splits.txt
