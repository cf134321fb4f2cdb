#!/bin/bash
# idle attention chunks leave before the dependency wait; 2-SM GEMM hints: tests + bench
O=gpurun_out/r02l; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_zz_no_alloc.py tests/test_gpu_model.py -m gpu -x -q > $O/tests.log 2>&1; tail -3 $O/tests.log
for spec in "1700 8" "1500 8"; do set -- $spec
  echo "ctx $1 max_splits $2" >> $O/period_vs_splits.txt
  timeout 200 python tools/step_ablation.py --period-only --ctx $1 --max-splits $2 --sustain-steps 240 2>&1 | grep "full step\|replays" >> $O/period_vs_splits.txt
done
cat $O/period_vs_splits.txt
timeout 120 python tools/prefill_gemm_time.py 12390 > $O/prefill_gemm_time.log 2>&1; tail -1 $O/prefill_gemm_time.log
timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $O/bench_s2.json 2> $O/bench_s2.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02l/bench_s2.json'))
print(d["value"], d["e2e"]); print(d["fixed"]); print(d["decode"]); print(d["prefill"])
PY
