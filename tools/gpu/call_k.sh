#!/bin/bash
# where the fixed schedule's 5.62 ms per step comes from: per-phase decode time, idle attention CTAs
O=gpurun_out/r02k; mkdir -p $O
for spec in "1700 7" "1700 8" "1500 6" "1500 8"; do set -- $spec
  echo "ctx $1 max_splits $2" >> $O/period_vs_splits.txt
  timeout 200 python tools/step_ablation.py --period-only --ctx $1 --max-splits $2 --sustain-steps 240 2>&1 | grep "full step\|replays" >> $O/period_vs_splits.txt
done
cat $O/period_vs_splits.txt
timeout 600 python bench.py --steps 1 --warmup 1 --no-cpu-baseline > $O/bench_s1.json 2> $O/bench_s1.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02k/bench_s1.json'))
print(d["fixed"]); print(d["decode"])
PY
