#!/bin/bash
# GPU call P (2 GPUs): Mistral TP=2 with the reordered decode-attention grid.
O=gpurun_out/r02p; mkdir -p $O
BCG_BENCH_WATCHDOG_S=300 timeout 330 python bench.py --gpus 2 --tp 2 --model mistral-22b --honest 16 --byzantine 4 --steps 2 --warmup 1 --no-cpu-baseline > $O/bench_mistral_tp2.json 2> $O/err.log; echo "rc=$?" >> $O/err.log
tail -n 3 $O/err.log; python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02p/bench_mistral_tp2.json').read().strip().splitlines()[-1])
print(d['value'], d['e2e']); print(d['decode']); print(d['fixed'])
PY
