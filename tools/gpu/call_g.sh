#!/bin/bash
# GPU call G (2 GPUs): Mistral TP=2 with qk-norm/RoPE folded into decode attention (one launch fewer per layer).
O=gpurun_out/r02g; mkdir -p $O
BCG_BENCH_WATCHDOG_S=300 timeout 330 python bench.py --gpus 2 --tp 2 --model mistral-22b --honest 16 --byzantine 4 --steps 1 --warmup 1 --no-cpu-baseline --fused-rope 1 > $O/bench_mistral_tp2_fusedrope.json 2> $O/err.log; echo "rc=$?" >> $O/err.log
tail -n 3 $O/err.log; python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02g/bench_mistral_tp2_fusedrope.json').read().strip().splitlines()[-1])
print(d['decode'], d['fixed'])
PY
