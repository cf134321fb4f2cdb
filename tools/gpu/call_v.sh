#!/bin/bash
# GPU call V (1 GPU): bench line sanity after the last bench.py edit.
O=gpurun_out/r02v; mkdir -p $O
timeout 400 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --fixed-rounds 0 > $O/bench.json 2> $O/err.log; echo "rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02v/bench.json'))
print(d["value"], d["e2e"]["value"], d["roofline"])
PY
