#!/bin/bash
# GPU call Z (1 GPU): bench line of the round's last commit.
O=gpurun_out/r02z; mkdir -p $O
timeout 200 python bench.py --steps 3 --warmup 1 --no-cpu-baseline > $O/bench.json 2> $O/err.log; echo "rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02z/bench.json'))
print(d["value"], d["e2e"]["value"], d["fixed"], d["per_step"]["host_ms"], d["decode"]["ms_per_step"])
PY
