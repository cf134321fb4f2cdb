#!/bin/bash
# GPU call Y (1 GPU): prompt memo in the engine — adapter / games tests and a short bench line.
O=gpurun_out/r02y; mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_adapter.py tests/test_gpu_games.py -m gpu -x -q > $O/tests.log 2>&1; echo "rc=$?" >> $O/tests.log; tail -2 $O/tests.log
timeout 300 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --fixed-rounds 0 > $O/bench.json 2> $O/err.log; echo "rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02y/bench.json'))
print(d["value"], d["e2e"]["value"], d["per_step"])
PY
