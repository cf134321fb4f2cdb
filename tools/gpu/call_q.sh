#!/bin/bash
# GPU call Q (1 GPU): full GPU suite, smoke, default bench line and reference arm with the final build.
O=gpurun_out/r02q; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/suite.log 2>&1; echo "rc=$?" >> $O/suite.log; tail -4 $O/suite.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/smoke.log 2>&1; tail -2 $O/smoke.log
timeout 900 python bench.py > $O/bench_default.json 2> $O/bench_default.err; echo "rc=$?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err; echo "rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02q/bench_default.json'))
print(d["value"], d["e2e"]); print(d["fixed"]); print(d["decode"]); print(d["prefill"]); print(d["roofline"]["frac"], d["clocks"])
PY
