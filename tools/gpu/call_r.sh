#!/bin/bash
# GPU call R (2 GPUs): the driver's N=2 launch of the bench with the final build (one game per GPU + the row-sharded single game).
O=gpurun_out/r02r; mkdir -p $O
BCG_BENCH_WATCHDOG_S=400 timeout 450 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 2 --warmup 1 --no-cpu-baseline > $O/bench_dp2.json 2> $O/err.log; echo "rc=$?" >> $O/err.log
tail -n 3 $O/err.log; python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02r/bench_dp2.json').read().strip().splitlines()[-1])
print(d['value'], d['e2e'], d['value_mean_over_ranks']); print(d['fixed']); print(d['single_game']); print(d['ranks'])
PY
