#!/bin/bash
# GPU call C4 (2 GPUs): TP=2 bench paths (spawned and torchrun-composed), the fixed schedule under TP, sweep tool smoke.
O=gpurun_out/r02c4; mkdir -p $O
timeout 900 python bench.py --gpus 2 --tp 2 --model mistral-22b --honest 16 --byzantine 4 --steps 1 --warmup 1 --no-cpu-baseline > $O/bench_mistral_tp2.json 2> $O/bench_mistral_tp2.err; echo "rc=$?" >> $O/bench_mistral_tp2.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --tp 2 --model qwen3-8b --honest 8 --byzantine 0 --steps 1 --warmup 1 > $O/bench_8b_tp2_torchrun.json 2> $O/bench_8b_tp2_torchrun.err; echo "rc=$?" >> $O/bench_8b_tp2_torchrun.err
timeout 600 python tools/tp_sweep.py --tp 2 --model qwen3-8b --counts 4,16 --out $O/sweep_8b_tp2.jsonl > $O/sweep_8b_tp2.log 2>&1; echo "rc=$?" >> $O/sweep_8b_tp2.log
for f in $O/*.err $O/sweep_8b_tp2.log; do echo "== $f"; grep -v "^\s*$" $f | tail -n 6; done; head -c 1200 $O/bench_mistral_tp2.json
