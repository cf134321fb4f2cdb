#!/bin/bash
# GPU call C9 (2 GPUs): TP replicas inside a torchrun job (with a watchdog), Mistral TP=2 bench incl. the fixed schedule.
O=gpurun_out/r02c9; mkdir -p $O
BCG_BENCH_WATCHDOG_S=150 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --tp 2 --model qwen3-8b --honest 8 --byzantine 0 --steps 1 --warmup 1 --fixed-rounds 0 > $O/bench_8b_tp2_torchrun.json 2> $O/bench_8b_tp2_torchrun.err; echo "rc=$?" >> $O/bench_8b_tp2_torchrun.err
BCG_BENCH_WATCHDOG_S=400 timeout 450 python bench.py --gpus 2 --tp 2 --model mistral-22b --honest 16 --byzantine 4 --steps 2 --warmup 1 --no-cpu-baseline > $O/bench_mistral_tp2.json 2> $O/bench_mistral_tp2.err; echo "rc=$?" >> $O/bench_mistral_tp2.err
grep -v "^\s*$" $O/bench_8b_tp2_torchrun.err | grep -v "^\[W\|OMP_NUM\|^\*\*\*" | tail -n 60; tail -n 3 $O/bench_mistral_tp2.err; head -c 300 $O/bench_8b_tp2_torchrun.json
