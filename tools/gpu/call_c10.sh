#!/bin/bash
# GPU call C10 (2 GPUs): LL-style push all-reduce: TP parity tests, Mistral TP=2 bench, TP replica inside torchrun.
O=gpurun_out/r02c10; mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_tp.py -m gpu -q -rs -s > $O/tp_tests.log 2>&1; echo "rc=$?" >> $O/tp_tests.log
BCG_BENCH_WATCHDOG_S=400 timeout 450 python bench.py --gpus 2 --tp 2 --model mistral-22b --honest 16 --byzantine 4 --steps 2 --warmup 1 --no-cpu-baseline > $O/bench_mistral_tp2.json 2> $O/bench_mistral_tp2.err; echo "rc=$?" >> $O/bench_mistral_tp2.err
BCG_BENCH_WATCHDOG_S=200 timeout 250 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --tp 2 --model qwen3-8b --honest 8 --byzantine 0 --steps 1 --warmup 1 > $O/bench_8b_tp2_torchrun.json 2> $O/bench_8b_tp2_torchrun.err; echo "rc=$?" >> $O/bench_8b_tp2_torchrun.err
grep -E "passed|failed|worst" $O/tp_tests.log | tail -n 8; tail -n 3 $O/bench_mistral_tp2.err; grep -v "^\s*$" $O/bench_8b_tp2_torchrun.err | grep -v "OMP_NUM\|^\*\*\*" | tail -n 25
