#!/bin/bash
# GPU call D (8 GPUs): TP=8 / TP=4 parity, Qwen3-32B TP=8 agent-count sweep (configs[4]), Mistral-22B TP=2 x 4 replicas (configs[3]).
O=gpurun_out/r02d; mkdir -p $O
nvidia-smi --query-gpu=index,name --format=csv > $O/gpus.txt 2>&1
timeout 240 python -m pytest tests/test_gpu_tp.py -m gpu -q -rs -s -k "tp8 or tp4" > $O/tp_tests.log 2>&1; echo "rc=$?" >> $O/tp_tests.log
timeout 420 python tools/tp_sweep.py --tp 8 --model qwen3-32b --counts 10,4,8,16,32,64 --out $O/sweep_32b_tp8.jsonl > $O/sweep_32b_tp8.log 2>&1; echo "rc=$?" >> $O/sweep_32b_tp8.log
BCG_BENCH_WATCHDOG_S=280 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 8 --tp 2 --model mistral-22b --honest 16 --byzantine 4 --steps 2 --warmup 1 > $O/bench_mistral_tp2x4.json 2> $O/bench_mistral_tp2x4.err; echo "rc=$?" >> $O/bench_mistral_tp2x4.err
grep -E "passed|failed|worst" $O/tp_tests.log | tail -n 6; grep -v "^\s*$" $O/sweep_32b_tp8.log | tail -n 9 | cut -c1-600; grep -v "^\s*$" $O/bench_mistral_tp2x4.err | grep -v "OMP_NUM\|^\*\*\*\|Warning\|warn" | tail -n 12; head -c 700 $O/bench_mistral_tp2x4.json
