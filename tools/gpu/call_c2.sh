#!/bin/bash
# GPU call C2 (2 GPUs): tensor-parallel parity (fused push-reduce, vocab-parallel sampler, NCCL path), TP=2 bench,
# data-parallel bench with the row-sharded single game, static-scratch entry on a second device.
O=gpurun_out/r02c2; mkdir -p $O
nvidia-smi --query-gpu=index,name --format=csv > $O/gpus.txt 2>&1
timeout 600 python -m pytest tests/test_gpu_tp.py tests/test_zz_no_alloc.py -m gpu -q -rs -s > $O/tp_tests.log 2>&1; echo "rc=$?" >> $O/tp_tests.log
timeout 900 python bench.py --gpus 2 --tp 2 --model mistral-22b --honest 16 --byzantine 4 --steps 2 --warmup 1 --no-cpu-baseline > $O/bench_mistral_tp2.json 2> $O/bench_mistral_tp2.err; echo "rc=$?" >> $O/bench_mistral_tp2.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 2 --warmup 1 > $O/bench_14b_dp2.json 2> $O/bench_14b_dp2.err; echo "rc=$?" >> $O/bench_14b_dp2.err
tail -n 12 $O/tp_tests.log; tail -n 3 $O/bench_mistral_tp2.err $O/bench_14b_dp2.err; head -c 1500 $O/bench_mistral_tp2.json
