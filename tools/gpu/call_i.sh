#!/bin/bash
# round-2 decode evidence: in-pipeline ablation (short and sustained period), incumbents with interleaved arms, configs[0]
O=gpurun_out/r02i; mkdir -p $O
timeout 600 python tools/step_ablation.py > $O/step_ablation.txt 2> $O/step_ablation.err
tail -12 $O/step_ablation.txt
timeout 900 python tools/incumbent_kernels.py --out $O/incumbent_kernels.json > $O/incumbent.log 2>&1
timeout 600 python bench.py --model qwen3-0.6b --honest 4 --byzantine 0 --steps 5 --warmup 3 --no-cpu-baseline > $O/bench_qwen3-0.6b_4h0b.json 2> $O/bench06.err
tail -c 600 $O/bench_qwen3-0.6b_4h0b.json
