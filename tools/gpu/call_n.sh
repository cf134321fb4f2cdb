#!/bin/bash
# small models are launch-bound: do the opt-in fusions (norm folded into the GEMMs, rope folded into attention) win there?
O=gpurun_out/r02n; mkdir -p $O
for m in "qwen3-0.6b 4" "qwen3-8b 8"; do set -- $m
  for f in 0 0x20000 0x80000 0xA0000; do
    echo "== $1 rows $2 flags $f" >> $O/fusion_flags.txt
    timeout 200 python tools/step_ablation.py --model $1 --rows $2 --period-only --base-flags $f --sustain-steps 480 2>&1 | grep "full step\|replays" >> $O/fusion_flags.txt
  done
done
cat $O/fusion_flags.txt
