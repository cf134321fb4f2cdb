#!/bin/bash
# GPU call W (1 GPU): device-side timeline and in-pipeline ablation of the decode step, final build.
O=gpurun_out/r02w; mkdir -p $O
timeout 200 python tools/timeline.py > $O/timeline.txt 2> $O/timeline.err; echo "rc=$?"
timeout 300 python tools/step_ablation.py --sustain-steps 480 > $O/step_ablation.txt 2> $O/step_ablation.err; echo "rc=$?"
tail -12 $O/step_ablation.txt
