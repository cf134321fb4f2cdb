#!/bin/bash
# GPU call C5 (1 GPU): dead row under the fixed schedule with the Mistral prompts — single GPU first.
O=gpurun_out/r02c5; mkdir -p $O
timeout 300 python tools/debug_dead_rows.py --model mistral-22b --honest 16 --byzantine 4 --layers 4 > $O/dead_L4.log 2>&1; echo "rc=$?" >> $O/dead_L4.log
timeout 300 python tools/debug_dead_rows.py --model qwen3-14b --honest 8 --byzantine 2 --layers 4 > $O/dead_qwen_L4.log 2>&1; echo "rc=$?" >> $O/dead_qwen_L4.log
tail -n 30 $O/dead_L4.log; tail -n 12 $O/dead_qwen_L4.log
