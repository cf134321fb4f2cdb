#!/bin/bash
O=gpurun_out/r02c6b; mkdir -p $O
timeout 200 python -m pytest tests/test_gpu_kernels.py -q -k "prefill_attn" > $O/prefill.log 2>&1; echo "rc=$?" >> $O/prefill.log
timeout 200 python tools/debug_dead_rows.py --model mistral-22b --honest 16 --byzantine 4 --layers 4 > $O/dead_L4.log 2>&1; echo "rc=$?" >> $O/dead_L4.log
grep -E "passed|failed|^E  " $O/prefill.log | head -n 10; tail -n 12 $O/dead_L4.log
