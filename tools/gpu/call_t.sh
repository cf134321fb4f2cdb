#!/bin/bash
# GPU call T (1 GPU): the printed numbers of the model-parity and game-replay tests with the final build.
O=gpurun_out/r02t; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_model.py tests/test_gpu_game_replay.py -m gpu -q -s > $O/parity_numbers.txt 2>&1; echo "rc=$?" >> $O/parity_numbers.txt
grep -c "" $O/parity_numbers.txt; tail -3 $O/parity_numbers.txt
