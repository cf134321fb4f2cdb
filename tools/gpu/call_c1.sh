#!/bin/bash
# GPU call C1 (1 GPU): new tests (JSON validity, games driver, local model dir), attention retime + ncu, suite with printed parity numbers.
O=gpurun_out/r02c1; mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_kernels.py -q -k "prefill_attn" > $O/prefill.log 2>&1; echo "rc=$?" >> $O/prefill.log
timeout 300 python tools/prefill_attn_time.py $O/prefill_attn_time.json > $O/prefill_attn_time.log 2>&1
timeout 1500 python -m pytest tests -m gpu -q -rs > $O/suite.log 2>&1; echo "rc=$?" >> $O/suite.log
timeout 900 python -m pytest tests/test_gpu_model.py tests/test_gpu_game_replay.py tests/test_gpu_json_validity.py -m gpu -q -s -k "full_width or full_depth or replays or stopped" > $O/parity_numbers.log 2>&1
timeout 200 python tools/prefill_attn_time.py > $O/ncu_plain1.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:prefill_attn -s 3 -c 1 -o $O/prof_prefill_attn python tools/prefill_attn_time.py > $O/ncu1.log 2>&1
timeout 600 python bench.py --steps 3 --warmup 1 --no-cpu-baseline > $O/bench_s3.json 2> $O/bench_s3.err
cat $O/prefill_attn_time.log; tail -n 6 $O/suite.log; grep -E "max \|dlogit\||generations|sampled rows" $O/parity_numbers.log | tail -n 20
