#!/bin/bash
# GPU call A (1 GPU): validate the tcgen05 prefill attention, run the suite, time incumbents, short bench.
O=gpurun_out/r02a; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > $O/gpu.txt 2>&1
timeout 300 python -m pytest tests/test_gpu_kernels.py -q -k "prefill_attn" > $O/prefill_new.log 2>&1
RC=$?; echo "rc=$RC" >> $O/prefill_new.log
if [ $RC -eq 0 ]; then
  timeout 1200 python -m pytest tests -m gpu -q > $O/suite.log 2>&1; echo "rc=$?" >> $O/suite.log
  timeout 300 python tools/prefill_attn_time.py $O/prefill_attn_time.json > $O/prefill_attn_time.log 2>&1
else
  BCG_PREFILL_ATTN_LEGACY=1 timeout 1200 python -m pytest tests -m gpu -q > $O/suite_legacy.log 2>&1; echo "rc=$?" >> $O/suite_legacy.log
fi
BCG_PREFILL_ATTN_LEGACY=1 timeout 300 python tools/prefill_attn_time.py $O/prefill_attn_time_legacy.json > $O/prefill_attn_time_legacy.log 2>&1
timeout 900 python tools/incumbent_kernels.py --out $O/incumbent_kernels.json > $O/incumbent.log 2>&1
if [ $RC -eq 0 ]; then
  timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $O/bench_s2.json 2> $O/bench_s2.err
else
  BCG_PREFILL_ATTN_LEGACY=1 timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu-baseline > $O/bench_s2_legacy.json 2> $O/bench_s2.err
fi
tail -5 $O/prefill_new.log; tail -3 $O/suite*.log; cat $O/prefill_attn_time*.log | tail -14
