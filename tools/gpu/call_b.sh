#!/bin/bash
# GPU call B (1 GPU): prefill attention v2 (BN 64 / 128), engine host-path changes, suite, ncu pages, bench.
O=gpurun_out/r02b; mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_kernels.py -q -k "prefill_attn" > $O/prefill_bn64.log 2>&1; RC64=$?; echo "rc=$RC64" >> $O/prefill_bn64.log
BCG_PREFILL_ATTN_BN=128 timeout 300 python -m pytest tests/test_gpu_kernels.py -q -k "prefill_attn" > $O/prefill_bn128.log 2>&1; RC128=$?; echo "rc=$RC128" >> $O/prefill_bn128.log
[ $RC64 -eq 0 ] && timeout 300 python tools/prefill_attn_time.py $O/prefill_attn_time_bn64.json > $O/prefill_attn_time_bn64.log 2>&1
[ $RC128 -eq 0 ] && BCG_PREFILL_ATTN_BN=128 timeout 300 python tools/prefill_attn_time.py $O/prefill_attn_time_bn128.json > $O/prefill_attn_time_bn128.log 2>&1
if [ $RC64 -eq 0 ]; then ENVX=""; elif [ $RC128 -eq 0 ]; then ENVX="BCG_PREFILL_ATTN_BN=128"; else ENVX="BCG_PREFILL_ATTN_LEGACY=1"; fi
echo "suite env: $ENVX" > $O/suite.log
env $ENVX timeout 1500 python -m pytest tests -m gpu -q >> $O/suite.log 2>&1; echo "rc=$?" >> $O/suite.log
env $ENVX timeout 600 python bench.py --steps 3 --warmup 1 --no-cpu-baseline > $O/bench_s3.json 2> $O/bench_s3.err
if [ $RC64 -eq 0 ] || [ $RC128 -eq 0 ]; then
  env $ENVX timeout 200 python tools/prefill_attn_time.py > $O/ncu_plain1.log 2>&1 && \
  env $ENVX timeout 600 ncu --set full --clock-control none --import-source on -k regex:prefill_attn -s 3 -c 1 -o $O/prof_prefill_attn python tools/prefill_attn_time.py > $O/ncu1.log 2>&1
  env $ENVX timeout 200 python tools/ncu_targets.py > $O/ncu_plain2.log 2>&1 && \
  env $ENVX timeout 900 ncu --set full --clock-control none --import-source on -k "regex:mask_sample|add_rmsnorm_row|qknorm_rope" -c 4 -o $O/prof_glue python tools/ncu_targets.py > $O/ncu2.log 2>&1
fi
tail -3 $O/prefill_bn64.log $O/prefill_bn128.log $O/suite.log; cat $O/prefill_attn_time_bn*.log | tail -14
