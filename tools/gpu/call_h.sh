#!/bin/bash
# GPU call H (1 GPU): round-2 evidence with the final kernels: default bench, reference arm, incumbents, ncu launch list and --set full pages.
O=gpurun_out/r02h; mkdir -p $O
timeout 900 python bench.py > $O/bench_default.json 2> $O/bench_default.err; echo "rc=$?" >> $O/bench_default.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_reference.json 2> $O/bench_reference.err
timeout 600 python tools/incumbent_kernels.py --out $O/incumbent_kernels.json > $O/incumbent.log 2>&1
timeout 200 python tools/prefill_attn_time.py $O/prefill_attn_time.json > $O/prefill_attn_time.log 2>&1
timeout 200 python tools/prefill_gemm_time.py 12390 > $O/prefill_gemm_time.log 2>&1
timeout 300 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --fixed-rounds 0 > $O/bench_plain_for_ncu.json 2> $O/plain.err && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $O/ncu_launches_bench_s1w1.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline --fixed-rounds 0 > $O/ncu_launches.log 2>&1
timeout 200 python tools/ncu_targets.py > $O/ncu_plain2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:gemm2sm|qknorm_rope_kvwrite_token|mask_sample" -c 12 -o $O/prof_final python tools/ncu_targets.py > $O/ncu2.log 2>&1
tail -n 2 $O/bench_default.err; head -c 400 $O/bench_default.json; echo; cat $O/prefill_gemm_time.log; tail -n 4 $O/ncu2.log
