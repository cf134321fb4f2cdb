#!/bin/bash
# GPU call F (1 GPU): whole suite and a short bench with the round's final kernels.
O=gpurun_out/r02f; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q -rs > $O/suite.log 2>&1; echo "rc=$?" >> $O/suite.log
timeout 600 python bench.py --steps 3 --warmup 1 --no-cpu-baseline > $O/bench_s3.json 2> $O/bench_s3.err; echo "rc=$?" >> $O/bench_s3.err
tail -n 8 $O/suite.log; tail -n 3 $O/bench_s3.err; head -c 300 $O/bench_s3.json
