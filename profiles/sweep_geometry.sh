#!/bin/bash
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5) > gpurun_out/s3_tests.log
for gt in "33 32" "65 32" "129 32" "17 64" "33 64" "65 64" "17 128" "33 128" "9 256"; do set -- $gt
  echo "G=$1 T=$2" >> gpurun_out/s3_sweep.log
  MRZ_PROFILE=1 MRZ_G=$1 MRZ_T=$2 timeout 300 python bench.py --size 268435456 --steps 1 --warmup 1 --no-cpu-baseline 2>> gpurun_out/s3_sweep.err | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['roofline']['ms_per_step'])" >> gpurun_out/s3_sweep.log 2>&1
done
MRZ_PROFILE=1 timeout 300 python bench.py --size 1073741824 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/s3_b1g.json 2> gpurun_out/s3_b1g.err
cat gpurun_out/s3_tests.log gpurun_out/s3_sweep.log; tail -2 gpurun_out/s3_b1g.err; cat gpurun_out/s3_b1g.json
