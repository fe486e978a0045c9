#!/bin/bash
# plan-offload sweep: automaton time vs number of helper CTAs (MRZ_HELPERS)
(timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3)
for h in 0 4 8 16 32; do for sz in 1073741824 4294967296; do
  echo "helpers=$h size=$sz"
  MRZ_HELPERS=$h MRZ_PROFILE=1 timeout 200 python bench.py --size $sz --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/s14_${h}_$sz.json 2> gpurun_out/s14_${h}_$sz.err
  tail -2 gpurun_out/s14_${h}_$sz.err | head -1 | cut -c1-330
  python -c "import json;d=json.load(open('gpurun_out/s14_${h}_$sz.json'));print(d['value'],d['roofline']['ms_per_step'])"
done; done
