#!/bin/bash
# Round-1 ncu evidence (run under gpurun, 1 GPU): launch list of one bench step + full-set captures of the
# dominant kernel (k_automaton) and of the streaming kernels.  Outputs land in gpurun_out/.
set -u
TAG=${1:-r1}
CMD="python bench.py --size 268435456 --steps 1 --warmup 1 --no-cpu-baseline"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_automaton -s 4 -c 2 -o gpurun_out/${TAG}_automaton -f $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
ncu --set full --clock-control none -k regex:'k_scan_count|k_scan_emit|k_emit_s1|k_crc_blocks' -c 6 -o gpurun_out/${TAG}_stream -f $CMD > gpurun_out/${TAG}_ncu3.log 2>&1
ls -la gpurun_out/
