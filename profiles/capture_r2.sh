#!/bin/bash
# Round-2 evidence (run under gpurun, 1 GPU): the bench lines, the reference arm, the ncu launch list of one bench step and
# full-set captures of k_automaton / k_tagmap / the CRC kernels at the benched size (configs[1], 4 GiB).  Outputs: gpurun_out/.
set -u
TAG=${1:-r2}
O=gpurun_out
mkdir -p $O
python bench.py > $O/${TAG}_bench_tar4g.json 2> $O/${TAG}_bench_tar4g.err
python bench.py --impl reference --steps 1 --warmup 1 > $O/${TAG}_bench_reference_arm.json 2> $O/${TAG}_bench_reference_arm.err
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-cli"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_launches_tar4g.csv $CMD > $O/${TAG}_ncu1.log 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_automaton -s 1 -c 1 -o $O/${TAG}_automaton_tar4g -f $CMD > $O/${TAG}_ncu2.log 2>&1
timeout 600 ncu --set full --clock-control none -k regex:'k_tagmap|k_crc|k_scan|k_emit' -c 8 -o $O/${TAG}_stream_tar4g -f $CMD > $O/${TAG}_ncu3.log 2>&1
for r in automaton stream; do
  ncu -i $O/${TAG}_${r}_tar4g.ncu-rep --page raw --csv > $O/${TAG}_${r}_tar4g_raw.csv 2>/dev/null
done
ls -la $O
