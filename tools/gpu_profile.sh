#!/bin/bash
# ncu captures of the render kernel of each workload (one GPU).  Every ncu run follows the same command run plainly.
# Summaries land in gpurun_out/r2_<scene>_{full,lines}.txt (copy them to profiles/), the launch list of the DEFAULT
# bench command in gpurun_out/r2_default_launches.txt.
set -u
mkdir -p gpurun_out
R=${ROUND:-r2}
python bench.py --steps 5 --warmup 3 > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_default.csv \
    python bench.py --steps 5 --warmup 3 > gpurun_out/ncu_launches.log 2>&1
python tools/ncu_summary.py launches gpurun_out/launches_default.csv > gpurun_out/${R}_default_launches.txt
SCENES=${SCENES:-"grid513_4k octree8_4k bunny_4k default_look_1080p grid65_1080p"}
for s in $SCENES; do
  python tools/ktime.py --reps 5 $s > gpurun_out/plain_$s.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:render -s 3 -c 1 -f -o gpurun_out/prof_$s \
      python tools/ktime.py --reps 5 $s > gpurun_out/ncu_full_$s.log 2>&1
  echo "$s rc=$?"
  # summarise on the box: the reports are ~11 MB each and gpurun_out/ only travels back up to 64 MiB
  python tools/ncu_summary.py report gpurun_out/prof_$s.ncu-rep > gpurun_out/${R}_${s}_full.txt 2>&1
  python tools/ncu_lines.py gpurun_out/prof_$s.ncu-rep 0.5 > gpurun_out/${R}_${s}_lines.txt 2>&1
  if [ "$s" != "grid513_4k" ]; then rm -f gpurun_out/prof_$s.ncu-rep; fi
done
ls -la gpurun_out | head -40
