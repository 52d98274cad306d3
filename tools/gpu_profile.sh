#!/bin/bash
# ncu captures of the render kernel of each workload (one GPU).  Every ncu run follows the same command run plainly.
set -u
mkdir -p gpurun_out
python bench.py --steps 5 --warmup 3 > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_grid65_1080p.csv \
    python bench.py --steps 5 --warmup 3 > gpurun_out/ncu_launches.log 2>&1
SCENES=${SCENES:-"grid65_1080p grid513_4k octree6_4k octree8_4k bunny_4k bunny_s2_4k cube_512"}
for s in $SCENES; do
  python bench.py --scene $s --steps 5 --warmup 3 > gpurun_out/plain_$s.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:render -s 3 -c 1 -f -o gpurun_out/prof_$s \
      python bench.py --scene $s --steps 5 --warmup 3 > gpurun_out/ncu_full_$s.log 2>&1
  echo "$s rc=$?"
  # summarise on the box: the reports are ~11 MB each and gpurun_out/ only travels back up to 64 MiB
  python tools/ncu_summary.py report gpurun_out/prof_$s.ncu-rep > gpurun_out/r1_${s}_full.txt 2>&1
  python tools/ncu_lines.py gpurun_out/prof_$s.ncu-rep 0.5 > gpurun_out/r1_${s}_lines.txt 2>&1
  if [ "$s" != "grid65_1080p" ]; then rm -f gpurun_out/prof_$s.ncu-rep; fi
done
python tools/ncu_summary.py launches gpurun_out/launches_grid65_1080p.csv > gpurun_out/r1_grid65_1080p_launches.txt
ls -la gpurun_out
