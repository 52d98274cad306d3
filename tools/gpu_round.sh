#!/bin/bash
# One gpurun pass: GPU parity tests, bench lines of every workload, the reference arm, the ncu launch list and one
# `ncu --set full` capture of the headline kernel.  Outputs under gpurun_out/ (summarised into profiles/ afterwards).
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
nproc > gpurun_out/nproc.txt
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
python bench.py > gpurun_out/bench_grid65_1080p.json 2> gpurun_out/bench_grid65_1080p.err
for s in cube_512 octree6_4k bunny_4k grid65_4k; do
  python bench.py --scene $s --steps 50 --warmup 5 > gpurun_out/bench_$s.json 2> gpurun_out/bench_$s.err
done
python bench.py --math fast > gpurun_out/bench_grid65_1080p_fast.json 2> gpurun_out/bench_fast.err
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err
# ncu: launch list of the bench command, then the full capture of the render kernel
python bench.py --steps 5 --warmup 3 > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_grid65_1080p.csv \
    python bench.py --steps 5 --warmup 3 > gpurun_out/ncu_launches.log 2>&1
python bench.py --steps 5 --warmup 3 > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:render -s 3 -c 2 -o gpurun_out/prof_grid65_1080p \
    python bench.py --steps 5 --warmup 3 > gpurun_out/ncu_full.log 2>&1
for s in octree6_4k bunny_4k; do
  python bench.py --scene $s --steps 5 --warmup 3 > gpurun_out/plain_$s.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:render -s 3 -c 1 -o gpurun_out/prof_$s \
      python bench.py --scene $s --steps 5 --warmup 3 > gpurun_out/ncu_full_$s.log 2>&1
done
ls -la gpurun_out
tail -3 gpurun_out/pytest_gpu.log; tail -2 gpurun_out/smoke.log; cat gpurun_out/bench_grid65_1080p.json
