#!/bin/bash
# Round 2, N = 1: GPU parity tests, smoke, measured roofs, one bench line per workload, reference arm.
# Usage (on the box, from the repo root): bash tools/gpu_round2.sh [quick]
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
tools/peaks > gpurun_out/r2_peaks.json 2> gpurun_out/r2_peaks.err; cp gpurun_out/r2_peaks.json ${RT_PEAKS_CACHE:-/tmp/rt_b200_peaks.json}
b() { name=$1; shift; python bench.py "$@" > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err; }
b grid513_4k --steps 50 --warmup 5
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err
if [ "${1:-}" != "quick" ]; then
b grid65_1080p --scene grid65_1080p --steps 100 --warmup 5
b default_look_1080p --scene default_look_1080p --steps 50 --warmup 5
b default_look_octree_1080p --scene default_look_octree_1080p --steps 50 --warmup 5
b default_look_bunny_1080p --scene default_look_bunny_1080p --steps 50 --warmup 5
b cube_512 --scene cube_512 --steps 100 --warmup 5
b cube_512_brute --scene cube_512 --steps 100 --warmup 5 --mesh-accel brute
b grid65_4k --scene grid65_4k --steps 50 --warmup 5
b grid257_4k --scene grid257_4k --steps 50 --warmup 5
b octree6_4k --scene octree6_4k --steps 50 --warmup 5
b octree8_4k --scene octree8_4k --steps 50 --warmup 5
b octree9_4k --scene octree9_4k --steps 50 --warmup 5
b bunny_4k --scene bunny_4k --steps 50 --warmup 5
b bunny_s2_4k --scene bunny_s2_4k --steps 50 --warmup 5
b bunny_s2_orbit64_4k --scene bunny_s2_orbit64_4k --steps 64 --warmup 5
b grid513_4k_fast --math fast --steps 50 --warmup 5
fi
tail -3 gpurun_out/pytest_gpu.log; tail -2 gpurun_out/smoke.log; cat gpurun_out/r2_peaks.json; python tools/bench_table.py
