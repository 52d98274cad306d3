#!/bin/bash
# N = 1: GPU parity tests, smoke, one bench line per workload (+ the variants the design choices rest on), reference arm.
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
python tools/pcie_probe.py > gpurun_out/pcie.json 2>&1
b() { name=$1; shift; python bench.py "$@" > gpurun_out/bench_$name.json 2> gpurun_out/bench_$name.err; }
b grid65_1080p
b grid65_1080p_fast --math fast
b grid65_1080p_persistent --sched persistent
b cube_512 --scene cube_512 --steps 100 --warmup 5
b grid65_4k --scene grid65_4k --steps 50 --warmup 5
b grid257_4k_linear --scene grid257_4k --grid-layout linear --steps 50 --warmup 5
b grid257_4k_bricks --scene grid257_4k --grid-layout bricks --steps 50 --warmup 5
b grid513_4k_linear --scene grid513_4k --grid-layout linear --steps 50 --warmup 5
b grid513_4k_bricks --scene grid513_4k --grid-layout bricks --steps 50 --warmup 5
b octree6_4k --scene octree6_4k --steps 50 --warmup 5
b octree6_4k_tiles --scene octree6_4k --steps 50 --warmup 5 --sched tiles
b octree6_4k_compact --scene octree6_4k --steps 50 --warmup 5 --sched compact
b octree8_4k --scene octree8_4k --steps 50 --warmup 5
b octree9_4k --scene octree9_4k --steps 50 --warmup 5
b bunny_4k --scene bunny_4k --steps 50 --warmup 5
b bunny_4k_persistent --scene bunny_4k --steps 50 --warmup 5 --sched persistent
b bunny_4k_lbvh_gpu --scene bunny_4k --steps 50 --warmup 5 --mesh-accel lbvh_gpu
b bunny_s2_4k --scene bunny_s2_4k --steps 50 --warmup 5
b bunny_s2_4k_lbvh_gpu --scene bunny_s2_4k --steps 50 --warmup 5 --mesh-accel lbvh_gpu
b bunny_s2_orbit64_4k --scene bunny_s2_orbit64_4k --steps 64 --warmup 5
python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err
python bench.py --impl reference --scene bunny_s2_orbit64_4k --steps 4 --warmup 1 > gpurun_out/bench_reference_orbit.json 2> gpurun_out/bench_reference_orbit.err
# host-call transfer modes side by side (copies / stores with in-out through copies / stores both ways)
E2E_SCENES=example_grid:1920x1080,example_grid:3840x2160,sdf_6:1920x1080,sdf_6:3840x2160,stanford-bunny:1920x1080,stanford-bunny:3840x2160 \
  python tools/e2e_ab.py 0,3,2 > gpurun_out/e2e_ab.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; tail -2 gpurun_out/smoke.log; python tools/bench_table.py
