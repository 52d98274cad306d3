#!/bin/bash
# usage: tools/gpu_multi2.sh N [tests]  (inside gpurun --gpus N): round-2 multi-GPU numbers
set -u
N=$1
mkdir -p gpurun_out
if [ "${2:-}" = "tests" ]; then
  python -m pytest tests/test_gpu_multi_device.py -q -x > gpurun_out/pytest_multi_device_$N.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_multi_device_$N.log
  tail -5 gpurun_out/pytest_multi_device_$N.log
fi
python tools/multi_e2e.py > gpurun_out/multi_e2e_n$N.txt 2> gpurun_out/multi_e2e_n$N.err; tail -25 gpurun_out/multi_e2e_n$N.txt; tail -5 gpurun_out/multi_e2e_n$N.err
RT_MULTI_BAND_ROWS=32 python tools/multi_e2e.py grid513_4k grid65_1080p > gpurun_out/multi_e2e_n${N}_band32.txt 2>&1; tail -9 gpurun_out/multi_e2e_n${N}_band32.txt
run() { name=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
   bench.py --gpus $N "$@" > gpurun_out/bench_multi${N}_$name.json 2> gpurun_out/bench_multi${N}_$name.err; }
run grid513_4k --steps 50 --warmup 5
RT_STAGE_STORES=2 run grid513_4k_staged --steps 50 --warmup 5
RT_STAGE_STORES=0 run grid513_4k_plain --steps 50 --warmup 5
run grid65_1080p --scene grid65_1080p --steps 100 --warmup 5
run octree8_4k --scene octree8_4k --steps 50 --warmup 5
RT_STAGE_STORES=2 run octree8_4k_staged --scene octree8_4k --steps 50 --warmup 5
run bunny_s2_orbit64_4k --scene bunny_s2_orbit64_4k --steps 64 --warmup 5
python bench.py --impl reference --gpus $N --steps 5 --warmup 1 > gpurun_out/bench_multi${N}_reference.json 2>/dev/null
python tools/bench_table.py | grep -E "file|multi${N}_"
for f in gpurun_out/bench_multi${N}_*.err; do if [ -s $f ]; then echo "== $f"; tail -5 $f; fi; done
