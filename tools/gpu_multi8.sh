#!/bin/bash
# inside `gpurun --gpus 8`: round-2 multi-GPU numbers (one process over n GPUs, one process per GPU, orbit partitions)
set -u
mkdir -p gpurun_out
python -m pytest tests/test_gpu_multi_device.py -q -x > gpurun_out/pytest_multi_device_8.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_multi_device_8.log
tail -3 gpurun_out/pytest_multi_device_8.log
python tools/multi_e2e.py grid513_4k grid65_1080p octree8_4k bunny_4k default_look_1080p > gpurun_out/multi_e2e_n8.txt 2> gpurun_out/multi_e2e_n8.err; tail -24 gpurun_out/multi_e2e_n8.txt; tail -3 gpurun_out/multi_e2e_n8.err
RT_STAGE_STORES=0 python tools/multi_e2e.py grid513_4k > gpurun_out/multi_e2e_n8_nostage.txt 2>&1; tail -6 gpurun_out/multi_e2e_n8_nostage.txt
python tools/orbit_frame_parallel.py > gpurun_out/orbit_partitions.txt 2> gpurun_out/orbit_partitions.err; tail -7 gpurun_out/orbit_partitions.txt; tail -3 gpurun_out/orbit_partitions.err
run() { n=$1; name=$2; shift 2; python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 \
   bench.py --gpus $n "$@" > gpurun_out/bench_multi${n}_$name.json 2> gpurun_out/bench_multi${n}_$name.err; }
run 8 grid513_4k --steps 50 --warmup 5
RT_STAGE_STORES=0 run 8 grid513_4k_plain --steps 50 --warmup 5
run 4 grid513_4k --steps 50 --warmup 5
RT_STAGE_STORES=2 run 4 grid513_4k_staged --steps 50 --warmup 5
run 8 grid65_1080p --scene grid65_1080p --steps 100 --warmup 5
run 8 octree8_4k --scene octree8_4k --steps 50 --warmup 5
run 8 bunny_s2_orbit64_4k --scene bunny_s2_orbit64_4k --steps 64 --warmup 5
python bench.py --impl reference --gpus 8 --steps 5 --warmup 1 > gpurun_out/bench_multi8_reference.json 2>/dev/null
python tools/bench_table.py | grep -E "file|multi[48]_"
for f in gpurun_out/bench_multi[48]_*.err; do if grep -q -i -E "error|traceback" $f; then echo "== $f"; tail -8 $f; fi; done
