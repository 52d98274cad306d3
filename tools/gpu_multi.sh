#!/bin/bash
# usage: tools/gpu_multi.sh N   (inside gpurun --gpus N): bench lines of the screen-tiled path on N GPUs
set -u
N=$1
mkdir -p gpurun_out
run() { name=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
   bench.py --gpus $N "$@" > gpurun_out/multi${N}_$name.json 2> gpurun_out/multi${N}_$name.err; }
run grid65_1080p --steps 100 --warmup 5
run grid65_4k --scene grid65_4k --steps 50 --warmup 5
run grid513_4k --scene grid513_4k --steps 50 --warmup 5
run octree6_4k --scene octree6_4k --steps 50 --warmup 5
run octree8_4k --scene octree8_4k --steps 50 --warmup 5
run bunny_4k --scene bunny_4k --steps 50 --warmup 5
run bunny_s2_orbit64_4k --scene bunny_s2_orbit64_4k --steps 64 --warmup 5
if [ "$N" = "2" ]; then run grid65_1080p_nccl --steps 50 --warmup 5 --gather nccl; fi
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/multi${N}_*.json")):
    try:
        j=json.loads(open(f).read().strip().splitlines()[-1])
        p=j.get("parity",{})
        print(f.split("/")[-1], round(j["value"]), round(j["ms_per_step"],4), p.get("bit_identical"), p.get("mask_mismatch"), p.get("depth_outliers"))
    except Exception as e:
        print(f, "ERR", e, open(f.replace(".json",".err")).read()[-800:])
PY
