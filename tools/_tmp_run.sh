set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log
python tools/ktime.py --check grid65_1080p grid513_4k grid65_4k bunny_4k cube_512 octree8_4k octree6_4k octree9_4k default_look_1080p default_look_octree_1080p 2>&1 | tee gpurun_out/ktime_after.txt
