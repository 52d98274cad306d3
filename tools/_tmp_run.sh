set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log
python tools/ktime.py --check grid513_4k grid257_4k grid65_4k grid65_1080p default_look_1080p
python tools/ktime.py --bands 8 grid513_4k
python tools/ktime.py --math fast --check grid513_4k grid65_1080p
