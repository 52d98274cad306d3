set -u
mkdir -p gpurun_out
HEAD_LIB=$PWD/triangles_sdf_cpu_raytracing_b200/variants/librt_b200_head.so
(
echo "# same box: committed kernels (librt_b200_head.so, no heavy-first code) vs the working tree with its defaults"
for i in 1 2; do
RT_B200_LIB=$HEAD_LIB python tools/ktime.py grid513_4k grid65_1080p bunny_4k grid65_4k default_look_1080p
python tools/ktime.py grid513_4k grid65_1080p bunny_4k grid65_4k default_look_1080p
done
RT_B200_LIB=$HEAD_LIB python tools/ktime.py --bands 8 grid513_4k bunny_4k bunny_s2_4k
python tools/ktime.py --bands 8 grid513_4k bunny_4k bunny_s2_4k
RT_B200_LIB=$HEAD_LIB python tools/ktime.py --bands 4 grid513_4k bunny_4k
python tools/ktime.py --bands 4 grid513_4k bunny_4k
RT_HEAVY_CYCLES=35000 python tools/ktime.py grid65_1080p default_look_1080p
RT_HEAVY_CYCLES=35000 python tools/ktime.py --bands 8 grid513_4k bunny_4k
) > gpurun_out/r2_heavy_ab.txt 2>&1
cat gpurun_out/r2_heavy_ab.txt
