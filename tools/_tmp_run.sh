set -u
mkdir -p gpurun_out
V=$PWD/triangles_sdf_cpu_raytracing_b200/variants
(
python tools/ktime.py --check grid513_4k grid65_1080p grid65_4k octree8_4k octree9_4k octree6_4k
for v in g12 g11; do RT_B200_LIB=$V/librt_b200_$v.so python tools/ktime.py --check grid513_4k grid65_1080p grid65_4k; done
RT_B200_LIB=$V/librt_b200_osm.so python tools/ktime.py --check octree8_4k octree9_4k octree6_4k
python tools/ktime.py grid513_4k octree8_4k
) > gpurun_out/r2_variants3.txt 2>&1
cat gpurun_out/r2_variants3.txt
