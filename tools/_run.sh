run() { echo "== $1 $2 $3"; RTGPU_LIB=$1 RT_CTA_PATHS=$2 RT_CTA_REFILL=16 timeout 300 python tools/mode_sweep.py $3 1920 1080 32 3 2>&1 | tail -1; }
for sc in sphere_scene cover_100k_motion; do
run ray_tracing_in_one_weekend_b200/librtgpu.so 1152 $sc
run ray_tracing_in_one_weekend_b200/librtgpu.so 960 $sc
run variants/t512b2/librtgpu.so 1280 $sc
run variants/t512b2/librtgpu.so 1024 $sc
run variants/t320b3/librtgpu.so 896 $sc
run variants/t320b3/librtgpu.so 640 $sc
done
