run() { echo "== $1 $2"; RTGPU_LIB=$1 timeout 300 python tools/mode_sweep.py $2 1920 1080 32 3 2>&1 | tail -1; }
for sc in sphere_scene simple_light_scene; do
run ray_tracing_in_one_weekend_b200/librtgpu.so $sc
run variants/pu2/librtgpu.so $sc
run variants/pu7/librtgpu.so $sc
done
