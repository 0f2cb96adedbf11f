# Run on the GPU box with a run-time pool size variant (tools/build_variant.py rtp -DRT_CTA_POOL_PATHS=0): pool size / carve-out per scene.
for c in "cover_100k_motion 1920 1080 32" "final_scene 1024 1024 64"; do
  for pc in "1120 -1" "928 -1" "736 -1" "736 60" "544 -1" "1120 100"; do
    set -- $c; a=($pc)
    RT_CTA_PATHS=${a[0]} RT_CTA_CARVEOUT=${a[1]} RTGPU_LIB=variants/rtp/librtgpu.so timeout 120 python tools/mode_sweep.py $1 $2 $3 $4 3 2>&1 | tail -1
  done
done
