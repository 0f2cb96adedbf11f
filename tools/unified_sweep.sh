#!/bin/bash
# Run on the GPU box with a -DRT_DEV_KNOBS variant (tools/build_variant.py rt): separate vs unified todo list per scene.
V=${1:-rt}
for c in "sphere_scene 1920 1080 64" "textured 1920 1080 64" "cover_defocus 1920 1080 64" "cover_4k_motion 1920 1080 64" "cover_100k_motion 1920 1080 32" "final_scene 1024 1024 64" "smoke_instances 1024 1024 64"; do
  set -- $c
  for u in "0 16" "1 8" "1 16" "1 24"; do
    set -- $c; uu=($u)
    RT_CTA_UNIFIED=${uu[0]} RT_CTA_REFILL_CAMERA=${uu[1]} RTGPU_LIB=variants/$V/librtgpu.so timeout 120 python tools/mode_sweep.py $1 $2 $3 $4 3 2>&1 | tail -1
  done
done
