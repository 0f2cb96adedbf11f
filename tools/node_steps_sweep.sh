#!/bin/bash
# Run on the GPU box with a -DRT_DEV_KNOBS variant: node visits per lane and round (0 = while-while) per scene.
V=${1:-rt}
for c in "sphere_scene 1920 1080 64" "textured 1920 1080 64" "cover_4k_motion 1920 1080 64" "cover_100k_motion 1920 1080 32" "final_scene 1024 1024 64"; do
  for k in 0 1 2 4; do
    set -- $c
    RT_CTA_NODE_STEPS=$k RTGPU_LIB=variants/$V/librtgpu.so timeout 120 python tools/mode_sweep.py $1 $2 $3 $4 3 2>&1 | tail -1
  done
done
