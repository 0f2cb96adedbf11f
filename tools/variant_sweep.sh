#!/bin/bash
# Run on the GPU box: time the default renderer of every variants/<name>/librtgpu.so on a few scenes.
# Usage: bash tools/variant_sweep.sh "<variant names>" ["scene w h spp;..."]
mkdir -p gpurun_out
CASES=${2:-"sphere_scene 1920 1080 64;cornell_box 1024 1024 64;cornell_box_solid 1024 1024 64;pbr_ibl 1920 1080 64"}
for v in $1; do
  IFS=';' read -ra CS <<< "$CASES"
  for c in "${CS[@]}"; do
    set -- $c
    echo -n "[$v] "
    RTGPU_LIB=variants/$v/librtgpu.so timeout 120 python tools/mode_sweep.py $1 $2 $3 $4 3 2>&1 | tail -1
  done
done
