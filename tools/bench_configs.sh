#!/bin/bash
# The five BASELINE.json configs through bench.py on ONE GPU (cfg5 at the per-GPU share of an 8-GPU split).
# Usage (GPU box): bash tools/bench_configs.sh <out_dir>
out=${1:-gpurun_out/configs}; mkdir -p $out
python bench.py --scene sphere_scene --width 1200 --height 800 --spp 16 > $out/cfg1_cover_1200x800_16spp.json 2> $out/cfg1.err
python bench.py --scene textured --width 1920 --height 1080 --spp 256 > $out/cfg2_textured_1080p_256spp.json 2> $out/cfg2.err
python bench.py --scene cornell_box --width 1024 --height 1024 --spp 1024 > $out/cfg3_cornell_1024_1024spp.json 2> $out/cfg3.err
python bench.py --scene cornell_box_solid --width 1024 --height 1024 --spp 1024 --no-cpu-baseline > $out/cfg3b_cornell_solid_1024_1024spp.json 2> $out/cfg3b.err
python bench.py --scene pbr_ibl --width 1920 --height 1080 --spp 1024 > $out/cfg4_pbr_ibl_1080p_1024spp.json 2> $out/cfg4.err
python bench.py --scene cover_100k_motion --width 3840 --height 2160 --spp 512 --steps 2 > $out/cfg5_cover100k_4k_512spp_share.json 2> $out/cfg5.err
for f in $out/*.json; do python - "$f" <<'PY'
import json, sys
j = json.load(open(sys.argv[1]))
cb = j.get("cpu_baseline", {})
print(f"{sys.argv[1].split('/')[-1]:48s} {j['value']:9.1f} Mpaths/s {j['mrays_s']:9.1f} Mrays/s e2e {j['e2e']['value']:9.1f} frac {j['roofline']['frac']:.3f} rays/path {j['rays_per_path']:.2f} cpu {cb.get('value', float('nan')):.2f} ({cb.get('cores')} thr)")
PY
done
