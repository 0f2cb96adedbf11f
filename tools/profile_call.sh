#!/bin/bash
# Run on the GPU box (gpurun): ncu --set full of ONE launch of the headline kernel and of the Cornell-box kernel,
# plus the launch list of a short bench run.  Usage: bash tools/profile_call.sh <tag>
tag=${1:-r02}
mkdir -p gpurun_out
B="--no-configs --no-strong --no-cpu-baseline --e2e-steps 1"
timeout 120 python bench.py --steps 2 --warmup 3 $B > gpurun_out/${tag}_plain.json 2> gpurun_out/${tag}_plain.err || exit 1
timeout 170 ncu --set full --clock-control none --import-source on -k regex:k_render_cta -s 4 -c 1 -f -o gpurun_out/${tag}_headline \
    python bench.py --steps 2 --warmup 3 $B > gpurun_out/${tag}_ncu_headline.log 2>&1
timeout 170 ncu --set full --clock-control none --import-source on -k regex:k_render_cta -s 4 -c 1 -f -o gpurun_out/${tag}_cornell \
    python bench.py --scene cornell_box --width 1024 --height 1024 --spp 64 --steps 2 --warmup 3 $B > gpurun_out/${tag}_ncu_cornell.log 2>&1
timeout 120 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 3 $B > gpurun_out/${tag}_ncu_launches.log 2>&1
ls -la gpurun_out/${tag}_*
