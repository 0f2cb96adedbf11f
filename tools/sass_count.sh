#!/bin/bash
# Development tool: compile the headline instantiation of the block-wavefront kernel (spheres only, no pbr.rs class)
# with extra flags and print its size: registers / spills (ptxas -v) and SASS instruction counts by opcode.
#   bash tools/sass_count.sh [-DRT_...=...]      (set F=4 etc. in the environment for another feature set)
set -e
F=${F:-0}
OUT=${OUT:-/tmp/sass_count}
mkdir -p $OUT
cd "$(dirname "$0")/../ray_tracing_in_one_weekend_b200/csrc"
nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=false -prec-div=true -prec-sqrt=true -ftz=false \
     -Xptxas -v -DRT_CTA_F=$F -DRT_CTA_CLS_MASK=0x3f -DRT_CTA_SUFFIX=_m3f "$@" -c cta_inst.cu -o $OUT/cta.o 2> $OUT/ptxas.log
grep -A2 "k_render_ctaILj${F}ELb0E" $OUT/ptxas.log | grep -E "Used|stack" | sed 's/^ptxas info *: //'
cuobjdump -sass $OUT/cta.o > $OUT/all.sass
awk -v f="k_render_ctaILj${F}ELb0E" '/Function :/ {on = index($0, f) > 0} on' $OUT/all.sass > $OUT/kernel.sass
grep -cE "^\s+/\*[0-9a-f]{4,5}\*/" $OUT/kernel.sass | sed 's/^/instructions: /'
grep -E "^\s+/\*[0-9a-f]{4,5}\*/" $OUT/kernel.sass | sed -E 's/^\s+\/\*[0-9a-f]+\*\/\s+(@!?U?P[0-9T]\s+)?//' | awk '{split($1,a,"."); c[a[1]]++} END {for (k in c) printf "%s:%d ", k, c[k]; print ""}' | tr ' ' '\n' | sort -t: -k2 -nr | head -24 | tr '\n' ' '; echo
