python bench.py > gpurun_out/c22_bench.json 2> gpurun_out/c22_bench.err; cut -c1-300 gpurun_out/c22_bench.json; tail -3 gpurun_out/c22_bench.err
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/c22_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01_cta_v4_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/c22_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_render_cta -s 2 -c 1 -o gpurun_out/prof_cta4_bench python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/c22_ncu2.log 2>&1
tail -2 gpurun_out/c22_ncu2.log
