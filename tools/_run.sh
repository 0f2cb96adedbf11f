timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/c6_pytest.log 2>&1; tail -3 gpurun_out/c6_pytest.log
python bench.py > gpurun_out/c6_bench.json 2> gpurun_out/c6_bench.err; cat gpurun_out/c6_bench.json; tail -3 gpurun_out/c6_bench.err
python tools/mode_sweep.py sphere_scene 1920 1080 8 3 > gpurun_out/c6_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_render_cta -s 1 -c 1 -o gpurun_out/prof_cta2 python tools/mode_sweep.py sphere_scene 1920 1080 8 3 > gpurun_out/c6_ncu.log 2>&1
tail -2 gpurun_out/c6_ncu.log
