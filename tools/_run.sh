timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/c31_pytest.log 2>&1; tail -2 gpurun_out/c31_pytest.log
for s in sphere_scene cover_100k_motion; do timeout 300 python tools/mode_sweep.py $s 1920 1080 32 3 2>&1 | tail -1; done
