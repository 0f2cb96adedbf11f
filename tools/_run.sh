timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/c40_pytest.log 2>&1; tail -3 gpurun_out/c40_pytest.log
timeout 600 python -m pytest tests/test_gpu_render.py -m gpu -q -s -k "cosine and cta" 2>&1 | grep "lum rel"
timeout 300 python tools/mode_sweep.py sphere_scene 1920 1080 32 3 2>&1 | tail -1
