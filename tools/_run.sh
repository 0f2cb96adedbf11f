timeout 900 python -m pytest tests/test_gpu_render.py -m gpu -x -q -s -k "roulette" 2>&1 | grep "rays/path\|passed\|failed\|Error\|assert" | head -20
timeout 300 python tools/mode_sweep.py sphere_scene 1920 1080 32 3 2>&1 | tail -1
