timeout 600 python -m pytest tests/test_gpu_render.py -m gpu -x -q -k "cta_wavefront or agree" > gpurun_out/c7_pytest.log 2>&1; tail -3 gpurun_out/c7_pytest.log
for s in sphere_scene cornell_box final_scene pbr_ibl cover_100k_motion; do timeout 300 python tools/mode_sweep.py $s 1920 1080 32 3 2>&1 | tee -a gpurun_out/c7_scenes.log; done
