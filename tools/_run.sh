timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/c13_pytest.log 2>&1; tail -2 gpurun_out/c13_pytest.log
for s in sphere_scene simple_light_scene cornell_box_solid cornell_box final_scene cover_100k_motion; do timeout 300 python tools/mode_sweep.py $s 1920 1080 32 3 2>&1 | tee -a gpurun_out/c13_scenes.log; done
for n in 768 960; do RT_CTA_PATHS=$n python tools/mode_sweep.py cover_100k_motion 1920 1080 16 3 2>&1 | tail -1; done
