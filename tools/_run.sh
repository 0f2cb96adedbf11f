timeout 900 python -m pytest tests -m gpu -x -q -k "cornell or final or shade or furnace" > gpurun_out/c17_pytest.log 2>&1; tail -2 gpurun_out/c17_pytest.log
for s in cornell_box final_scene cornell_box_solid; do timeout 300 python tools/mode_sweep.py $s 1920 1080 32 3 2>&1 | tee -a gpurun_out/c17_scenes.log; done
