for lm in 1 2 4; do for s in final_scene cornell_box pbr_ibl textured; do RT_LEAF_MAX=$lm timeout 300 python tools/mode_sweep.py $s 1920 1080 32 3 2>&1 | tee -a gpurun_out/c20_leaf.log; done; done
RT_LEAF_MAX=1 timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/c20_pytest.log 2>&1; tail -2 gpurun_out/c20_pytest.log
