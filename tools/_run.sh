for s in sphere_scene final_scene cover_100k_motion; do for f in 0 1; do RT_SCENE_FLAGS=$f timeout 300 python tools/mode_sweep.py $s 1920 1080 16 3 2>&1 | tail -1; done; done
