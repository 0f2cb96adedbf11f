timeout 900 python -m pytest tests/test_gpu_bvh_build.py -m gpu -x -q -s 2>&1 | grep -v "^$" | tail -25
