timeout 900 python -m pytest tests/test_gpu_sgm_general.py -m gpu -x -q 2>&1 | tail -15
