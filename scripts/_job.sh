timeout 1500 python -m pytest tests/test_gpu_configs.py -m gpu -x -q --durations=3 -k c2 2>&1 | tail -8
