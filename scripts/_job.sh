timeout 900 python -m pytest tests/test_gpu_stream.py tests/test_facade.py -m gpu -x -q 2>&1 | tail -8
