mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_hierarchical.py tests/test_facade.py -m gpu -q 2>&1 | grep -E "^E  |Error|passed|failed|^FAILED" | head -20
