mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_facade.py tests/test_gpu_image_refinement.py tests/test_gpu_aggregation.py -m gpu -q 2>&1 | grep -E "^E  |Error|passed|failed|^FAILED" | head -40
