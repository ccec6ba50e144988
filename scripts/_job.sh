mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_image_refinement.py tests/test_gpu_patchmatch.py -m gpu -q -k "full_size" 2>&1 | grep -E "^E  |Error|passed|failed|^FAILED" | head -20
