mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_image_refinement.py tests/test_facade.py -m gpu -q 2>&1 | grep -E "^E  |Error|passed|failed|^FAILED" | head -40
timeout 300 python scripts/quick_bench_ibr.py 2>&1 | tee gpurun_out/stage_ibr.jsonl | tail -12
