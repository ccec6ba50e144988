mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_patchmatch.py tests/test_facade.py -m gpu -q 2>&1 | grep -E "^E  |Error|passed|failed|^FAILED" | head -20
timeout 300 python scripts/quick_bench_patchmatch.py 2>&1 | tee gpurun_out/stage_patchmatch.jsonl | tail -5
