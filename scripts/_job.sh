mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_patchmatch.py -m gpu -q 2>&1 | grep -E "^E  |Error|passed|failed|^FAILED" | head -20
timeout 600 python scripts/quick_bench_patchmatch.py 2>&1 | tee gpurun_out/stage_patchmatch.jsonl | tail -5
