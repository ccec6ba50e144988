mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_aggregation.py -m gpu -q -x 2>&1 | grep -E "^E  |Error|passed|failed|^FAILED" | head -20
timeout 300 python scripts/quick_bench_aggregation.py aggregation 2>&1 | tee gpurun_out/stage_aggregation.jsonl | tail -6
