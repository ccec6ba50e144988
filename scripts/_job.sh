mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_filters.py tests/test_facade.py -m gpu -x -q 2>&1 | tail -12
timeout 300 python scripts/quick_bench_filters.py 2>&1 | tee gpurun_out/stage_filters.jsonl | tail -6
