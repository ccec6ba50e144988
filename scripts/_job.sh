mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_search2d.py tests/test_facade.py -m gpu -q -x 2>&1 | grep -E "^E  |Error|passed|failed|^FAILED" | head -40
timeout 300 python scripts/quick_bench_aggregation.py search2d 2>&1 | tee gpurun_out/stage_search2d.jsonl | tail -6
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
