timeout 900 python -m pytest tests/test_gpu_sgm_general.py -m gpu -x -q 2>&1 | tail -3
SGM_ITERS=3 timeout 300 python scripts/quick_bench_sgm_general.py all 2>&1 | python -c "import sys,json; [print(json.loads(l)['case'],json.loads(l)['ms'], json.loads(l)['kernels']) for l in sys.stdin if l.startswith('{')]"
