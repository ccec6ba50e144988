set -x
mkdir -p gpurun_out/r2b
timeout 900 python -m pytest tests/test_gpu_sgm_general.py -m gpu -x -q > gpurun_out/r2b/pytest_sgm.log 2>&1
tail -15 gpurun_out/r2b/pytest_sgm.log
timeout 300 python scripts/quick_bench_sgm_general.py all > gpurun_out/r2b/sgm_general_new.jsonl 2> gpurun_out/r2b/sgm_general_new.err
cat gpurun_out/r2b/sgm_general_new.jsonl; tail -5 gpurun_out/r2b/sgm_general_new.err
