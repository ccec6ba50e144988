mkdir -p gpurun_out/r2i
timeout 900 python -m pytest tests/test_gpu_sgm_general.py -m gpu -x -q 2>&1 | tail -5
SGM_ITERS=3 timeout 300 python scripts/quick_bench_sgm_general.py all 2>&1 | python -c "import sys,json; [print(json.loads(l)['case'], json.loads(l)['ms'], json.loads(l)['kernels']) for l in sys.stdin if l.startswith('{')]"
SGM_ITERS=1 timeout 900 ncu --metrics smsp__inst_executed.sum,gpu__time_duration.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:k_sgm_sweep_t -s 6 -c 2 python scripts/quick_bench_sgm_general.py score_f32 2>&1 | grep -E "k_sgm|inst_executed|time_duration|issue_active|dram__bytes|barrier"
