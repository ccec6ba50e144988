mkdir -p gpurun_out/r2m
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python scripts/quick_bench_sgm_general.py all > gpurun_out/r2m/stage_sgm.jsonl 2>gpurun_out/r2m/stage_sgm.err; cat gpurun_out/r2m/stage_sgm.jsonl
SGM_ITERS=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_sgm_sweep -s 9 -c 3 -o gpurun_out/r2m/ncu_sgm_sweep_final python scripts/quick_bench_sgm_general.py score_f32 > gpurun_out/r2m/ncu.log 2>&1
SGM_ITERS=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r2m/launches_sgm_general.csv python scripts/quick_bench_sgm_general.py score_f32 > /dev/null 2>&1
