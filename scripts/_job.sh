mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-secondary --no-cpu-baseline"
$CMD > gpurun_out/b_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_r2_c2.csv $CMD > gpurun_out/ncu_launch.log 2>&1
$CMD > gpurun_out/b_plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"k_sgm_apply_rows|k_hamming_rows" -s 4 -c 4 -o gpurun_out/c2_full -f $CMD > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log; ls -la gpurun_out/*.ncu-rep gpurun_out/launches_r2_c2.csv
