mkdir -p gpurun_out
N=${NGPU:-2}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --workload c4 --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_c4_${N}gpu.json 2> gpurun_out/bench_c4_${N}gpu.err
tail -c 1800 gpurun_out/bench_c4_${N}gpu.json; tail -3 gpurun_out/bench_c4_${N}gpu.err
