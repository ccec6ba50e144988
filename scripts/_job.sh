mkdir -p gpurun_out/r2n
timeout 600 python -m pytest tests/test_gpu_stream.py -m gpu -x -q 2>&1 | tail -5
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r2n/bench_c2.json 2> gpurun_out/r2n/bench_c2.err; tail -3 gpurun_out/r2n/bench_c2.err; python -c "
import json; l=json.load(open('gpurun_out/r2n/bench_c2.json'))
print({k:l[k] for k in ('value','ms_per_step','e2e','sustained','roofline','cpu_baseline','gpu_launches')})
for k,v in l['secondary'].items(): print(k, json.dumps(v)[:900])"
