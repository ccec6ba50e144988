mkdir -p gpurun_out
N=${NGPU:-2}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_default_${N}gpu.json 2> gpurun_out/bench_default_${N}gpu.err
echo rc=$?
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_default_${N}gpu.json').read().strip().splitlines()[-1])
print('value',d['value'],'e2e',d['e2e']['value'],'n',d['n_gpus'])
for k,v in d.get('secondary',{}).items():
    print(k, v.get('value'), v.get('ms_per_step'), (v.get('result_sharded') or {}).get('ms_per_step'), (v.get('batch') or {}).get('pairs_per_s'))
PY
tail -3 gpurun_out/bench_default_${N}gpu.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus $N --steps 1 --warmup 0 2>/dev/null | tail -c 300
