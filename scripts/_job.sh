mkdir -p gpurun_out/r2r
python bench.py --workload c5 --pairs 1024 --result-dtype u8 > gpurun_out/r2r/c5_1gpu_u8.json 2>/dev/null
python bench.py --workload c5 --pairs 1024 > gpurun_out/r2r/c5_1gpu_narrowwire.json 2>/dev/null
python bench.py --workload c5 --pairs 1024 --wide-wire > gpurun_out/r2r/c5_1gpu_widewire.json 2>/dev/null
for f in gpurun_out/r2r/c5*.json; do python -c "
import json
for ln in open('$f'):
    if ln.startswith('{'):
        b=json.loads(ln)['batch']; print('$f', round(b['pairs_per_s']), round(b['ms_per_pair_per_gpu'],3), b['pcie_GBps_per_gpu'])"; done
