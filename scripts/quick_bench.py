"""Scratch stage timing on the C2 configuration (device-resident operands, CUDA events on the ctx stream)."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from libstevi_b200 import api, synth

H, W, D = 1080, 1920, 256
if len(sys.argv) > 3:
    H, W, D = map(int, sys.argv[1:4])
L, R, gt = synth.stereo_pair(H, W, 2, d_bg=8, d_sq=32, d_max=200)
dl, dr = torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda()
ctx = api.Context(0, torch_stream=True)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

def timeit(fn, n=10):
    for _ in range(3): fn()
    ts = []
    for _ in range(n):
        flush.zero_()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts)), float(np.min(ts))

cells = H * W * D
cv = api.unfoldBasedCostVolume(api.CENSUS, dl, dr, 4, 3, D, ctx=ctx)
sg = api.sgmCostVolume(cv, 10, 120, ctx=ctx)
idx = api.extractSelectedIndex(sg, api.COST, ctx=ctx)
outs = (cv, sg, torch.empty((H, W), dtype=torch.int32, device="cuda"))
res = {}
res["cv"] = timeit(lambda: api.unfoldBasedCostVolume(api.CENSUS, dl, dr, 4, 3, D, ctx=ctx))
res["sgm"] = timeit(lambda: api.sgmCostVolume(cv, 10, 120, ctx=ctx))
res["wta"] = timeit(lambda: api.extractSelectedIndex(sg, api.COST, ctx=ctx))
res["pipeline_materialised"] = timeit(lambda: api.censusSgmPipeline(dl, dr, 4, 3, D, 10, 120, out=outs, ctx=ctx))
res["pipeline_disp_only"] = timeit(lambda: api.censusSgmPipeline(dl, dr, 4, 3, D, 10, 120, out=outs[2], ctx=ctx))
GB = {"cv": cells * 2, "sgm": cells * 6, "wta": cells * 4, "pipeline_materialised": cells * 12, "pipeline_disp_only": cells * 12}
for k, (med, mn) in res.items():
    print(f"{k:24s} median {med:8.3f} ms  min {mn:8.3f} ms  {cells / med / 1e6:9.1f} Gdisp/s  algorithmic {GB[k] / med / 1e6:8.1f} GB/s")
print("workspace MB", ctx.workspace_bytes / 2**20, "launches", ctx.launches)
