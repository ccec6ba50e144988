"""Stage timing of meanFilter2D / sigmaFilter on a device-resident 1920x1080 f32 image (CUDA events on the ctx stream, 512 MB
L2 flush between iterations). One JSON line per case with the per-kernel breakdown."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from libstevi_b200 import api

ctx = api.Context(0, torch_stream=True)
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
H, W = 1080, 1920
img = torch.rand((H, W), device="cuda") * 2 - 1


def timed(fn, n=7, warm=3):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(n):
        flush.zero_()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts))


def breakdown(fn):
    ctx.set_timing(True); ctx.timing_report(); fn(); rep = ctx.timing_report(); ctx.set_timing(False)
    return [[n.strip("()"), round(ms, 4)] for n, ms in rep]


for hr, vr in ((2, 2), (4, 4)):
    mean = api.meanFilter2D(hr, vr, img, ctx=ctx)
    for name, fn, alg in (("meanFilter2D", lambda: api.meanFilter2D(hr, vr, img, ctx=ctx), H * W * 8),
                          ("sigmaFilter", lambda: api.sigmaFilter(hr, vr, mean, img, ctx=ctx), H * W * 12)):
        ms = timed(fn)
        print(json.dumps({"stage": name, "H": H, "W": W, "window": f"{2*hr+1}x{2*vr+1}", "ms": round(ms, 4),
                          "algorithmic_bytes": alg, "algorithmic_GBps": round(alg / ms / 1e6, 1), "kernels": breakdown(fn)}), flush=True)
