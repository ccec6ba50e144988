"""Stage timing of the feature-domain refinements (image_based_refinement.h) at 1080p, F = 81 (9x9 windows), device-resident."""
import json
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from libstevi_b200 import api

H, W, F, D = 1080, 1920, 81, 192
dev = torch.device("cuda:0")
ctx = api.Context(0, torch_stream=True)
g = torch.Generator(device=dev).manual_seed(1)
fl = torch.rand((H, W, F), device=dev, generator=g) - 0.5
fr = torch.roll(fl, -7, dims=1) * 0.7 + torch.roll(fl, -8, dims=1) * 0.3
idx = torch.full((H, W), 7, dtype=torch.int32, device=dev)
idx2 = torch.zeros((H, W, 2), dtype=torch.int32, device=dev)
idx2[..., 1] = 7
cv = torch.rand((H, W, 16), device=dev, generator=g)

def timed(name, fn, bytes_):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): fn()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(json.dumps({"stage": name, "ms": round(ms, 4), "algorithmic_GB": round(bytes_ / 1e9, 3), "GBps": round(bytes_ / ms / 1e6, 1)}))

vol = H * W * F * 4
timed("refineBarycentricDisp ZNCC 1080p F=81", lambda: api.refineBarycentricDisp(api.ZNCC, fl, fr, idx, ctx=ctx), 2 * vol)
timed("refineBarycentricDisp SSD 1080p F=81", lambda: api.refineBarycentricDisp(api.SSD, fl, fr, idx, ctx=ctx), 2 * vol)
timed("refineBarycentricDisp SAD 1080p F=81", lambda: api.refineBarycentricDisp(api.SAD, fl, fr, idx, ctx=ctx), 2 * vol)
timed("refineBarycentricSymmetricDisp<1> ZNCC", lambda: api.refineBarycentricSymmetricDisp(api.ZNCC, fl, fr, idx, D, 1, ctx=ctx), 2 * vol)
timed("refineBarycentricSymmetricDisp<2> SSD", lambda: api.refineBarycentricSymmetricDisp(api.SSD, fl, fr, idx, D, 2, ctx=ctx), 2 * vol)
timed("refineBarycentric2dDisp Queen ZNCC", lambda: api.refineBarycentric2dDisp(api.ZNCC, fl, fr, idx2, (-2, 2, 0, 15), api.QUEEN, ctx=ctx), 2 * vol)
timed("refineBarycentricSymmetric2dDisp Queen SSD", lambda: api.refineBarycentric2dDisp(api.SSD, fl, fr, idx2, (-2, 2, 0, 15), api.QUEEN, symmetric=True, ctx=ctx), 2 * vol)
timed("refineCostSymmetricDisp ZNCC", lambda: api.refineCostSymmetricDisp(api.ZNCC, fl, fr, idx, cv, ctx=ctx), 2 * vol)
