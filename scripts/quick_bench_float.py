"""Scratch stage timing of the float cost volumes on the C3 configuration (1080x1920, D=192, 9x9), device-resident
operands, per-kernel CUDA-event times from the ctx hooks."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from libstevi_b200 import api, synth

H, W, D, hr, vr = 1080, 1920, 192, 4, 4
if len(sys.argv) > 5:
    H, W, D, hr, vr = map(int, sys.argv[1:6])
names = sys.argv[6].split(",") if len(sys.argv) > 6 else ["ZNCC", "SSD", "SAD"]
L, R, gt = synth.stereo_pair(H, W, 3, d_bg=8, d_sq=32, d_max=120, dtype=np.float32)
dl, dr = torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda()
ctx = api.Context(0, torch_stream=True)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
cells = H * W * D
for name in names:
    mf = getattr(api, name)
    cv = api.unfoldBasedCostVolume(mf, dl, dr, hr, vr, D, ctx=ctx)
    for _ in range(3):
        api.unfoldBasedCostVolume(mf, dl, dr, hr, vr, D, ctx=ctx)
    ctx.set_timing(True)
    ctx.timing_report()
    tot = []
    per = {}
    for _ in range(10):
        flush.zero_()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); api.unfoldBasedCostVolume(mf, dl, dr, hr, vr, D, ctx=ctx); e1.record(); torch.cuda.synchronize()
        tot.append(e0.elapsed_time(e1))
        for k, ms in ctx.timing_report():
            per.setdefault(k, []).append(ms)
    ctx.set_timing(False)
    med = float(np.median(tot))
    print(f"{name:5s} total {med:7.3f} ms  {cells / med / 1e6:8.1f} Gdisp/s  store {cells * 4 / med / 1e6:7.1f} GB/s")
    for k, v in per.items():
        n = len(v) // 10
        print(f"      {k[:60]:60s} x{n}  {np.median(v) * n:7.3f} ms")
