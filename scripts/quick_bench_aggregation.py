"""Stage timing of variableCostVolumeAggregation and the 2-D search chain on full-size operands, device resident, CUDA events on
the ctx stream (torch's current stream), 512 MB L2 flush between iterations. Prints JSON lines.
  aggregation: 1080x1920, D=192 f32 volume in the reference's layout {W*D, 1, W}, 3x3 / 5x5 / 7x7 windows
  search2d:    1080x1920, 9x9 ZNCC windows, search window 5 x 33 (165 candidates): volume, winner, 3x3 neighbourhood, refinement"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from libstevi_b200 import api, synth

PEAK = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json"))).get("hbm_gbs", 6547.8)
ctx = api.Context(0, torch_stream=True)
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")


def timed(fn, n=10, warm=3):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(n):
        flush.zero_()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts))


H, W, D = 1080, 1920, 192
which = sys.argv[1] if len(sys.argv) > 1 else "all"
if which in ("all", "aggregation"):
    cv = torch.rand((H, D, W), device="cuda").permute(0, 2, 1)            # {W*D, 1, W}
    out = torch.empty((H, D, W), device="cuda").permute(0, 2, 1)
    for r in ([int(x) for x in sys.argv[2].split(',')] if len(sys.argv) > 2 else (1, 2, 3)):
        win = torch.rand((H, W, 2 * r + 1, 2 * r + 1), device="cuda") + 0.01
        ms = timed(lambda: api.variableCostVolumeAggregation(cv, win, out=out, ctx=ctx))
        alg = H * W * D * 8 + win.numel() * 4
        print(json.dumps({"stage": "variableCostVolumeAggregation", "window": f"{2*r+1}x{2*r+1}", "H": H, "W": W, "D": D, "ms": round(ms, 4),
                          "Gcells_per_s": round(H * W * D / ms / 1e6, 1), "algorithmic_GBps": round(alg / ms / 1e6, 1),
                          "frac_of_hbm_peak": round(alg / ms / 1e6 / PEAK, 3)}))
if which in ("all", "search2d"):
    L, R, _ = synth.stereo_pair(H, W, 3, d_bg=8, d_sq=16, d_max=30, dtype=np.float32)
    dl, dr = torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda()
    search = (-2, 2, 0, 32)
    n = 5 * 33
    vol = api.unfoldBased2dDisparityCostVolume(api.ZNCC, dl, dr, 4, 4, search, ctx=ctx)
    ms_cv = timed(lambda: api.unfoldBased2dDisparityCostVolume(api.ZNCC, dl, dr, 4, 4, search, ctx=ctx), n=5)
    ms_cv_exact = timed(lambda: api.unfoldBased2dDisparityCostVolume(api.ZNCC, dl, dr, 4, 4, search, exact_sums=True, ctx=ctx), n=5)
    print(json.dumps({"stage": "search2d volume, box sums (default) vs the reference's channel-order sums", "box_ms": round(ms_cv, 3),
                      "exact_sums_ms": round(ms_cv_exact, 3), "cells": H * W * n, "box_store_GBps": round(H * W * n * 4 / ms_cv / 1e6, 1),
                      "box_frac_of_hbm_peak": round(H * W * n * 4 / ms_cv / 1e6 / PEAK, 3)}))
    idx = api.extractSelected2dIndex(vol, api.SCORE, ctx=ctx)
    ms_wta = timed(lambda: api.extractSelected2dIndex(vol, api.SCORE, ctx=ctx), n=5)
    t = api.truncatedBidirectionaCostVolume(vol, idx, 1, 1, ctx=ctx)
    ms_t = timed(lambda: api.truncatedBidirectionaCostVolume(vol, idx, 1, 1, ctx=ctx))
    ms_r = timed(lambda: api.refineDisp2dCostInterpolation(t, idx, api.PARABOLA, api.ANISOTROPIC, ctx=ctx))
    ms_p = timed(lambda: api.refineDisp2dCostPatchInterpolation(t, idx, api.PARABOLA, ctx=ctx))
    print(json.dumps({"stage": "search2d ZNCC 9x9, 5x33 candidates", "H": H, "W": W, "cost_volume_ms": round(ms_cv, 3),
                      "cost_volume_Gcells_per_s": round(H * W * n / ms_cv / 1e6, 1), "cost_volume_store_GBps": round(H * W * n * 4 / ms_cv / 1e6, 1),
                      "wta_2d_ms": round(ms_wta, 3), "wta_2d_read_GBps": round(H * W * n * 4 / ms_wta / 1e6, 1),
                      "truncated_3x3_ms": round(ms_t, 4), "refine_anisotropic_ms": round(ms_r, 4), "refine_patch_ms": round(ms_p, 4)}))
