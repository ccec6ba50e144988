"""Stage timing of PatchMatch over on-demand costs (patchmatch.h) at 1080p: ZNCC, 5x5 windows (F = 25), 1-D search of 128 and 2-D
search of 5 x 64, nIter = 5, nRandomSearch = 4 — device-resident feature volumes; the exhaustive cost volume beside it."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from libstevi_b200 import api, synth

H, W = 1080, 1920
ctx = api.Context(0, torch_stream=True)
L, R, gt = synth.stereo_pair(H, W, 3, d_bg=8, d_sq=32, d_max=60, dtype=np.float32)
dl, dr = torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda()
fl, fr = api.unfold(2, 2, dl, ctx=ctx), api.unfold(2, 2, dr, ctx=ctx)


def timed(fn, n=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts))


for name, search, n_iter in (("1-D search of 128", (0, 127), 5), ("2-D search of 5 x 64", (-2, 2, 0, 63), 5)):
    out = api.patchMatch(api.ZNCC, fr, fl, search, n_iter, 4, seed=1, ctx=ctx)
    ms = timed(lambda: api.patchMatch(api.ZNCC, fr, fl, search, n_iter, 4, seed=1, ctx=ctx))
    d = out[..., -1].cpu().numpy()
    inner = (slice(8, -8), slice(8, W - 136))
    print(json.dumps({"stage": "patchMatch ZNCC 5x5 " + name, "H": H, "W": W, "nIter": n_iter, "nRandomSearch": 4, "ms": round(ms, 2),
                      "equal_to_ground_truth": round(float((d[inner] == gt[inner]).mean()), 4)}))
ctx.set_timing(True)
api.patchMatch(api.ZNCC, fr, fl, (0, 127), 5, 4, seed=1, ctx=ctx)
per = {}
for name, ms in ctx.timing_report():
    per.setdefault(name.split("<")[0].strip("()"), []).append(ms)
ctx.set_timing(False)
print(json.dumps({"stage": "patchMatch 1-D kernels (ms, summed over the 5 iterations)", **{k: round(sum(v), 3) for k, v in per.items()}}))
ms_cv = timed(lambda: api.unfoldBasedCostVolume(api.ZNCC, dl, dr, 2, 2, 128, ctx=ctx))
cv = api.unfoldBasedCostVolume(api.ZNCC, dl, dr, 2, 2, 128, ctx=ctx)
ms_wta = timed(lambda: api.extractSelectedIndex(cv, api.SCORE, ctx=ctx))
idx = api.extractSelectedIndex(cv, api.SCORE, ctx=ctx).cpu().numpy()
print(json.dumps({"stage": "exhaustive: unfoldBasedCostVolume ZNCC 5x5 D=128 + extractSelectedIndex", "ms": round(ms_cv + ms_wta, 3),
                  "equal_to_ground_truth": round(float((idx[8:-8, 8:W - 136] == gt[8:-8, 8:W - 136]).mean()), 4)}))

# the reference's own patchMatch on the host cores of this box (test infrastructure used as the timed CPU baseline, like bench.py's)
try:
    import time
    import oracle as O
    if O.ref_ibr is not None and O.ref_ibr.has("patch_match"):
        fs, ft = O.port.unfold(R, 2, 2), O.port.unfold(L, 2, 2)
        t0 = time.time()
        out = O.ref_ibr.patch_match(O.ZNCC, fs, ft, (0, 127), 5, 4)
        dt = time.time() - t0
        print(json.dumps({"stage": "reference CPU patchMatch ZNCC 5x5 1-D search of 128 (oracle/_ref, all host cores)", "ms": round(dt * 1e3, 1),
                          "cores": os.cpu_count(), "equal_to_ground_truth": round(float((out[8:-8, 8:W - 136, 0] == gt[8:-8, 8:W - 136]).mean()), 4)}))
except Exception as e:  # the checker is optional here
    print(json.dumps({"stage": "reference CPU patchMatch", "skipped": str(e)}))
