"""Stage timing of the general SGM recurrence (sgm.h:186-404 as a real recurrence: Score strategy, float volumes, CORRECTED
semantics) on full-size operands, device resident, CUDA events on the ctx stream, 512 MB L2 flush between iterations.
  score_f32:      1080x1920, D=256 f32 volume of ZNCC-like scores in the reference's layout {W*D, 1, W}, sgmCostVolume<8, Score>
                  (the reference's own benchmark case, benchmarkCrossCorrelationAlgorithms.cpp:282-303)
  corrected_u16:  the census cost volume of C2 (u16), sgmCostVolume<8, Cost> with STEVI_CORRECTED (textbook recurrence)
  cost_f32:       float costs, COMPAT Cost mode (Q3 evaluated per path)
Prints one JSON line per case with the per-kernel breakdown of the ctx's event hooks."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from libstevi_b200 import api

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
try:
    PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 6547.8)
except Exception:
    PEAK = 6547.8
ctx = api.Context(0, torch_stream=True)
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")


def timed(fn, n=5, warm=2):
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
    ctx.set_timing(True)
    ctx.timing_report()
    fn()
    rep = ctx.timing_report()
    ctx.set_timing(False)
    return [[name.strip("()"), round(ms, 4)] for name, ms in rep]   # in launch order


H, W, D = (int(x) for x in os.environ.get("SGM_SHAPE", "1080,1920,256").split(","))
which = sys.argv[1] if len(sys.argv) > 1 else "all"
ndir = int(os.environ.get("SGM_NDIR", "8"))
n_iter = int(os.environ.get("SGM_ITERS", "5"))
cases = []
if which in ("all", "score_f32"):
    g = torch.Generator(device="cuda"); g.manual_seed(1)
    v = (torch.rand((H, D, W), device="cuda", generator=g) * 2 - 1).permute(0, 2, 1)   # {W*D, 1, W}
    cases.append(("score_f32", v, dict(P1=0.001, P2=0.01, Pout=100.0, strategy=api.SCORE, semantics=api.COMPAT), 4))
if which in ("all", "corrected_u16"):
    g = torch.Generator(device="cuda"); g.manual_seed(2)
    v = torch.randint(0, 33, (H, D, W), device="cuda", generator=g, dtype=torch.int32).to(torch.uint16).permute(0, 2, 1)
    cases.append(("corrected_u16", v, dict(P1=10, P2=120, Pout=100.0, strategy=api.COST, semantics=api.CORRECTED), 2))
if which in ("all", "cost_f32"):
    g = torch.Generator(device="cuda"); g.manual_seed(3)
    v = (torch.rand((H, D, W), device="cuda", generator=g) * 10).permute(0, 2, 1)
    cases.append(("cost_f32", v, dict(P1=0.5, P2=2.0, Pout=100.0, strategy=api.COST, semantics=api.COMPAT), 4))
for name, v, kw, isz in cases:
    fn = lambda: api.sgmCostVolume(v, kw["P1"], kw["P2"], Pout=kw["Pout"], nDirections=ndir, strategy=kw["strategy"],
                                   semantics=kw["semantics"], ctx=ctx)
    ms = timed(fn, n=n_iter)
    alg = H * W * D * (isz + 4)
    print(json.dumps({"stage": "sgmCostVolume general recurrence", "case": name, "H": H, "W": W, "D": D, "directions": ndir,
                      "ms": round(ms, 3), "Gcells_per_s": round(H * W * D / ms / 1e6, 1), "algorithmic_bytes": alg,
                      "algorithmic_GBps": round(alg / ms / 1e6, 1), "frac_of_hbm_peak": round(alg / ms / 1e6 / PEAK, 4),
                      "workspace_MB": round(ctx.workspace_bytes / 1e6, 1), "kernels": breakdown(fn)}), flush=True)
    del v
    torch.cuda.empty_cache()
