"""A/B proxy for the tensor-core question of the ZNCC cost volume (north_star: "a tensor-core banded-tile contraction only where ncu
shows it beats the CUDA-core path"). The banded contraction of C3 (1080p, 9x9 => F = 81, D = 192) as the cuBLAS batched GEMM it
would be — per image row and 128-column tile: [128 x 81] . [81 x 320] (the 128 + 192 target columns the tile's disparities reach) —
timed alone in TF32 and in FP32, i.e. WITHOUT building the two 81-channel feature volumes it reads (2 x 0.67 GB written and read)
and WITHOUT extracting the band into the [H][D][W] volume (1.6 GB written). Printed beside the box-sum kernel's measured time."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from libstevi_b200 import api, synth

H, W, D, F = 1080, 1920, 192, 81
tiles = W // 128
A = torch.randn(H * tiles, 128, F, device="cuda")
B = torch.randn(H * tiles, F, 128 + D, device="cuda")


def timed(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts))


out = torch.empty(H * tiles, 128, 128 + D, device="cuda")
res = {}
for name, tf32 in (("tf32", True), ("fp32", False)):
    torch.backends.cuda.matmul.allow_tf32 = tf32
    res[name] = timed(lambda: torch.bmm(A, B, out=out))
flop = 2.0 * H * tiles * 128 * (128 + D) * F
L, R, _ = synth.stereo_pair(H, W, 3, d_bg=8, d_sq=16, d_max=30, dtype=np.float32)
dl, dr = torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda()
ctx = api.Context(0, torch_stream=True)
cv = api.unfoldBasedCostVolume(api.ZNCC, dl, dr, 4, 4, D, ctx=ctx)
box = timed(lambda: api.unfoldBasedCostVolume(api.ZNCC, dl, dr, 4, 4, D, ctx=ctx))
print(json.dumps({"stage": "C3 ZNCC 9x9 cost volume: banded contraction as cuBLAS batched GEMM (proxy) vs the box-sum path",
                  "gemm_shape": [H * tiles, 128, 128 + D, F], "gemm_gflop": round(flop / 1e9, 1),
                  "bmm_tf32_ms": round(res["tf32"], 3), "bmm_fp32_ms": round(res["fp32"], 3),
                  "bmm_tf32_TFLOPs": round(flop / res["tf32"] / 1e9, 1),
                  "useful_fraction_of_the_products": round(D / (128 + D) * 1.0, 2),
                  "not_included_in_the_gemm_time": "two 81-channel feature volumes (2 x 0.67 GB written + read), band extraction into [H][D][W] (1.6 GB)",
                  "box_sum_path_ms_incl_padding_and_statistics": round(box, 3)}))
