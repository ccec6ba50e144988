"""One disparity-only census+SGM call at 1080p D=256 on device operands (profiling target: ncu -k regex:k_hamming_rows)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from libstevi_b200 import api, synth

L, R, _ = synth.stereo_pair(1080, 1920, 2, d_bg=8, d_sq=32, d_max=200)
dl, dr = torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda()
ctx = api.Context(0, torch_stream=True)
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 2):
    d = api.censusSgmPipeline(dl, dr, 4, 3, 256, 10, 120, 100, 8, ctx=ctx)
torch.cuda.synchronize()
print(int(d.sum()))
