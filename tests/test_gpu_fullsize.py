"""Full-size runs of the BASELINE.json configurations on the GPU, checked through size-independent properties
(the CPU oracle needs minutes to hours at these sizes):

* the fused pipelines equal the staged API calls bit for bit;
* two independent GPU implementations agree (collapsed vs general SGM recurrence; box-sum vs direct cost volume);
* a band of rows, cut out with its halo, equals the CPU oracle run on that band (every cost-volume stage is row-local);
* the recovered disparities match the synthetic ground truth.
Operands stay on the device (torch tensors) so the runs take seconds."""
import numpy as np
import pytest

from libstevi_b200 import synth

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.fixture(scope="module")
def tctx(api):
    c = api.Context(0, torch_stream=True)
    yield c
    c.close()


@pytest.mark.parametrize("H,W,D,seed", [(1080, 1920, 256, 2), (720, 1280, 128, 1000), (720, 1280, 128, 1001)])
def test_census_sgm_full_size(api, tctx, oracle, H, W, D, seed):
    L, R, gt = synth.stereo_pair(H, W, seed, d_bg=8, d_sq=32, d_max=200 if D == 256 else 100)
    dl, dr = dev(L), dev(R)
    cv, sg, disp = api.censusSgmPipeline(dl, dr, 4, 3, D, 10, 120, 100, 8, return_volumes=True, ctx=tctx)
    # (1) staged calls give the same three arrays
    cv2 = api.unfoldBasedCostVolume(api.CENSUS, dl, dr, 4, 3, D, ctx=tctx)
    assert torch.equal(cv, cv2)
    sg2 = api.sgmCostVolume(cv2, 10, 120, Pout=100, nDirections=8, ctx=tctx)
    assert torch.equal(sg, sg2)
    idx = api.extractSelectedIndex(sg2, api.COST, ctx=tctx)
    assert torch.equal(disp, api.selectedIndexToDisp(idx, ctx=tctx))
    # (1b) the disparity-only call (aggregated costs evaluated in registers, never stored) picks the same winners
    assert torch.equal(disp, api.censusSgmPipeline(dl, dr, 4, 3, D, 10, 120, 100, 8, ctx=tctx))
    # (2) the general recurrence kernel (float volume) reproduces the collapsed evaluation
    sg3 = api.sgmCostVolume(cv2.to(torch.float32), 10, 120, Pout=100, nDirections=8, ctx=tctx)
    assert torch.equal(sg, sg3)
    del sg2, sg3
    # (3) a row band of the cost volume against the CPU oracle (rows are independent given a v_radius halo)
    i0, n = H // 2 - 8, 16
    want = oracle.port.cost_volume(oracle.CENSUS, L[i0 - 3:i0 + n + 3], R[i0 - 3:i0 + n + 3], 4, 3, D)[3:-3]
    assert np.array_equal(cv[i0:i0 + n].cpu().numpy(), want)
    # (4) disparities against the ground truth away from the borders / occlusions
    d = disp.cpu().numpy()
    inner = (slice(16, -16), slice(16, W - D))
    # (the compat-mode census is effectively 32 comparisons against the window's top-left pixel, with Q2's rounding of
    # the target words, and compat SGM adds no smoothness term: ~60 % on pure noise texture is the reference's own figure)
    assert (np.abs(d[inner] - gt[inner]) <= 1).mean() > 0.5
    # the winner of the aggregated volume is its row-wise minimum, last one on ties (Q5)
    s = sg[H // 3].cpu().numpy()
    assert np.array_equal(D - 1 - np.argmin(s[:, ::-1], axis=1), d[H // 3])


def test_zncc_refine_full_size(api, tctx, oracle):
    H, W, D, hr, vr = 1080, 1920, 192, 4, 4
    L, R, gt = synth.stereo_pair(H, W, 3, d_bg=8, d_sq=32, d_max=120, dtype=np.float32)
    dl, dr = dev(L), dev(R)
    cv, idx, tcv, refined = api.localMatchingPipeline(api.ZNCC, dl, dr, hr, vr, D, api.PARABOLA, return_all=True, ctx=tctx)
    # staged calls agree bit for bit
    cv2 = api.unfoldBasedCostVolume(api.ZNCC, dl, dr, hr, vr, D, ctx=tctx)
    assert torch.equal(cv, cv2)
    idx2 = api.extractSelectedIndex(cv2, api.SCORE, ctx=tctx)
    assert torch.equal(idx, idx2)
    tcv2 = api.truncatedCostVolume(cv2, idx2, hr, vr, 1, ctx=tctx)
    assert torch.equal(torch.nan_to_num(tcv, nan=7.0), torch.nan_to_num(tcv2, nan=7.0))
    ref2 = api.refineDispCostInterpolation(tcv2, idx2, api.PARABOLA, ctx=tctx)
    assert torch.equal(torch.nan_to_num(refined, nan=7.0), torch.nan_to_num(ref2, nan=7.0))
    del cv2, tcv2
    # normalised scores stay in [-1, 1]
    assert float(cv.abs().max()) <= 1.0 + 1e-4
    # a row band against the CPU oracle (tolerance of SURVEY.md §8(c): 1e-4 absolute)
    i0, n = 500, 6
    want = oracle.port.cost_volume(oracle.ZNCC, L[i0 - vr:i0 + n + vr], R[i0 - vr:i0 + n + vr], hr, vr, D)[vr:-vr]
    got = cv[i0:i0 + n].cpu().numpy()
    assert np.nanmax(np.abs(got - want)) <= 1e-4
    # the box-sum path against the direct O(F) path on a cropped pair (same band, fewer columns)
    sl = (slice(i0 - 20, i0 + 44), slice(300, 812))
    a = api.unfoldBasedCostVolume(api.ZNCC, dev(L[sl]), dev(R[sl]), hr, vr, 64, ctx=tctx)
    b = api.unfoldBasedCostVolume(api.ZNCC, dev(L[sl]), dev(R[sl]), hr, vr, 64, exact_sums=True, ctx=tctx)
    assert float((a - b).abs().max()) <= 1e-4
    # winners and refined disparities against the ground truth
    d, rf = idx.cpu().numpy(), refined.cpu().numpy()
    inner = (slice(16, -16), slice(16, W - D))
    assert (d[inner] == gt[inner]).mean() > 0.9
    ok = (d[inner] == gt[inner]) & np.isfinite(rf[inner])
    assert np.abs(rf[inner][ok] - gt[inner][ok]).max() <= 0.5  # the parabola never moves the winner by more than half a pixel


def test_hierarchical_full_size(api, tctx):
    H, W, D, r, depth = 3072, 4096, 512, 3, 3
    L, R, gt = synth.stereo_pair(H, W, 4, d_bg=24, d_sq=96, d_max=0, dtype=np.float32)
    tcv, disp = api.hiearchicalTruncatedCostVolume(api.ZNCC, depth, dev(L), dev(R), r, r, D, 2, ctx=tctx)
    assert tuple(tcv.shape) == (H, W, 5) and tuple(disp.shape) == (H, W)
    d = disp.cpu().numpy()
    m = 2 * sum(2 ** k * 2 for k in range(depth + 1))
    assert d.min() >= -m and d.max() <= D + m          # testCorrelationHierarchical.cpp:22-139, range criterion
    side = min(H, W) // 3
    i0, j0 = (H - side) // 2, (W - side) // 2
    inner = (slice(i0 + 32, i0 + side - 32), slice(j0 + 32, j0 + side - 32))
    assert (d[inner] == gt[inner]).mean() >= 0.95      # ... and its accuracy criterion on the interior of the square
    t = tcv.cpu().numpy()[inner]
    assert np.nanargmax(t.reshape(-1, 5), axis=1).mean() == pytest.approx(2, abs=0.2)  # the window is centred on the winner
