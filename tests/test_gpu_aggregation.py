"""GPU parity of variableCostVolumeAggregation (guided_cost_filtering.h:102-176) against the CPU oracle
(bit-identical to the compiled reference) and the golden vector. Float sums in the reference's order with separately rounded
multiplies and adds: the comparison is bit-exact."""
import os

import numpy as np
import pytest

from libstevi_b200 import synth
from test_oracle_pinning import bilateral_windows

pytestmark = pytest.mark.gpu

G2 = np.load(os.path.join(os.path.dirname(__file__), "golden", "reference_outputs_2d.npz"))


@pytest.mark.parametrize("H,W,D,vr,hr", [(23, 150, 9, 1, 1), (40, 300, 33, 2, 2), (17, 129, 5, 3, 3), (12, 140, 7, 1, 3), (9, 131, 4, 4, 0),
                                         (5, 4, 3, 3, 3), (1, 1, 2, 2, 2), (11, 512, 6, 2, 2), (9, 516, 35, 1, 1), (8, 20, 3, 2, 2),
                                         (7, 1030, 4, 2, 2), (21, 203, 19, 3, 3), (33, 260, 21, 3, 3), (19, 77, 6, 2, 2), (9, 64, 40, 3, 3)])
def test_aggregation_matches_oracle(api, ctx, oracle, H, W, D, vr, hr):
    rng = np.random.default_rng(H * W)
    cv = rng.uniform(-2, 2, (H, W, D)).astype(np.float32)
    win = rng.uniform(0.01, 1, (H, W, 2 * vr + 1, 2 * hr + 1)).astype(np.float32)
    want = oracle.port.variable_cost_volume_aggregation(cv, win)
    assert np.array_equal(api.variableCostVolumeAggregation(cv, win, ctx=ctx), want)
    ref_layout = np.ascontiguousarray(cv.transpose(0, 2, 1)).transpose(0, 2, 1)          # {W*D, 1, W}
    got = api.variableCostVolumeAggregation(ref_layout, win, layout="ref", ctx=ctx)
    assert got.transpose(0, 2, 1).flags.c_contiguous and np.array_equal(got, want)


@pytest.mark.parametrize("r", [1, 2, 3])
def test_aggregation_on_column_sliced_views(api, ctx, oracle, r):
    """A column-sliced view of a volume in the reference's layout (pitch wider than the view, width not a multiple of 4): the
    16-byte paths must not be taken for it, and every column is written."""
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(70 + r)
    H, Wfull, W, D = 14, 48, 45, 10
    full = rng.uniform(-1, 1, (H, D, Wfull)).astype(np.float32)
    win = rng.uniform(0.01, 1, (H, W, 2 * r + 1, 2 * r + 1)).astype(np.float32)
    tctx = api.Context(0, torch_stream=True)
    view = torch.from_numpy(full).cuda().permute(0, 2, 1)[:, 1:1 + W]          # strides {Wfull*D, 1, Wfull}, first column off alignment
    out = torch.full((H, D, Wfull), float("nan"), device="cuda").permute(0, 2, 1)[:, 1:1 + W]
    api.variableCostVolumeAggregation(view, torch.from_numpy(win).cuda(), out=out, ctx=tctx)
    torch.cuda.synchronize()
    want = oracle.port.variable_cost_volume_aggregation(np.ascontiguousarray(full.transpose(0, 2, 1)[:, 1:1 + W]), win)
    assert np.array_equal(out.cpu().numpy(), want)
    tctx.close()


def test_aggregation_golden_and_pipeline(api, ctx, oracle):
    H, W = int(G2["H"]), int(G2["W"])
    cv = np.ascontiguousarray(G2[f"cv2d_{oracle.SSD}_1"][:, :, :, 2])
    win = bilateral_windows(synth.random_image(H, W, int(G2["seed_r"])), 2, 2)
    assert np.array_equal(api.variableCostVolumeAggregation(cv, win, ctx=ctx), G2["aggregated_ssd_5x5"])
    # adjacent to the cost volume on the device: volume -> aggregation -> winner, nothing leaves the GPU in between
    torch = pytest.importorskip("torch")
    L, R = synth.random_image(90, 260, 61), synth.random_image(90, 260, 62)
    wins = bilateral_windows(R, 2, 2)
    tctx = api.Context(0, torch_stream=True)
    vol = api.unfoldBasedCostVolume(api.ZSSD, torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda(), 2, 2, 24, ctx=tctx)
    agg = api.variableCostVolumeAggregation(vol, torch.from_numpy(wins).cuda(), layout="ref", ctx=tctx)
    idx = api.extractSelectedIndex(agg, api.COST, ctx=tctx)
    torch.cuda.synchronize()
    want = oracle.port.variable_cost_volume_aggregation(np.ascontiguousarray(vol.cpu().numpy()), wins)
    assert np.array_equal(agg.cpu().numpy(), want)
    assert np.array_equal(idx.cpu().numpy(), oracle.port.wta(want, oracle.COST))
    tctx.close()


def test_aggregation_shape_rules(api, ctx):
    cv = np.zeros((9, 12, 5), np.float32)
    for bad in ((8, 12, 3, 3), (9, 11, 3, 3), (9, 12, 2, 3), (9, 12, 3, 4)):        # guided_cost_filtering.h:108-130: empty array
        assert api.variableCostVolumeAggregation(cv, np.ones(bad, np.float32), ctx=ctx) is None
    with pytest.raises(api.SteviError):
        api.variableCostVolumeAggregation(cv, np.ones((9, 12, 3, 3), np.float32), out=cv, ctx=ctx)
