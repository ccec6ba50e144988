"""GPU parity of the 2-D search functions: unfoldBased2dDisparityCostVolume,
extractSelected2dIndex, selected2dIndexToDisp, against the CPU oracle, the golden vectors of the compiled reference and
restatements of the reference's own tests (test/unittests/testCorrelation2d.cpp).

Tolerance: the kernel sums in the reference's channel order (the `exact_sums` path of the 1-D volumes): 2e-6 of the
largest magnitude (1.0 for the normalised correlations), NaNs in the same cells. Indices and disparities are bit-exact.
"""
import os

import numpy as np
import pytest

from libstevi_b200 import synth
from test_oracle_pinning import check_2d_disparity, check_2d_matching

pytestmark = pytest.mark.gpu

MFS = ["CC", "NCC", "ZCC", "ZNCC", "SSD", "ZSSD", "SAD", "ZSAD"]
G2 = np.load(os.path.join(os.path.dirname(__file__), "golden", "reference_outputs_2d.npz"))


def close(name, got, want):
    assert got.shape == want.shape and got.dtype == np.float32
    assert np.array_equal(np.isnan(got), np.isnan(want))
    scale = 1.0 if name in ("NCC", "ZNCC") else max(1.0, float(np.nanmax(np.abs(want))))
    err = float(np.nanmax(np.abs(got - want))) / scale
    assert err <= 2e-6, (name, err)


@pytest.mark.parametrize("name", MFS)
@pytest.mark.parametrize("search", [(0, 2, 0, 4), (-3, -1, 2, 6), (-2, 2, -5, 0), (4, 4, -1, -1)])
def test_cost_volume_2d_matches_oracle(api, ctx, oracle, name, search):
    L, R = synth.random_image(21, 37, 3), synth.random_image(21, 37, 4)
    for direction in (0, 1):
        got = api.unfoldBased2dDisparityCostVolume(getattr(api, name), L, R, 2, 1, search, dDir=direction, ctx=ctx)
        close(name, np.ascontiguousarray(got), oracle.port.cost_volume_2d(getattr(oracle, name), L, R, 2, 1, search, direction))


@pytest.mark.parametrize("name", MFS)
def test_cost_volume_2d_golden(api, ctx, oracle, name):
    L, R = synth.random_image(int(G2["H"]), int(G2["W"]), int(G2["seed_l"])), synth.random_image(int(G2["H"]), int(G2["W"]), int(G2["seed_r"]))
    search = tuple(int(v) for v in G2["search"])
    for direction in (0, 1):
        got = api.unfoldBased2dDisparityCostVolume(getattr(api, name), L, R, 2, 1, search, dDir=direction, ctx=ctx)
        close(name, np.ascontiguousarray(got), G2[f"cv2d_{getattr(oracle, name)}_{direction}"])


def test_layouts_windows_and_device_operands(api, ctx, oracle):
    torch = pytest.importorskip("torch")
    L, R = synth.random_image(33, 150, 8), synth.random_image(33, 150, 9)   # more than one 128-column block
    search = (-1, 1, -6, 9)
    want = oracle.port.cost_volume_2d(oracle.ZNCC, L, R, 4, 3, search)
    ref_layout = api.unfoldBased2dDisparityCostVolume(api.ZNCC, L, R, 4, 3, search, ctx=ctx)
    assert ref_layout.strides == (150 * 16 * 3 * 4, 4, 150 * 16 * 4, 150 * 4)      # {W*dW*dH, 1, W*dW, W} elements
    c_layout = api.unfoldBased2dDisparityCostVolume(api.ZNCC, L, R, 4, 3, search, layout="c", ctx=ctx)
    close("ZNCC", np.ascontiguousarray(ref_layout), want)
    assert np.array_equal(np.ascontiguousarray(ref_layout), c_layout, equal_nan=True)
    tctx = api.Context(0, torch_stream=True)
    dev = api.unfoldBased2dDisparityCostVolume(api.ZNCC, torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda(), 4, 3, search, ctx=tctx)
    idx = api.extractSelected2dIndex(dev, api.SCORE, ctx=tctx)
    disp = api.selected2dIndexToDisp(idx, search, ctx=tctx)
    torch.cuda.synchronize()
    assert np.array_equal(dev.cpu().numpy(), np.ascontiguousarray(ref_layout), equal_nan=True)
    assert np.array_equal(disp.cpu().numpy(), oracle.port.index2d_to_disp(oracle.port.wta_2d(want, oracle.SCORE), search))
    tctx.close()
    # windows the box kernels do not cover (wider than 9, zero radius) go through the same path
    for hr, vr in ((6, 0), (0, 2)):
        got = api.unfoldBased2dDisparityCostVolume(api.SAD, L, R, hr, vr, (0, 1, -2, 2), ctx=ctx)
        close("SAD", np.ascontiguousarray(got), oracle.port.cost_volume_2d(oracle.SAD, L, R, hr, vr, (0, 1, -2, 2)))


def test_row_offset_zero_equals_the_1d_volume(api, ctx):
    """The slice at row offset 0 of a search window starting at column offset 0 is the 1-D right-to-left volume of the
    reference-order kernel, bit for bit (same kernel, same arithmetic)."""
    L, R = synth.random_image(64, 200, 21), synth.random_image(64, 200, 22)
    cv2 = api.unfoldBased2dDisparityCostVolume(api.ZNCC, L, R, 3, 3, (-2, 2, 0, 11), layout="c", ctx=ctx)
    cv1 = api.unfoldBasedCostVolume(api.ZNCC, L, R, 3, 3, 12, exact_sums=True, layout="c", ctx=ctx)
    assert np.array_equal(cv2[:, :, 2, :], cv1, equal_nan=True)


def test_selection_2d(api, ctx, oracle):
    rng = np.random.default_rng(5)
    v = rng.integers(0, 4, size=(9, 31, 3, 5)).astype(np.float32)      # many ties
    v[2, 3, 0, 0] = np.nan
    v[4, 5, 1, 2] = np.nan
    search = (-1, 1, -2, 2)
    for strategy in (0, 1):
        want = oracle.port.wta_2d(v, strategy)
        for vol in (v, np.ascontiguousarray(v.transpose(0, 2, 3, 1)).transpose(0, 3, 1, 2)):
            got = api.extractSelected2dIndex(vol, strategy, ctx=ctx)
            assert np.array_equal(got, want)
        assert np.array_equal(api.selected2dIndexToDisp(want, search, ctx=ctx), oracle.port.index2d_to_disp(want, search))
    for mf, strategy in ((oracle.ZNCC, 1), (oracle.SSD, 0)):
        gsearch = tuple(int(x) for x in G2["search"])
        idx = api.extractSelected2dIndex(G2[f"cv2d_{mf}_1"], strategy, ctx=ctx)
        assert np.array_equal(idx, G2[f"wta2d_{mf}"])
        assert np.array_equal(api.selected2dIndexToDisp(idx, gsearch, ctx=ctx), G2[f"disp2d_{mf}"])


def test_reference_shape_rules(api, ctx):
    a, b = np.zeros((5, 6), np.float32), np.zeros((5, 7), np.float32)
    with pytest.raises(api.SteviError) as e:       # cross_correlations.h:808-810: empty array
        api.unfoldBased2dDisparityCostVolume(api.ZNCC, a, b, 1, 1, (0, 1, 0, 1), ctx=ctx)
    assert e.value.status == api.E_SHAPE
    with pytest.raises(api.SteviError) as e:       # :336-338: empty search window
        api.unfoldBased2dDisparityCostVolume(api.ZNCC, a, a, 1, 1, (1, 0, 0, 1), ctx=ctx)
    assert e.value.status == api.E_SHAPE
    with pytest.raises(api.SteviError) as e:
        api.unfoldBased2dDisparityCostVolume(api.CENSUS, a, a, 1, 1, (0, 1, 0, 1), ctx=ctx)
    assert e.value.status == api.E_UNSUPPORTED


@pytest.mark.parametrize("name", ["NCC", "ZNCC", "SSD", "ZSSD", "SAD", "ZSAD"])
@pytest.mark.parametrize("hr,vr,dw,dh", [(1, 1, 5, 5), (3, 2, 4, 6)])
def test_ref_test2dMatching_and_test2dDisparity(api, ctx, oracle, name, hr, vr, dw, dh):
    """testCorrelation2d.cpp:73-194 run against the CUDA functions."""
    def cv2d(mf, l, r, hr_, vr_, search):
        return api.unfoldBased2dDisparityCostVolume(mf, l, r, hr_, vr_, search, ctx=ctx)
    mf = getattr(oracle, name)   # oracle and api share the matching-function codes
    assert mf == getattr(api, name)
    check_2d_matching(cv2d, mf, hr, vr, dw, dh, 77)
    check_2d_disparity(cv2d, lambda v, s: api.extractSelected2dIndex(v, s, ctx=ctx),
                       lambda i, s: api.selected2dIndexToDisp(i, s, ctx=ctx), mf, hr, vr, 91)
