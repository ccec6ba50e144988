"""GPU parity of winner-neighbourhood extraction and cost-based subpixel refinement:
selectedCost, truncatedCostVolume, extractInBoundDomain, refineDispCostInterpolation and the fused refinement.
Gathers are bit-exact; the refined disparity is compared to 1e-5 px (logf in the Gaussian kernel: 1e-4)."""
import numpy as np
import pytest

from libstevi_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def volume(oracle):
    L, R, _ = synth.stereo_pair(37, 61, 5, d_bg=3, d_sq=8, dtype=np.float32)
    cv = oracle.port.cost_volume(oracle.ZNCC, L, R, 2, 2, 14)
    idx = oracle.port.wta(cv, oracle.SCORE)
    return cv, idx


def same(a, b):
    return a.shape == b.shape and np.array_equal(a, b, equal_nan=True)


def test_selected_cost(api, ctx, oracle, volume):
    cv, idx = volume
    assert same(api.selectedCost(cv, idx, ctx=ctx), oracle.port.selected_cost(cv, idx))
    cv16 = (np.abs(cv) * 1000).astype(np.uint16)
    got = api.selectedCost(cv16, idx, ctx=ctx)
    assert got.dtype == np.uint16 and np.array_equal(got, np.take_along_axis(cv16, idx[..., None], 2)[..., 0])


@pytest.mark.parametrize("direction", [0, 1])
@pytest.mark.parametrize("tdir", [0, 1, 2])
@pytest.mark.parametrize("radius", [1, 2, 3])
def test_truncated_cost_volume_and_in_bound_domain(api, ctx, oracle, volume, direction, tdir, radius):
    cv, idx = volume
    got = api.truncatedCostVolume(cv, idx, 2, 2, radius, dDir=direction, sdir=tdir, ctx=ctx)
    assert same(got, oracle.port.truncated_cost_volume(cv, idx, 2, 2, radius, direction, tdir))
    got = api.extractInBoundDomain(idx, cv.shape[2], 2, 2, radius, dDir=direction, sdir=tdir, ctx=ctx)
    assert same(got, oracle.port.in_bound_domain(idx, cv.shape[2], 2, 2, radius, direction, tdir))


def test_truncated_cost_volume_strided_volume(api, ctx, oracle, volume):
    cv, idx = volume
    cvt = np.ascontiguousarray(cv.transpose(0, 2, 1)).transpose(0, 2, 1)  # the reference's [H][D][W] layout
    assert same(api.truncatedCostVolume(cvt, idx, 2, 2, 1, ctx=ctx), oracle.port.truncated_cost_volume(cv, idx, 2, 2, 1))


@pytest.mark.parametrize("kernel", [0, 1, 2])
def test_refine_disp_cost_interpolation(api, ctx, oracle, volume, kernel):
    cv, idx = volume
    src = np.abs(cv) + np.float32(0.05) if kernel == 2 else cv  # the Gaussian kernel takes logs
    tcv = oracle.port.truncated_cost_volume(src, idx, 2, 2, 2)
    got = api.refineDispCostInterpolation(tcv, idx, kernel, ctx=ctx)
    want = oracle.port.refine(tcv, idx, kernel)
    assert np.array_equal(np.isnan(got), np.isnan(want))
    ok = np.isfinite(want) & (np.abs(want - idx) < 4)  # ill-conditioned triplets (curvature ~ 0) amplify the last ulp
    assert np.nanmax(np.abs(got[ok] - want[ok])) <= (1e-4 if kernel == 2 else 1e-5)
    # fused: straight from the cost volume (radius 1)
    tcv1, fused = api.refineFromCostVolume(src, idx, 2, 2, kernel, return_tcv=True, ctx=ctx)
    want_tcv = oracle.port.truncated_cost_volume(src, idx, 2, 2, 1)
    assert same(tcv1, want_tcv)
    want1 = oracle.port.refine(want_tcv, idx, kernel)
    assert np.array_equal(np.isnan(fused), np.isnan(want1))
    ok = np.isfinite(want1) & (np.abs(want1 - idx) < 4)
    assert np.nanmax(np.abs(fused[ok] - want1[ok])) <= (1e-4 if kernel == 2 else 1e-5)
    assert np.array_equal(api.refineFromCostVolume(src, idx, 2, 2, kernel, ctx=ctx), fused, equal_nan=True)


def test_refine_rejects_even_depth_like_the_reference(api, ctx):
    tcv = np.zeros((4, 5, 4), np.float32)
    idx = np.zeros((4, 5), np.int32)
    assert api.refineDispCostInterpolation(tcv, idx, ctx=ctx) is None  # cost_based_refinement.h:141-143
    assert api.refineDispCostInterpolation(tcv[:, :, :1], idx, ctx=ctx) is None


def test_parabola_known_answer(api, ctx):
    # testCostRefinement.cpp:33-57: tcv = (a-b, 0, a+b) => refined = idx - b/(2a)
    rng = np.random.default_rng(1)
    a, b = rng.uniform(0.5, 2.0, (6, 9)), rng.uniform(-0.4, 0.4, (6, 9))
    tcv = np.stack([a - b, np.zeros_like(a), a + b], -1).astype(np.float32)
    idx = rng.integers(0, 20, (6, 9)).astype(np.int32)
    got = api.refineDispCostInterpolation(tcv, idx, api.PARABOLA, ctx=ctx)
    assert np.max(np.abs(got - (idx - b / (2 * a)))) < 1e-5
