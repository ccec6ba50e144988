"""GPU parity of the coarse-to-fine path: averagePoolingDownsample, computeGuidedCV and
hiearchicalTruncatedCostVolume against the CPU oracle. Pooling is bit-exact. The guided volume's comparisons are summed
in a different order than the reference's (raw window pixels instead of normalised feature vectors), so the picked
disparity may differ where two candidates are equal to ~1e-6: the gate is equality on >= 99.9 % of the pixels, the
volume compared (1e-4) where the disparities agree, and the reference's own statistical criterion."""
import numpy as np
import pytest

from libstevi_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("h,w", [(40, 64), (25, 43), (2, 3), (31, 2)])
def test_avgpool2(api, ctx, oracle, h, w):
    img = synth.random_image(h, w, 3)
    got = api.averagePoolingDownsample(img, 2, ctx=ctx)
    want = oracle.port.avgpool2(img)
    assert got.shape == want.shape and np.array_equal(got, want)
    got = api.averagePoolingDownsample(np.asfortranarray(img), 2, ctx=ctx)
    assert np.array_equal(got, want)


def test_avgpool_general_window(api, ctx):
    # downsampling.h:67-118 restated: window 3 on a 9x12 image (no remainder, so the reference's swapped offsets are 0)
    img = synth.random_image(9, 12, 4)
    got = api.averagePoolingDownsample(img, 3, ctx=ctx)
    want = img.reshape(3, 3, 4, 3).transpose(0, 2, 1, 3).reshape(3, 4, 9)
    acc = np.zeros((3, 4), np.float32)
    for k in range(9):
        acc = acc + want[..., k]
    assert np.array_equal(got, acc / np.float32(9))


def agree(got, want, frac=0.999):
    (gt, gd), (wt, wd) = got, want
    assert gd.shape == wd.shape and gt.shape == wt.shape
    eq = gd == wd
    assert eq.mean() >= frac, eq.mean()
    a, b = gt[eq], wt[eq]
    assert np.array_equal(np.isnan(a), np.isnan(b))
    scale = max(1.0, float(np.nanmax(np.abs(b))))
    assert np.nanmax(np.abs(a - b)) / scale <= 1e-4


@pytest.mark.parametrize("name", ["ZNCC", "NCC", "CC", "ZCC", "SSD", "ZSSD", "SAD", "ZSAD"])
@pytest.mark.parametrize("direction", [0, 1])
def test_guided_cv(api, ctx, oracle, name, direction):
    L, R, gt = synth.stereo_pair(36, 58, 6, d_bg=4, d_sq=10, dtype=np.float32)
    guide = (gt[::2, ::2] // 2).astype(np.int32)
    if direction == 0:
        guide = -guide  # deeper levels hand disp_estimate = sign * d down (hierarchical.h:230)
    for radius in (1, 2):
        got = api.computeGuidedCV(getattr(api, name), L, R, 2, 1, guide, radius, dDir=direction, ctx=ctx)
        want = oracle.port.guided_cv(getattr(oracle, name), L, R, 2, 1, guide, radius, direction)
        agree(got, want)


def test_guided_cv_rejects_one_pixel_axes(api, ctx):
    L = synth.random_image(1, 20, 1)
    assert api.computeGuidedCV(api.ZNCC, L, L, 1, 1, np.zeros((2, 2), np.int32), ctx=ctx) is None


@pytest.mark.parametrize("depth", [1, 2, 3])
@pytest.mark.parametrize("name", ["ZNCC", "SAD"])
def test_hierarchical(api, ctx, oracle, depth, name):
    h, w, D, r = 48, 64, 16, 2
    L, R, gt = synth.stereo_pair(h, w, 12, d_bg=2, d_sq=6, dtype=np.float32)
    got = api.hiearchicalTruncatedCostVolume(getattr(api, name), depth, L, R, r, r, D, 2, ctx=ctx)
    want = oracle.port.hierarchical(getattr(oracle, name), depth, L, R, r, r, D, 2)
    agree(got, want, 0.995)
    # the reference's own criterion (testCorrelationHierarchical.cpp:22-139)
    disp = got[1]
    m = 2 * sum(2 ** k * r for k in range(depth + 1))
    assert disp.min() >= -m and disp.max() <= D + m
    side = min(h, w) // 3
    i0, j0 = (h - side) // 2, (w - side) // 2
    inner = (slice(i0 + r + 2, i0 + side - r - 2), slice(j0 + r + 2, j0 + side - r - 2))
    assert (disp[inner] == gt[inner]).mean() >= 0.95


def test_hierarchical_odd_sizes_and_per_level_radii(api, ctx, oracle):
    L, R, _ = synth.stereo_pair(45, 71, 8, d_bg=3, d_sq=7, dtype=np.float32)
    got = api.hiearchicalTruncatedCostVolume(api.ZNCC, 2, L, R, 2, 2, 14, 1, ctx=ctx)
    want = oracle.port.hierarchical(oracle.ZNCC, 2, L, R, 2, 2, 14, 1)
    agree(got, want, 0.995)
    got = api.hiearchicalTruncatedCostVolume(api.ZNCC, 2, L, R, (1, 2, 3), (1, 1, 2), 14, 2, ctx=ctx)
    assert got[0].shape == (45, 71, 5) and np.isfinite(got[0]).mean() > 0.9


@pytest.mark.parametrize("name", ["ZNCC", "SSD", "ZSAD", "CC", "SAD"])
@pytest.mark.parametrize("ddir", ["r2l", "l2r"])
def test_guided_cv_on_feature_volumes(api, ctx, oracle, name, ddir):
    """computeGuidedCV with the reference's own operands (hierarchical.h:73-233): caller-supplied feature volumes and a guide of
    half the size. Disparities equal; volumes equal to 2e-6 of the largest magnitude (sums in the reference's channel order)."""
    rng = np.random.default_rng(17)
    L, R = synth.random_image(37, 90, 51), synth.random_image(37, 86, 52)
    mf = getattr(oracle, name)
    fl, fr = oracle.port.features_for_matchfunc(mf, oracle.port.unfold(L, 2, 1)), oracle.port.features_for_matchfunc(mf, oracle.port.unfold(R, 2, 1))
    guide = rng.integers(-3, 9, (19, 45)).astype(np.int32)
    d_api, d_o = (api.RIGHT_TO_LEFT, oracle.R2L) if ddir == "r2l" else (api.LEFT_TO_RIGHT, oracle.L2R)
    for radius in (1, 2, 3):
        tcv, disp = api.computeGuidedCVFromFeatures(getattr(api, name), fl, fr, guide, radius, d_api, ctx=ctx)
        checkers = [oracle.port]
        if oracle.ref_ibr is not None and oracle.ref_ibr.has("guided_cv_features_f32") and name in ("ZNCC", "SSD", "ZSAD"):
            checkers.append(oracle.ref_ibr)
        for chk in checkers:
            wt, wd = chk.guided_cv_features(mf, fl, fr, guide, radius, d_o)
            assert np.array_equal(disp, wd)
            assert np.array_equal(np.isnan(tcv), np.isnan(wt))
            assert np.nanmax(np.abs(tcv - wt)) <= 2e-6 * max(1.0, float(np.nanmax(np.abs(wt))))
    assert api.computeGuidedCVFromFeatures(getattr(api, name), fl[:20], fr, guide, 2, d_api, ctx=ctx) is None   # heights differ: empty volume
