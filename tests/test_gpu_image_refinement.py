"""Feature-domain subpixel refinement (correlation/image_based_refinement.h): the GPU entry points
against (1) the reference's own known-answer tests, restated (test/unittests/testCorrelationFilters.cpp:590-1108: a source
vector built as a known affine mix of two neighbouring target vectors must come back at that subpixel position, 1e-4), and
(2) the reference's unmodified code compiled into oracle/_ref/libstevi_ref_ibr.so, whose Eigen calls run on the stand-in
least-squares routine of oracle/small_qr.h (float Householder QR with full pivoting).

Tolerance: both sides compute the same estimator with different rounding (fp64 normal equations on the GPU, float QR in the
checker), and the refinement takes hard decisions on nearly equal scores, so the gate is: |gpu - ref| <= 1e-3 px on >= 99.5 %
of the pixels, every pixel within the one-pixel range the refinement can move, NaN patterns identical."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

BARY = ["NCC", "ZNCC", "SSD", "ZSSD"]
L1 = ["SAD", "ZSAD"]


def smooth_pair(h, w, shift, seed):
    """A band-limited random texture and a copy displaced by a smooth sub-pixel field around `shift` columns."""
    rng = np.random.default_rng(seed)
    big = rng.uniform(-1, 1, (h, w + 64)).astype(np.float64)
    k = np.array([1, 4, 6, 4, 1], np.float64) / 16
    for _ in range(2):
        big = np.apply_along_axis(lambda r: np.convolve(r, k, mode="same"), 1, big)
        big = np.apply_along_axis(lambda c: np.convolve(c, k, mode="same"), 0, big)
    big += 0.02 * rng.standard_normal(big.shape)
    left = big[:, 16:16 + w]
    jj = np.arange(w)[None, :] + 16.0
    field = shift + 0.45 * np.sin(np.arange(h)[:, None] / 7.0) * np.cos(np.arange(w)[None, :] / 11.0)
    x = jj + field                                   # right(i,j) = left(i, j + field): RightToLeft disparity = field
    x0 = np.floor(x).astype(int)
    a = x - x0
    rows = np.arange(h)[:, None]
    right = (1 - a) * big[rows, x0] + a * big[rows, x0 + 1]
    return left.astype(np.float32), right.astype(np.float32), field


def features(oracle, mf_name, left, right, hr=2, vr=2, processed=True):
    fl, fr = oracle.port.unfold(left, hr, vr), oracle.port.unfold(right, hr, vr)
    if not processed:
        return fl, fr
    mf = getattr(oracle, mf_name)
    return oracle.port.features_for_matchfunc(mf, fl), oracle.port.features_for_matchfunc(mf, fr)


def agree(got, want, frac=0.995, tol=1e-3, reach=1.0 + 1e-3):
    assert got.shape == want.shape
    assert np.array_equal(np.isnan(got), np.isnan(want))
    ok = np.isfinite(want)
    diff = np.abs(got[ok] - want[ok])
    assert (diff <= tol).mean() >= frac, f"only {(diff <= tol).mean():.4f} of the pixels within {tol}"
    assert diff.max() <= 2 * reach
    return float((diff <= tol).mean())


def need_ref(oracle):
    if oracle.ref_ibr is None:
        pytest.skip("oracle/_ref/libstevi_ref_ibr.so was not built")
    return oracle.ref_ibr


# ---- the reference's known-answer tests ------------------------------------------------------------------------------------

@pytest.mark.parametrize("name", BARY + L1)
@pytest.mark.parametrize("c_radius,adj", [(1, -0.3), (1, 0.3), (3, -0.25), (3, 0.25), (5, -0.3), (5, 0.3)])
def test_barycentric_known_answer(api, ctx, name, c_radius, adj):
    """testBarycentricNccRefinement / Ssd / Sad (testCorrelationFilters.cpp:590-836): refined position of the first source
    vector = 1 + adjustment, within 1e-4."""
    rng = np.random.default_rng(100 * c_radius + int(adj * 100) + 7)
    f = (2 * c_radius + 1) ** 2
    tgt = rng.uniform(-1, 1, (1, 3, f)).astype(np.float32)
    src = np.zeros((1, 3, f), np.float32)
    src[0, 0] = (1 + adj) * tgt[0, 1] - adj * tgt[0, 0] if adj < 0 else (1 - adj) * tgt[0, 1] + adj * tgt[0, 2]
    idx = np.array([[1, 0, 0]], np.int32)
    out = api.refineBarycentricDisp(getattr(api, name), tgt, src, idx, ctx=ctx)
    assert abs(out[0, 0] - (1 + adj)) < 1e-4
    assert out[0, 2] == 0   # target column at the border: the index is kept (:214)


@pytest.mark.parametrize("name", BARY)
def test_barycentric_symmetric_known_answer(api, ctx, name):
    """testBarycentricSymmetric*Refinement (testCorrelationFilters.cpp:838-1108): a source that is an affine mix (weights summing
    to one) of the three target vectors around the winner comes back at the barycentre of the weights."""
    rng = np.random.default_rng(11)
    for w in ((0.2, 0.7, 0.1), (0.05, 0.6, 0.35), (0.3, 0.4, 0.3)):
        f = 49
        tgt = rng.uniform(-1, 1, (1, 5, f)).astype(np.float32)
        src = np.zeros((1, 5, f), np.float32)
        src[0, 1] = w[0] * tgt[0, 1] + w[1] * tgt[0, 2] + w[2] * tgt[0, 3]   # pixel 1 matches column 2: index 1
        idx = np.array([[0, 1, 0, 0, 0]], np.int32)
        out = api.refineBarycentricSymmetricDisp(getattr(api, name), tgt, src, idx, 4, ctx=ctx)
        assert abs(out[0, 1] - (1 + w[2] - w[0])) < 1e-4
        assert out[0, 0] == 0   # index 0 is never refined (:68)


# ---- against the reference's code ------------------------------------------------------------------------------------------

@pytest.mark.parametrize("name", BARY + L1)
@pytest.mark.parametrize("ddir", ["r2l", "l2r"])
def test_refine_barycentric_disp(api, ctx, oracle, name, ddir):
    ref = need_ref(oracle)
    left, right, field = smooth_pair(40, 96, 5.0, 21)
    fl, fr = features(oracle, name, left, right)
    mf = getattr(oracle, name)
    if ddir == "r2l":
        cv = oracle.port.aggregate_cost(mf, fl, fr, 0, 11, oracle.R2L)
        idx = oracle.port.wta(cv, oracle.strategy_of(mf))
        got = api.refineBarycentricDisp(getattr(api, name), fl, fr, idx, api.RIGHT_TO_LEFT, ctx=ctx)
        want = ref.refine_barycentric(mf, fl, fr, idx, oracle.R2L)
        port = oracle.port.refine_barycentric(mf, fl, fr, idx, oracle.R2L)
    else:
        cv = oracle.port.aggregate_cost(mf, fr, fl, 0, 11, oracle.L2R)   # source = the first volume, displaced the other way
        idx = oracle.port.wta(cv, oracle.strategy_of(mf))
        got = api.refineBarycentricDisp(getattr(api, name), fr, fl, idx, api.LEFT_TO_RIGHT, ctx=ctx)
        want = ref.refine_barycentric(mf, fr, fl, idx, oracle.L2R)
        port = oracle.port.refine_barycentric(mf, fr, fl, idx, oracle.L2R)
    if name in L1:
        # zero-padded window cells give 0/0 break points, which the reference hands to std::sort as NaN (unspecified order):
        # compare where every window involved lies inside the image
        jd = np.arange(idx.shape[1])[None, :] + (idx if ddir == "r2l" else -idx)
        inside = (jd >= 3) & (jd < idx.shape[1] - 3)
        inside[:2] = False
        inside[-2:] = False
        inside[:, :2] = False
        inside[:, -2:] = False
        got, want, port = np.where(inside, got, 0), np.where(inside, want, 0), np.where(inside, port, 0)
    agree(got, want)
    agree(got, port)   # and the C restatement (oracle/stevi_oracle.c)
    if ddir == "r2l" and name in ("ZNCC", "ZSSD"):   # and the refinement does its job on this pair
        inner = (slice(4, -4), slice(8, -16))
        assert np.abs(got[inner] - field[inner]).mean() < np.abs(idx[inner] - field[inner]).mean()


@pytest.mark.parametrize("name", BARY)
@pytest.mark.parametrize("radius", [1, 2])
def test_refine_barycentric_symmetric_disp(api, ctx, oracle, name, radius):
    ref = need_ref(oracle)
    left, right, _ = smooth_pair(32, 80, 4.0, 33)
    fl, fr = features(oracle, name, left, right)
    mf = getattr(oracle, name)
    cv = oracle.port.aggregate_cost(mf, fl, fr, 0, 9, oracle.R2L)
    idx = oracle.port.wta(cv, oracle.strategy_of(mf))
    got = api.refineBarycentricSymmetricDisp(getattr(api, name), fl, fr, idx, 10, radius, ctx=ctx)
    want = ref.refine_barycentric_symmetric(mf, radius, fl, fr, idx, 10)
    agree(got, want, frac=0.99 if radius == 2 else 0.995)
    agree(got, oracle.port.refine_barycentric_symmetric(mf, radius, fl, fr, idx, 10), frac=0.99 if radius == 2 else 0.995)


@pytest.mark.parametrize("name", BARY)
@pytest.mark.parametrize("contiguity", ["QUEEN", "ROOK", "BISHOP"])
@pytest.mark.parametrize("symmetric", [False, True])
def test_refine_barycentric_2d_disp(api, ctx, oracle, name, contiguity, symmetric):
    ref = need_ref(oracle)
    left, right, _ = smooth_pair(36, 64, 2.0, 45)
    right = np.roll(right, 1, axis=0)   # and one row of vertical displacement
    fl, fr = features(oracle, name, left, right)
    mf = getattr(oracle, name)
    search = (-2, 2, -1, 5)
    cv = oracle.port.cost_volume_2d(mf, left, right, 2, 2, search)
    idx = oracle.port.index2d_to_disp(oracle.port.wta_2d(cv, oracle.strategy_of(mf)), search)
    idx[3, 5] = (7, 0)   # outside the search window: kept as it is (:717-722)
    cont = getattr(api, contiguity)
    got = api.refineBarycentric2dDisp(getattr(api, name), fl, fr, idx, search, cont, symmetric=symmetric, ctx=ctx)
    want = ref.refine_barycentric_2d(mf, cont, symmetric, fl, fr, idx, search)
    agree(got, want, frac=0.99)
    agree(got, oracle.port.refine_barycentric_2d(mf, cont, symmetric, fl, fr, idx, search), frac=0.99)
    assert tuple(got[3, 5]) == (7.0, 0.0)


@pytest.mark.parametrize("name", ["CC", "NCC", "ZCC", "ZNCC", "SSD", "ZSSD", "SAD", "ZSAD"])
def test_refine_cost_symmetric_disp(api, ctx, oracle, name):
    ref = need_ref(oracle) if name not in ("CC", "ZCC") else oracle.port   # the reference wrapper instantiates the other six
    left, right, _ = smooth_pair(28, 72, 3.0, 57)
    fl, fr = features(oracle, name, left, right)
    mf = getattr(oracle, name)
    cv = oracle.port.aggregate_cost(mf, fl, fr, 0, 9, oracle.R2L)
    idx = oracle.port.wta(cv, oracle.strategy_of(mf))
    got = api.refineCostSymmetricDisp(getattr(api, name), fl, fr, idx, cv, ctx=ctx)
    want = ref.refine_cost_symmetric(mf, fl, fr, idx, cv)
    assert np.array_equal(np.isnan(got), np.isnan(want))
    ok = np.isfinite(want) & (np.abs(want - idx) < 4)   # flat parabolas give unbounded offsets on both sides
    assert (np.abs(got[ok] - want[ok]) <= 1e-3).mean() >= 0.995


@pytest.mark.parametrize("name", BARY)
@pytest.mark.parametrize("prenormalize", [False, True])
def test_refined_barycentric_disp_from_images(api, ctx, oracle, name, prenormalize):
    """refinedBarycentricDisp (image_based_refinement.h:2502-2544): unfold -> features -> cost volume -> winner -> refinement,
    the whole chain on the GPU against the whole chain of the reference."""
    ref = need_ref(oracle)
    left, right, field = smooth_pair(30, 88, 6.0, 69)
    got = api.refinedBarycentricDisp(getattr(api, name), left, right, 2, 2, 12, prenormalize, ctx=ctx)
    want = ref.refined_barycentric_disp(getattr(oracle, name), left, right, 2, 2, 12, prenormalize)
    agree(got, want, frac=0.99)
    assert api.refinedBarycentricDisp(getattr(api, name), left[:10], right, 2, 2, 12, ctx=ctx) is None   # heights differ: empty array


def test_unsupported_functions_are_refused(api, ctx):
    f = np.zeros((4, 8, 9), np.float32)
    idx = np.zeros((4, 8), np.int32)
    for name in ("CC", "ZCC", "CENSUS"):
        with pytest.raises(api.SteviError):
            api.refineBarycentricDisp(getattr(api, name), f, f, idx, ctx=ctx)
    with pytest.raises(api.SteviError):   # the reference's randomised L1 descent is not provided
        api.refineBarycentricSymmetricDisp(api.SAD, f, f, idx, 4, ctx=ctx)


def test_full_size_refinement_properties(api, ctx):
    """1080p, 9x9 windows (F = 81), device-resident: the refinements move the winners towards the true sub-pixel field, stay
    within one pixel of them, and host and device operands give the same maps."""
    torch = pytest.importorskip("torch")
    left, right, field = smooth_pair(1080, 1920, 9.0, 5)
    tctx = api.Context(0, torch_stream=True)
    dl, dr = torch.from_numpy(left).cuda(), torch.from_numpy(right).cuda()
    fl, fr = api.unfold(4, 4, dl, ctx=tctx), api.unfold(4, 4, dr, ctx=tctx)
    zl, zr = (api.getFeatureVolumeForMatchFuncFromFeatures(api.ZNCC, f, ctx=tctx) for f in (fl, fr))
    cv = api.aggregateCost(api.ZNCC, zl, zr, 20, ctx=tctx)
    idx = api.extractSelectedIndex(cv, api.SCORE, ctx=tctx)
    inner = (slice(8, -8), slice(16, -40))
    raw = idx.cpu().numpy().astype(np.float32)
    base_err = np.abs(raw[inner] - field[inner]).mean()
    # (the cost-symmetric form averages two parabolas and is unbounded on flat ones: no one-pixel bound for it)
    for bounded, fn in ((True, lambda: api.refineBarycentricDisp(api.ZNCC, zl, zr, idx, ctx=tctx)),
                        (True, lambda: api.refineBarycentricSymmetricDisp(api.ZNCC, zl, zr, idx, 20, 1, ctx=tctx)),
                        (False, lambda: api.refineCostSymmetricDisp(api.ZNCC, zl, zr, idx, cv, ctx=tctx))):
        out = fn().cpu().numpy()
        assert np.isfinite(out[inner]).mean() > 0.999
        ok = np.isfinite(out)
        if bounded:
            assert np.abs(out[ok] - raw[ok]).max() <= 1.0 + 1e-5
        assert np.nanmean(np.abs(out[inner] - field[inner])) < 0.8 * base_err
    small = (slice(500, 540), slice(900, 1000))
    host = api.refineBarycentricDisp(api.ZNCC, zl[small].cpu().numpy().copy(), zr[small].cpu().numpy().copy(), raw[small].astype(np.int32), ctx=ctx)
    dev = api.refineBarycentricDisp(api.ZNCC, zl[small].contiguous(), zr[small].contiguous(), idx[small].contiguous(), ctx=tctx).cpu().numpy()
    assert np.array_equal(host, dev, equal_nan=True)
    tctx.close()
