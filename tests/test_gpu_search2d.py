"""GPU parity of the 2-D search functions: unfoldBased2dDisparityCostVolume,
extractSelected2dIndex, selected2dIndexToDisp, truncatedBidirectionaCostVolume, refineDisp2dCostInterpolation and
refineDisp2dCostPatchInterpolation, against the CPU oracle, the golden vectors of the compiled reference and
restatements of the reference's own tests (test/unittests/testCorrelation2d.cpp).

Tolerance: the kernel sums in the reference's channel order (the `exact_sums` path of the 1-D volumes): 2e-6 of the
largest magnitude (1.0 for the normalised correlations), NaNs in the same cells. Indices, disparities and the truncated
volume (a gather) are bit-exact. refineDisp2dCostInterpolation: bit-exact against the oracle and the golden vectors of the
compiled reference for the Equiangular and Parabola kernels (correctly rounded operations in the same order); 2e-6 absolute
for the Gaussian kernel (logf of the device vs glibc, the 1-D refinement's bound; 1e-4 for its anisotropic estimate, whose line
intersection amplifies those last-bit differences). Patch interpolation: 1e-4 absolute
against the float64 oracle on patches with a clear extremum (closed-form float32 solve of the normal equations).
"""
import os

import numpy as np
import pytest

from libstevi_b200 import synth
from test_oracle_pinning import check_2d_cost_refinement, check_2d_disparity, check_2d_matching, lstsq_patch_deltas

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
        got = api.unfoldBased2dDisparityCostVolume(getattr(api, name), L, R, 2, 1, search, dDir=direction, exact_sums=True, ctx=ctx)
        close(name, np.ascontiguousarray(got), oracle.port.cost_volume_2d(getattr(oracle, name), L, R, 2, 1, search, direction))


@pytest.mark.parametrize("name", ["ZNCC", "SSD", "ZSAD", "CC"])
@pytest.mark.parametrize("W", [40, 132, 516])
@pytest.mark.parametrize("search", [(0, 1, 0, 6), (-1, 1, -3, 2), (-2, -1, -9, -5), (1, 2, 5, 5), (0, 0, -4, 3), (-1, 0, 1, 8)])
def test_cost_volume_2d_aligned_rows_match_oracle(api, ctx, oracle, name, W, search):
    """Widths that are multiples of 4 take the kernel with four source pixels x four column offsets per thread; its groups of
    offsets start at a multiple of 4 at or below lower1, so unaligned and negative windows exercise the partial groups."""
    L, R = synth.random_image(13, W, 5), synth.random_image(13, W, 6)
    got = api.unfoldBased2dDisparityCostVolume(getattr(api, name), L, R, 2, 1, search, exact_sums=True, ctx=ctx)
    want = oracle.port.cost_volume_2d(getattr(oracle, name), L, R, 2, 1, search)
    close(name, np.ascontiguousarray(got), want)
    c_layout = api.unfoldBased2dDisparityCostVolume(getattr(api, name), L, R, 2, 1, search, layout="c", exact_sums=True, ctx=ctx)
    assert np.array_equal(np.ascontiguousarray(got), c_layout, equal_nan=True)
    # the same cells through the four-candidate kernel (odd width: the last column dropped) agree bit for bit
    narrow = api.unfoldBased2dDisparityCostVolume(getattr(api, name), L[:, :W - 1].copy(), R[:, :W - 1].copy(), 2, 1, search, layout="c", exact_sums=True, ctx=ctx)
    inner = slice(8, W - 12)     # away from the dropped column's reach (window radius 2 + offsets up to 9)
    assert np.array_equal(narrow[:, inner], c_layout[:, inner], equal_nan=True)


@pytest.mark.parametrize("name", MFS)
def test_cost_volume_2d_golden(api, ctx, oracle, name):
    L, R = synth.random_image(int(G2["H"]), int(G2["W"]), int(G2["seed_l"])), synth.random_image(int(G2["H"]), int(G2["W"]), int(G2["seed_r"]))
    search = tuple(int(v) for v in G2["search"])
    for direction in (0, 1):
        got = api.unfoldBased2dDisparityCostVolume(getattr(api, name), L, R, 2, 1, search, dDir=direction, exact_sums=True, ctx=ctx)
        close(name, np.ascontiguousarray(got), G2[f"cv2d_{getattr(oracle, name)}_{direction}"])


def test_layouts_windows_and_device_operands(api, ctx, oracle):
    torch = pytest.importorskip("torch")
    L, R = synth.random_image(33, 150, 8), synth.random_image(33, 150, 9)   # more than one 128-column block
    search = (-1, 1, -6, 9)
    want = oracle.port.cost_volume_2d(oracle.ZNCC, L, R, 4, 3, search)
    ref_layout = api.unfoldBased2dDisparityCostVolume(api.ZNCC, L, R, 4, 3, search, exact_sums=True, ctx=ctx)
    assert ref_layout.strides == (150 * 16 * 3 * 4, 4, 150 * 16 * 4, 150 * 4)      # {W*dW*dH, 1, W*dW, W} elements
    c_layout = api.unfoldBased2dDisparityCostVolume(api.ZNCC, L, R, 4, 3, search, layout="c", exact_sums=True, ctx=ctx)
    close("ZNCC", np.ascontiguousarray(ref_layout), want)
    assert np.array_equal(np.ascontiguousarray(ref_layout), c_layout, equal_nan=True)
    tctx = api.Context(0, torch_stream=True)
    dev = api.unfoldBased2dDisparityCostVolume(api.ZNCC, torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda(), 4, 3, search, exact_sums=True, ctx=tctx)
    idx = api.extractSelected2dIndex(dev, api.SCORE, ctx=tctx)
    disp = api.selected2dIndexToDisp(idx, search, ctx=tctx)
    torch.cuda.synchronize()
    assert np.array_equal(dev.cpu().numpy(), np.ascontiguousarray(ref_layout), equal_nan=True)
    assert np.array_equal(disp.cpu().numpy(), oracle.port.index2d_to_disp(oracle.port.wta_2d(want, oracle.SCORE), search))
    tctx.close()
    # windows the box kernels do not cover (wider than 9, zero radius) go through the same path
    for hr, vr in ((6, 0), (0, 2)):
        got = api.unfoldBased2dDisparityCostVolume(api.SAD, L, R, hr, vr, (0, 1, -2, 2), exact_sums=True, ctx=ctx)
        close("SAD", np.ascontiguousarray(got), oracle.port.cost_volume_2d(oracle.SAD, L, R, hr, vr, (0, 1, -2, 2)))


@pytest.mark.parametrize("name", ["CC", "NCC", "ZCC", "ZNCC", "SSD", "ZSSD", "SAD"])
@pytest.mark.parametrize("hr,vr", [(1, 1), (2, 3), (4, 2)])
@pytest.mark.parametrize("W", [37, 132, 300])
@pytest.mark.parametrize("search", [(0, 2, 0, 4), (-3, -1, 2, 6), (-2, 2, -5, 0), (1, 1, -13, -9), (-1, 0, -3, 40), (4, 4, 7, 7)])
def test_cost_volume_2d_box_sums_match_oracle(api, ctx, oracle, name, hr, vr, W, search):
    """The default path: the O(1) box sums of the 1-D kernel, one pass per row offset with the target planes read shifted.
    Tolerance of the box form: 1e-4 absolute for the normalised correlations, 1e-5 of the largest magnitude
    otherwise; NaNs (flat windows) in the same cells. Column windows that start off a multiple of 8, negative and far offsets,
    rows and columns of the target outside the image, several 128-column blocks, both layouts."""
    L, R = synth.random_image(19, W, 3 + W), synth.random_image(19, W, 4 + W)
    L[5:12, 8:20] = 0.25   # a flat patch: NaN for the normalised zero-mean functions (Q6)
    want = oracle.port.cost_volume_2d(getattr(oracle, name), L, R, hr, vr, search)
    for layout in ("ref", "c"):
        got = np.ascontiguousarray(api.unfoldBased2dDisparityCostVolume(getattr(api, name), L, R, hr, vr, search, layout=layout, ctx=ctx))
        assert got.shape == want.shape and np.array_equal(np.isnan(got), np.isnan(want))
        tol = 1e-4 if name in ("NCC", "ZNCC") else 1e-5 * max(1.0, float(np.nanmax(np.abs(want))))
        assert np.nanmax(np.abs(got - want)) <= tol
    l2r = np.ascontiguousarray(api.unfoldBased2dDisparityCostVolume(getattr(api, name), L, R, hr, vr, search, dDir=api.LEFT_TO_RIGHT, ctx=ctx))
    want = oracle.port.cost_volume_2d(getattr(oracle, name), L, R, hr, vr, search, oracle.L2R)
    assert np.array_equal(np.isnan(l2r), np.isnan(want))
    assert np.nanmax(np.abs(l2r - want)) <= (1e-4 if name in ("NCC", "ZNCC") else 1e-5 * max(1.0, float(np.nanmax(np.abs(want)))))


def test_row_offset_zero_equals_the_1d_volume(api, ctx):
    """The slice at row offset 0 of a search window starting at column offset 0 is the 1-D right-to-left volume of the
    reference-order kernel, bit for bit (same kernel, same arithmetic)."""
    L, R = synth.random_image(64, 200, 21), synth.random_image(64, 200, 22)
    cv2 = api.unfoldBased2dDisparityCostVolume(api.ZNCC, L, R, 3, 3, (-2, 2, 0, 11), layout="c", exact_sums=True, ctx=ctx)
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
        api.unfoldBased2dDisparityCostVolume(api.ZNCC, a, b, 1, 1, (0, 1, 0, 1), exact_sums=True, ctx=ctx)
    assert e.value.status == api.E_SHAPE
    with pytest.raises(api.SteviError) as e:       # :336-338: empty search window
        api.unfoldBased2dDisparityCostVolume(api.ZNCC, a, a, 1, 1, (1, 0, 0, 1), exact_sums=True, ctx=ctx)
    assert e.value.status == api.E_SHAPE
    with pytest.raises(api.SteviError) as e:
        api.unfoldBased2dDisparityCostVolume(api.CENSUS, a, a, 1, 1, (0, 1, 0, 1), exact_sums=True, ctx=ctx)
    assert e.value.status == api.E_UNSUPPORTED


@pytest.mark.parametrize("name", ["NCC", "ZNCC", "SSD", "ZSSD", "SAD", "ZSAD"])
@pytest.mark.parametrize("hr,vr,dw,dh", [(1, 1, 5, 5), (3, 2, 4, 6)])
def test_ref_test2dMatching_and_test2dDisparity(api, ctx, oracle, name, hr, vr, dw, dh):
    """testCorrelation2d.cpp:73-194 run against the CUDA functions."""
    def cv2d(mf, l, r, hr_, vr_, search):
        return api.unfoldBased2dDisparityCostVolume(mf, l, r, hr_, vr_, search, exact_sums=True, ctx=ctx)
    mf = getattr(oracle, name)   # oracle and api share the matching-function codes
    assert mf == getattr(api, name)
    check_2d_matching(cv2d, mf, hr, vr, dw, dh, 77)
    check_2d_disparity(cv2d, lambda v, s: api.extractSelected2dIndex(v, s, ctx=ctx),
                       lambda i, s: api.selected2dIndexToDisp(i, s, ctx=ctx), mf, hr, vr, 91)


# ---- truncatedBidirectionaCostVolume, refineDisp2dCostInterpolation, refineDisp2dCostPatchInterpolation ------------------

def refined_close(got, want, kernel, oracle, aniso=0):
    assert got.shape == want.shape and got.dtype == np.float32
    if kernel == oracle.GAUSSIAN:
        # a delta next to the |delta| > 1 cut may fall on either side of it: compare where both kept their estimate
        kept = (got != np.round(got)) & (want != np.round(want))
        # the line intersection of the anisotropic estimate divides by 1 - a0*a1 and amplifies the last-bit differences of logf
        assert np.max(np.abs(got - want)[kept], initial=0.0) <= (1e-4 if aniso else 2e-6)
        assert np.mean((got != want) & ~kept) <= 0.01
    else:
        assert np.array_equal(got, want)


@pytest.mark.parametrize("name,strategy", [("ZNCC", 1), ("SSD", 0)])
def test_truncation_and_refinement_2d_golden(api, ctx, oracle, name, strategy):
    mf = getattr(oracle, name)
    cv, idx, disp = G2[f"cv2d_{mf}_1"], G2[f"wta2d_{mf}"], G2[f"disp2d_{mf}"]
    for r0, r1 in ((1, 1), (2, 1), (1, 0)):
        t = api.truncatedBidirectionaCostVolume(cv, idx, r0, r1, ctx=ctx)
        assert np.array_equal(t, G2[f"tcv2d_{mf}_{r0}_{r1}"], equal_nan=True)
        for kernel in (oracle.EQUIANGULAR, oracle.PARABOLA, oracle.GAUSSIAN):
            for aniso in (0, 1):
                got = api.refineDisp2dCostInterpolation(t, disp, kernel, aniso, ctx=ctx)
                refined_close(got, G2[f"refined2d_{mf}_{r0}_{r1}_{kernel}_{aniso}"], kernel, oracle, aniso)


@pytest.mark.parametrize("name,strategy", [("ZNCC", 1), ("NCC", 1), ("SAD", 0), ("ZSSD", 0)])
def test_truncation_and_refinement_2d_match_oracle(api, ctx, oracle, name, strategy):
    L, R = synth.random_image(37, 150, 41), synth.random_image(37, 150, 42)
    search = (-2, 3, -4, 4)
    cv = oracle.port.cost_volume_2d(getattr(oracle, name), L, R, 2, 2, search)
    idx = oracle.port.wta_2d(cv, strategy)
    ref_layout = np.ascontiguousarray(cv.transpose(0, 2, 3, 1)).transpose(0, 3, 1, 2)     # the reference's cost-volume strides
    for r0, r1 in ((1, 1), (0, 0), (3, 0), (1, 2), (2, 5)):
        want = oracle.port.truncated_cost_volume_2d(cv, idx, r0, r1)
        for vol in (cv, ref_layout):
            t = api.truncatedBidirectionaCostVolume(vol, idx, r0, r1, ctx=ctx)
            assert t.shape == want.shape and np.array_equal(t, want, equal_nan=True)
        for kernel in (oracle.EQUIANGULAR, oracle.PARABOLA, oracle.GAUSSIAN):
            for aniso in (0, 1):
                refined_close(api.refineDisp2dCostInterpolation(want, idx, kernel, aniso, ctx=ctx), oracle.port.refine_2d(want, idx, kernel, aniso), kernel, oracle, aniso)
        # strided operands: the neighbourhood axes first in memory
        ts = np.ascontiguousarray(want.transpose(2, 3, 0, 1)).transpose(2, 3, 0, 1)
        assert np.array_equal(api.refineDisp2dCostInterpolation(ts, idx, oracle.PARABOLA, 1, ctx=ctx), oracle.port.refine_2d(want, idx, oracle.PARABOLA, 1))


def test_refinement_2d_reads_the_centre_pixel_for_the_volume_kind(api, ctx, oracle):
    """cost_based_refinement.h:184-206: scores vs costs is decided from the image's centre pixel alone; the anisotropic
    estimate of every other pixel follows that decision."""
    rng = np.random.default_rng(3)
    t = rng.uniform(0.1, 1.0, (6, 9, 5, 5)).astype(np.float32)
    raw = rng.integers(-3, 4, (6, 9, 2)).astype(np.int32)
    for centre_is_peak in (False, True):
        t[3, 4] = rng.uniform(0.4, 0.6, (5, 5))
        t[3, 4, 2, 2] = 0.9 if centre_is_peak else 0.1
        want = oracle.port.refine_2d(t, raw, oracle.PARABOLA, 1)
        assert np.array_equal(api.refineDisp2dCostInterpolation(t, raw, oracle.PARABOLA, 1, ctx=ctx), want)
    t[0, 0, 1, 2] = np.nan      # NaN in a triplet: both deltas dropped
    got = api.refineDisp2dCostInterpolation(t, raw, oracle.EQUIANGULAR, 0, ctx=ctx)
    assert np.array_equal(got[0, 0], raw[0, 0].astype(np.float32))
    assert np.array_equal(got, oracle.port.refine_2d(t, raw, oracle.EQUIANGULAR, 0))


def test_ref_test2dCostRefinement(api, ctx):
    """testCostRefinement.cpp:59-227 run against the CUDA functions."""
    check_2d_cost_refinement(lambda t, raw, k, a: api.refineDisp2dCostInterpolation(t, raw, k, a, ctx=ctx),
                             lambda t, raw, k: api.refineDisp2dCostPatchInterpolation(t, raw, k, ctx=ctx), 5)


def test_patch_refinement_matches_oracle(api, ctx, oracle):
    rng = np.random.default_rng(12)
    t = rng.uniform(0.5, 2.0, (40, 150, 5, 3)).astype(np.float32)
    t[:, :, 2, 1] = 0.1
    raw = rng.integers(-5, 6, (40, 150, 2)).astype(np.int32)
    for kernel in (oracle.PARABOLA, oracle.GAUSSIAN):
        want = oracle.port.refine_2d_patch(t, raw, kernel)
        got = api.refineDisp2dCostPatchInterpolation(t, raw, kernel, ctx=ctx)
        kept = (got != raw) .all(axis=2) & (want != raw).all(axis=2)
        assert kept.mean() > 0.6
        assert np.max(np.abs(got - want)[kept]) <= 1e-4
        if oracle.ref_ibr is not None:      # the reference's own refineCostPatch (its Eigen inverses on the stand-in's Gauss-Jordan)
            ref = oracle.ref_ibr.refine_2d_patch(t, raw, kernel)
            assert np.array_equal((got != raw).all(axis=2), (ref != raw).all(axis=2))
            assert np.max(np.abs(got - ref)) <= 1e-4
        d = got[0, 0] - raw[0, 0]
        ref64 = lstsq_patch_deltas(np.log(t[0, 0, 1:4]) if kernel == oracle.GAUSSIAN else t[0, 0, 1:4])
        if np.all(np.abs(ref64) < 0.99):
            assert np.allclose(d, ref64, atol=1e-4)
    with pytest.raises(api.SteviError) as e:       # cost_based_refinement.h:82 static_assert
        api.refineDisp2dCostPatchInterpolation(t, raw, oracle.EQUIANGULAR, ctx=ctx)
    assert e.value.status == api.E_ARG


def test_refinement_2d_shape_rules_and_device_operands(api, ctx, oracle):
    torch = pytest.importorskip("torch")
    z = np.zeros((3, 3, 2), np.int32)
    for shape in ((3, 3, 4, 3), (3, 3, 3, 1), (3, 3, 3, 6)):      # cost_based_refinement.h:180-182, :386-388: empty array
        assert api.refineDisp2dCostInterpolation(np.zeros(shape, np.float32), z, ctx=ctx) is None
        assert api.refineDisp2dCostPatchInterpolation(np.zeros(shape, np.float32), z, ctx=ctx) is None
    L, R = synth.random_image(64, 200, 51), synth.random_image(64, 200, 52)
    search = (-2, 2, -3, 3)
    tctx = api.Context(0, torch_stream=True)
    cv = api.unfoldBased2dDisparityCostVolume(api.ZNCC, torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda(), 2, 2, search, exact_sums=True, ctx=tctx)
    idx = api.extractSelected2dIndex(cv, api.SCORE, ctx=tctx)
    t = api.truncatedBidirectionaCostVolume(cv, idx, 1, 1, ctx=tctx)
    disp = api.selected2dIndexToDisp(idx, search, ctx=tctx)
    ref = api.refineDisp2dCostInterpolation(t, disp, api.PARABOLA, api.ANISOTROPIC, ctx=tctx)
    torch.cuda.synchronize()
    cvh = cv.cpu().numpy()
    want_t = oracle.port.truncated_cost_volume_2d(np.ascontiguousarray(cvh), idx.cpu().numpy(), 1, 1)
    assert np.array_equal(t.cpu().numpy(), want_t, equal_nan=True)
    assert np.array_equal(ref.cpu().numpy(), oracle.port.refine_2d(want_t, disp.cpu().numpy(), oracle.PARABOLA, 1))
    tctx.close()
