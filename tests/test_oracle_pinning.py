"""Pins the CPU oracle (oracle/stevi_oracle.c, test infrastructure) to the reference.

1. against tests/golden/reference_outputs.npz — outputs of the UNMODIFIED reference headers (generator:
   tests/golden/make_golden.py), so this runs anywhere, including the GPU box where /root/reference is absent;
2. against the compiled reference itself (oracle/_ref) on further seeded inputs, when it is present;
3. against Qt-free restatements of the reference's own unit tests (analytic expectations):
   testCorrelationFilters.cpp:332-370 (ZNCC at the interior pixel), :384-445 (unfold), testCostRefinement.cpp:33-57
   (parabola), testCorrelationHierarchical.cpp:22-139 (statistical criterion).
"""
import os

import numpy as np
import pytest

import oracle as O
from libstevi_b200 import synth

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "reference_outputs.npz"))
H, W, D = int(G["H"]), int(G["W"]), int(G["D"])
L8, R8, _ = synth.stereo_pair(H, W, int(G["seed_u8"]), d_bg=2, d_sq=5, d_max=6)
Lf, Rf, _ = synth.stereo_pair(H, W, int(G["seed_f32"]), d_bg=2, d_sq=5, d_max=6, dtype=np.float32)
port = O.port


def same(a, b):
    return a.shape == b.shape and np.array_equal(a, b, equal_nan=True)


@pytest.mark.parametrize("hr,vr", [(4, 3), (3, 3), (2, 2), (6, 5)])
def test_golden_census(hr, vr):
    assert same(port.census(L8, hr, vr), G[f"census_u8_{hr}_{vr}"])
    assert same(port.census(Lf, hr, vr), G[f"census_f32_{hr}_{vr}"])
    for direction in (0, 1):
        assert same(port.cost_volume(O.CENSUS, L8, R8, hr, vr, D, direction), G[f"censuscv_u8_{hr}_{vr}_{direction}"])


def test_golden_sgm_wta():
    cv = G["censuscv_u8_4_3_1"]
    for ndir in (4, 8, 16):
        for strategy in (0, 1):
            assert same(port.sgm(cv, ndir, strategy, 10, 120, 100), G[f"sgm_u16_{ndir}_{strategy}"]), (ndir, strategy)
    assert same(port.sgm(cv, 8, 0, 10, 120, 100, (2, 1, 3, 2)), G["sgm_u16_8_0_margins"])
    assert same(port.wta(G["sgm_u16_8_0"], 0), G["wta_cost"])
    assert same(port.wta(cv, 0), G["wta_cv_u16"])
    zn = G[f"cv_f32_{O.ZNCC}_1"]
    assert same(port.sgm(zn, 8, 1, 0.001, 0.01, 100), G["sgm_f32_8_1"])
    assert same(port.sgm(G[f"cv_f32_{O.SSD}_1"], 8, 0, 0.5, 2.0, 100), G["sgm_f32_8_0"])


@pytest.mark.parametrize("mf", [O.CC, O.NCC, O.SSD, O.SAD, O.ZCC, O.ZNCC, O.ZSSD, O.ZSAD])
def test_golden_float_cost_volumes(mf):
    for direction in (0, 1):
        assert same(port.cost_volume(mf, Lf, Rf, 2, 2, D, direction), G[f"cv_f32_{mf}_{direction}"])


def test_golden_selection_refinement_hierarchy():
    zn = G[f"cv_f32_{O.ZNCC}_1"]
    idx = port.wta(zn, 1)
    assert same(idx, G["wta_score"])
    assert same(port.index_to_disp(idx, 1, 3), G["disp"])
    assert same(port.selected_cost(zn, idx), G["selected_cost"])
    for tdir in (0, 1, 2):
        assert same(port.truncated_cost_volume(zn, idx, 2, 2, 1, 1, tdir), G[f"tcv_{tdir}"])
    for kernel in (0, 1, 2):
        got, want = port.refine(G["tcv_0"], idx, kernel), G[f"refined_{kernel}"]
        assert np.allclose(got, want, rtol=1e-6, atol=1e-6, equal_nan=True)  # libm logf may differ in the last ulp
    assert same(port.avgpool2(Lf), G["avgpool"])
    assert same(port.avgpool2(np.ascontiguousarray(Lf[:25, :43])), G["avgpool_odd"])
    for depth in (1, 2):
        tcv, disp = port.hierarchical(O.ZNCC, depth, Lf, Rf, 2, 2, D, 2)
        assert same(disp, G[f"hier_disp_{depth}"])
        assert np.allclose(tcv, G[f"hier_tcv_{depth}"], atol=1e-6, equal_nan=True)
    assert same(port.unfold(L8[:6, :7], 2, 1), G["unfold_u8"])
    assert np.allclose(port.feature_volume(O.ZNCC, Lf[:8, :9], 1, 1), G["features_zncc"], atol=1e-7, equal_nan=True)


@pytest.mark.skipif(O.ref is None, reason="compiled reference (oracle/_ref) not present on this machine")
@pytest.mark.parametrize("seed", [7, 8])
def test_port_equals_compiled_reference(seed):
    ref = O.ref
    L, R, _ = synth.stereo_pair(30, 52, seed, d_bg=3, d_sq=7, d_max=8)
    for hr, vr in ((4, 3), (5, 1)):
        cv = port.cost_volume(O.CENSUS, L, R, hr, vr, 12)
        assert same(cv, ref.cost_volume(O.CENSUS, L, R, hr, vr, 12))
    for ndir in (4, 8, 16):
        for strategy in (0, 1):
            assert same(port.sgm(cv, ndir, strategy, 7, 50, 64, (1, 0, 2, 1)), ref.sgm(cv, ndir, strategy, 7, 50, 64, (1, 0, 2, 1)))
    Lq, Rq, _ = synth.stereo_pair(30, 52, seed, d_bg=3, d_sq=7, dtype=np.float32)
    for mf in (O.ZNCC, O.SAD):
        a, b = port.cost_volume(mf, Lq, Rq, 3, 2, 9), ref.cost_volume(mf, Lq, Rq, 3, 2, 9)
        assert np.allclose(a, b, atol=1e-6, equal_nan=True)
    a, b = port.hierarchical(O.ZNCC, 2, Lq, Rq, 2, 2, 12), ref.hierarchical(O.ZNCC, 2, Lq, Rq, 2, 2, 12)
    assert same(a[1], b[1])
    # winner neighbourhoods: truncatedCostVolume and its mask twin extractInBoundDomain, all direction combinations
    zn = port.cost_volume(O.ZNCC, Lq, Rq, 2, 2, 9)
    idx = port.wta(zn, O.SCORE)
    for direction in (0, 1):
        for tdir in (0, 1, 2):
            for radius in (1, 2):
                assert same(port.truncated_cost_volume(zn, idx, 2, 2, radius, direction, tdir),
                            ref.truncated_cost_volume(zn, idx, 2, 2, radius, direction, tdir))
                assert same(port.in_bound_domain(idx, 9, 2, 2, radius, direction, tdir),
                            ref.in_bound_domain(idx, 9, 2, 2, radius, direction, tdir))


# ---- restatements of the reference's own unit tests ------------------------------------------------------------------

def naive_zncc(win_s, win_t):
    """test/test_correlation_utils.h:52-98 InneficientZeromeanNormalizedCrossCorrelation"""
    s, t = win_s.astype(np.float64).ravel(), win_t.astype(np.float64).ravel()
    s, t = s - s.mean(), t - t.mean()
    return float((s * t).sum() / np.sqrt((s * s).sum() * (t * t).sum()))


@pytest.mark.parametrize("hr,vr", [(1, 1), (2, 3), (5, 5)])
@pytest.mark.parametrize("D", [1, 5])
def test_ref_testNCCFilter(hr, vr, D):
    # testCorrelationFilters.cpp:332-370: source window from the right image at column 0, target windows from the left
    # image at column d; only the single fully interior pixel is compared; tolerance 1e-3
    h, w = 2 * vr + 1, 2 * hr + D
    left, right = synth.random_image(h, w, 5, stream=1), synth.random_image(h, w, 5, stream=2)
    cv = port.cost_volume(O.ZNCC, left, right, hr, vr, D)
    for d in range(D):
        want = naive_zncc(right[:, 0:2 * hr + 1], left[:, d:d + 2 * hr + 1])
        assert abs(cv[vr, hr, d] - want) < 1e-3


def test_ref_testUnfoldOperator():
    # testCorrelationFilters.cpp:384-445 (auto padding): shape H x W x F and the centre pixel's features are the image
    hr, vr = 2, 1
    img = synth.random_image(2 * vr + 1, 2 * hr + 1, 9)
    fv = port.unfold(img, hr, vr)
    assert fv.shape == (3, 5, 15)
    assert np.array_equal(np.sort(fv[vr, hr]), np.sort(img.ravel()))
    assert np.array_equal(fv[vr, hr], img.ravel())  # Rotate0 channel order c = (2*hr+1)*k + l (unfold.h:180)


def test_ref_test1dCostParabola():
    # testCostRefinement.cpp:33-57: tcv = (a - b, 0, a + b)  =>  refined = idx - b / (2a)
    rng = np.random.default_rng(0)
    for _ in range(20):
        a, b = rng.uniform(0.5, 2.0), rng.uniform(-0.4, 0.4)
        tcv = np.array([[[a - b, 0.0, a + b]]], np.float32)
        got = port.refine(tcv, np.array([[3]], np.int32), O.PARABOLA)
        assert abs(got[0, 0] - (3 - b / (2 * a))) < 1e-5
    assert port.refine(np.zeros((1, 1, 4), np.float32), np.zeros((1, 1), np.int32)) is None  # even depth: empty array


@pytest.mark.parametrize("depth", [1, 2, 3])
def test_ref_testCorrelationHierarchical(depth):
    # testCorrelationHierarchical.cpp:22-139: 48x64 parallax-square pair, radii 2, D=16: all disparities within
    # [-m, D+m], m = 2*sum 2^k*r, and >= 95 % of the interior square pixels equal the ground truth
    h, w, D_, r = 48, 64, 16, 2
    L, R, gt = synth.stereo_pair(h, w, 12, d_bg=2, d_sq=6, dtype=np.float32)
    tcv, disp = port.hierarchical(O.ZNCC, depth, L, R, r, r, D_, 2)
    assert disp.shape == (h, w) and tcv.shape == (h, w, 5)
    m = 2 * sum(2 ** k * r for k in range(depth + 1))
    assert disp.min() >= -m and disp.max() <= D_ + m
    side = min(h, w) // 3
    i0, j0 = (h - side) // 2, (w - side) // 2
    inner = (slice(i0 + r + 2, i0 + side - r - 2), slice(j0 + r + 2, j0 + side - r - 2))
    assert (disp[inner] == gt[inner]).mean() >= 0.95


MASK = np.array([[1, 1, 2, 2, 2],
                 [1, 3, 3, 3, 2],
                 [0, 3, 7, 3, 0],
                 [5, 3, 3, 3, 4],
                 [5, 5, 0, 4, 4]], np.int32)


@pytest.mark.skipif(O.ref is None, reason="compiled reference (oracle/_ref) not present on this machine")
def test_port_equals_compiled_reference_unfold_family():
    ref = O.ref
    img2 = synth.random_image(11, 14, 21)
    img3 = np.stack([synth.random_image(11, 14, 22 + k) for k in range(3)], -1)
    for img in (img2, img3):
        for orientation in range(4):
            assert same(port.unfold_general(img, 2, 1, None, orientation), ref.unfold_general(img, 2, 1, None, orientation))
        for pad in ((0, 0, 0, 0), (3, 1, 0, 2)):
            assert same(port.unfold_general(img, 1, 2, pad, 0), ref.unfold_general(img, 1, 2, pad, 0))
        for pad in (None, (0, 0, 0, 0), (1, 2, 3, 0)):
            a, b = port.unfold_compressed(MASK, img, pad), ref.unfold_compressed(MASK, img, pad)
            assert a.shape == b.shape and np.allclose(a, b, atol=1e-6)
        a, b = port.unfold_compressed(MASK[:3, 1:4], img), ref.unfold_compressed(MASK[:3, 1:4], img)
        assert a.shape == b.shape and np.allclose(a, b, atol=1e-6)
    fv = port.unfold_general(img2, 3, 2)
    assert same(port.census_features(fv), ref.census_features(fv))
    assert same(port.census_features(fv), port.census(img2, 3, 2))


# ---- 2-D search ---------------------------------------------------------------------------------

G2 = np.load(os.path.join(os.path.dirname(__file__), "golden", "reference_outputs_2d.npz"))


@pytest.mark.parametrize("mf", [O.CC, O.NCC, O.SSD, O.SAD, O.ZCC, O.ZNCC, O.ZSSD, O.ZSAD])
def test_golden_2d_cost_volumes(mf):
    Lq, Rq = synth.random_image(int(G2["H"]), int(G2["W"]), int(G2["seed_l"])), synth.random_image(int(G2["H"]), int(G2["W"]), int(G2["seed_r"]))
    search = tuple(int(v) for v in G2["search"])
    for direction in (0, 1):
        assert same(port.cost_volume_2d(mf, Lq, Rq, 2, 1, search, direction), G2[f"cv2d_{mf}_{direction}"])


def test_golden_2d_selection():
    search = tuple(int(v) for v in G2["search"])
    for mf, strategy in ((O.ZNCC, 1), (O.SSD, 0)):
        idx = port.wta_2d(G2[f"cv2d_{mf}_1"], strategy)
        assert same(idx, G2[f"wta2d_{mf}"])
        assert same(port.index2d_to_disp(idx, search), G2[f"disp2d_{mf}"])
    # ties: the last extremum in (d1, d2) row-major order wins; a NaN at (0,0) sticks
    v = np.zeros((1, 2, 2, 3), np.float32)
    v[0, 1, 0, 0] = np.nan
    for strategy in (0, 1):
        got = port.wta_2d(v, strategy)
        assert got[0, 0].tolist() == [1, 2] and got[0, 1].tolist() == [0, 0]
    assert port.cost_volume_2d(O.ZNCC, np.zeros((5, 6), np.float32), np.zeros((5, 7), np.float32), 1, 1, (0, 1, 0, 1)) is None
    assert port.cost_volume_2d(O.ZNCC, np.zeros((5, 6), np.float32), np.zeros((5, 6), np.float32), 1, 1, (1, 0, 0, 1)) is None


@pytest.mark.skipif(O.ref is None, reason="compiled reference (oracle/_ref) not present on this machine")
def test_port_equals_compiled_reference_2d():
    ref = O.ref
    Lq, Rq = synth.random_image(14, 21, 31), synth.random_image(14, 21, 32)
    for search in ((0, 2, 0, 4), (-3, -1, 2, 6), (-2, 2, -5, 0)):
        for mf in (O.ZNCC, O.ZSAD, O.SSD, O.CC):
            for direction in (0, 1):
                a, b = port.cost_volume_2d(mf, Lq, Rq, 1, 2, search, direction), ref.cost_volume_2d(mf, Lq, Rq, 1, 2, search, direction)
                assert same(a, b)
            for strategy in (0, 1):
                assert same(port.wta_2d(b, strategy), ref.wta_2d(b, strategy))
        i2 = port.wta_2d(b, 0)
        assert same(port.index2d_to_disp(i2, search), ref.index2d_to_disp(i2, search))


def test_golden_2d_truncation_and_refinement():
    """truncatedBidirectionaCostVolume + refineDisp2dCostInterpolation of the port against outputs of the compiled reference."""
    for mf in (O.ZNCC, O.SSD):
        for r0, r1 in ((1, 1), (2, 1), (1, 0)):
            t = port.truncated_cost_volume_2d(G2[f"cv2d_{mf}_1"], G2[f"wta2d_{mf}"], r0, r1)
            assert same(t, G2[f"tcv2d_{mf}_{r0}_{r1}"])
            for kernel in (O.EQUIANGULAR, O.PARABOLA, O.GAUSSIAN):
                for aniso in (0, 1):
                    assert same(port.refine_2d(t, G2[f"disp2d_{mf}"], kernel, aniso), G2[f"refined2d_{mf}_{r0}_{r1}_{kernel}_{aniso}"])
    # the radii are uint8_t in the reference: 0 -> 1 on axis 0, 0 -> "as axis 0" on axis 1
    assert port.truncated_cost_volume_2d(G2[f"cv2d_{O.SSD}_1"], G2[f"wta2d_{O.SSD}"], 0, 0).shape[2:] == (3, 3)
    assert port.truncated_cost_volume_2d(G2[f"cv2d_{O.SSD}_1"], G2[f"wta2d_{O.SSD}"], 2, 0).shape[2:] == (5, 5)
    # cost_based_refinement.h:180-182: even or too small sides -> empty array
    z = np.zeros((3, 3, 2), np.int32)
    for shape in ((3, 3, 4, 3), (3, 3, 3, 1), (3, 3, 1, 3), (3, 3, 3, 6)):
        assert port.refine_2d(np.zeros(shape, np.float32), z) is None
        assert port.refine_2d_patch(np.zeros(shape, np.float32), z) is None


@pytest.mark.skipif(O.ref is None, reason="compiled reference (oracle/_ref) not present on this machine")
def test_port_equals_compiled_reference_2d_refinement():
    ref = O.ref
    for seed, (H, W) in enumerate(((13, 17), (9, 40))):
        Lq, Rq = synth.random_image(H, W, 301 + seed), synth.random_image(H, W, 352 + seed)
        search = (-2, 3, -3, 3)
        for mf, strategy in ((O.ZNCC, 1), (O.SAD, 0), (O.SSD, 0)):
            cv = ref.cost_volume_2d(mf, Lq, Rq, 2, 1, search, 1)
            idx = ref.wta_2d(cv, strategy)
            for r0, r1 in ((1, 1), (2, 1), (0, 0), (2, 3)):
                a, b = ref.truncated_cost_volume_2d(cv, idx, r0, r1), port.truncated_cost_volume_2d(cv, idx, r0, r1)
                assert same(a, b)
                for kernel in (O.EQUIANGULAR, O.PARABOLA, O.GAUSSIAN):
                    for aniso in (0, 1):
                        assert same(ref.refine_2d(a, idx, kernel, aniso), port.refine_2d(a, idx, kernel, aniso))
    assert ref.refine_2d(np.zeros((3, 3, 4, 3), np.float32), np.zeros((3, 3, 2), np.int32)) is None


def quadric_patch(n, A, b, s=1.0):
    """cost(dx, dy) = s * (p - b)^T A (p - b) on an n x n neighbourhood centred on 0 (testCostRefinement.cpp:78-82, :139-142)."""
    r = (n - 1) // 2
    t = np.zeros((1, 1, n, n), np.float32)
    for i in range(n):
        for j in range(n):
            p = np.array([i - r - b[0], j - r - b[1]], np.float64)
            t[0, 0, i, j] = s * float(p @ A @ p)
    return t


def check_2d_cost_refinement(refine_2d, refine_2d_patch, seed):
    """testCostRefinement.cpp:59-227: test2dCostIsotropicParabola, test2dCostAnisotropicParabola, test2dCostFullMatchingParabola.
    The reference compares with QCOMPARE (float fuzzy compare, 1e-5 relative) / 1e-6 absolute on random draws; the same
    bounds with a margin for the draw: 2e-5 absolute."""
    rng = np.random.default_rng(seed)
    raw = np.zeros((1, 1, 2), np.int32)
    for _ in range(20):
        s = rng.uniform(-1, 1)
        s = 0.5 if abs(s) < 1e-4 else s
        e = rng.uniform(-1, 1, 2)
        got = refine_2d(quadric_patch(3, np.eye(2), e, s), raw, O.PARABOLA, 0)
        assert np.allclose(got[0, 0], e, atol=2e-5), (got, e)
        s = rng.uniform(-1, 1)
        s = np.copysign(0.5, s) if abs(s) < 0.5 else s
        d = s * rng.uniform(0.7, 1, 2)
        b = rng.uniform(-0.5, 0.5, 2)
        for alpha, fn, n in ((0.1 * rng.uniform(-1, 1), lambda t: refine_2d(t, raw, O.PARABOLA, 1), 5),
                             (3 * rng.uniform(-1, 1), lambda t: refine_2d_patch(t, raw, O.PARABOLA), 3)):
            rot = np.array([[np.cos(alpha), -np.sin(alpha)], [np.sin(alpha), np.cos(alpha)]])
            A = rot @ np.diag(d) @ rot.T
            got = fn(quadric_patch(n, A, b))
            assert np.allclose(got[0, 0], b, atol=2e-5), (got, b, alpha)


def lstsq_patch_deltas(t):
    """refineCostPatch (cost_based_refinement.h:71-126) in float64 with a general least-squares solve."""
    v, h = np.meshgrid([-1.0, 0.0, 1.0], [-1.0, 0.0, 1.0], indexing="ij")
    v, h = v.ravel(), h.ravel()
    A = np.stack([v * v, v * h, h * h, v, h, np.ones(9)], axis=1)
    f = np.linalg.lstsq(A, t.astype(np.float64).ravel(), rcond=None)[0]
    M = np.array([[2 * f[0], f[1]], [f[1], 2 * f[2]]])
    return np.linalg.solve(M, -f[3:5])


def test_ref_test2dCostRefinement():
    check_2d_cost_refinement(port.refine_2d, port.refine_2d_patch, 5)


def test_patch_refinement_equals_a_general_least_squares_solve():
    rng = np.random.default_rng(9)
    for kernel in (O.PARABOLA, O.GAUSSIAN):
        n = 0
        for _ in range(200):
            t = rng.uniform(0.5, 2.0, (1, 1, 3, 3)).astype(np.float32)
            t[0, 0, 1, 1] = 0.1                      # a clear minimum at the centre
            want = lstsq_patch_deltas(np.log(t[0, 0]) if kernel == O.GAUSSIAN else t[0, 0])
            got = port.refine_2d_patch(t, np.array([[[3, -2]]], np.int32), kernel)[0, 0] - np.array([3, -2], np.float32)
            if np.all(np.abs(want) < 0.99):
                n += 1
                assert np.allclose(got, want, atol=1e-5), (got, want)
            elif np.any(np.abs(want) > 1.01):
                assert np.all(got == 0)
        assert n > 100
    assert port.refine_2d_patch(np.ones((1, 1, 3, 3), np.float32), np.zeros((1, 1, 2), np.int32), O.EQUIANGULAR) is None


def naive_match(mf, win_s, win_t):
    """test/test_correlation_utils.h InneficientMatchingFunction: the textbook formula on two windows."""
    s, t = win_s.astype(np.float64).ravel(), win_t.astype(np.float64).ravel()
    if mf in (O.ZCC, O.ZNCC, O.ZSSD, O.ZSAD):
        s, t = s - s.mean(), t - t.mean()
    if mf in (O.NCC, O.ZNCC):
        s, t = s / np.sqrt((s * s).sum()), t / np.sqrt((t * t).sum())
    if mf in (O.CC, O.NCC, O.ZCC, O.ZNCC):
        return float((s * t).sum())
    if mf in (O.SSD, O.ZSSD):
        return float(((s - t) ** 2).sum())
    return float(np.abs(s - t).sum())


def check_2d_matching(cost_volume_2d, mf, hr, vr, disp_w, disp_h, seed):
    """testCorrelation2d.cpp:73-124 test2dMatching: cell (v_radius, h_radius, i, j) of the volume is the matching function of
    the right image's first window against the left image's window shifted by (i, j); tolerance 1e-3 as in the reference."""
    h, w = 2 * vr + disp_h + 1, 2 * hr + disp_w + 1
    left, right = synth.random_image(h, w, seed) / 127.5 - 1.0, synth.random_image(h, w, seed + 1) / 127.5 - 1.0
    left, right = left.astype(np.float32), right.astype(np.float32)
    cv = cost_volume_2d(mf, left, right, hr, vr, (0, disp_h, 0, disp_w))
    win1 = right[0:2 * vr + 1, 0:2 * hr + 1]
    for i in range(disp_h):
        for j in range(disp_w):
            want = naive_match(mf, win1, left[i:i + 2 * vr + 1, j:j + 2 * hr + 1])
            assert abs(want - float(cv[vr, hr, i, j])) < 1e-3, (mf, i, j)


def check_2d_disparity(cost_volume_2d, wta_2d, index2d_to_disp, mf, hr, vr, seed):
    """testCorrelation2d.cpp:128-194 test2dDisparity: a window planted at offset (-1,-1) / (disp_h-1, disp_w-1) is found."""
    disp_w = disp_h = 3
    h, w = 2 * vr + disp_h + 1, 2 * hr + disp_w + 1
    rnd = lambda k: (synth.random_image(h, w, seed + k) / 127.5 - 1.0).astype(np.float32)
    source, target0, targetmax = rnd(0), rnd(1), rnd(2)
    for i in range(2 * vr + 1):
        for j in range(2 * hr + 1):
            target0[i, j] = source[i + 1, j + 1]
            targetmax[i + disp_h, j + disp_w] = source[i + 1, j + 1]
    search = (-1, disp_h - 1, -1, disp_w - 1)
    strategy = O.strategy_of(mf)
    d0 = index2d_to_disp(wta_2d(cost_volume_2d(mf, target0, source, hr, vr, search), strategy), search)
    assert d0[vr + 1, hr + 1].tolist() == [-1, -1]
    dm = index2d_to_disp(wta_2d(cost_volume_2d(mf, targetmax, source, hr, vr, search), strategy), search)
    assert dm[vr + 1, hr + 1].tolist() == [disp_h - 1, disp_w - 1]


@pytest.mark.parametrize("mf", [O.NCC, O.ZNCC, O.SSD, O.ZSSD, O.SAD, O.ZSAD])
@pytest.mark.parametrize("hr,vr,dw,dh", [(1, 1, 5, 5), (3, 2, 4, 6)])
def test_ref_test2dMatching_and_test2dDisparity(mf, hr, vr, dw, dh):
    check_2d_matching(port.cost_volume_2d, mf, hr, vr, dw, dh, 77)
    check_2d_disparity(port.cost_volume_2d, port.wta_2d, port.index2d_to_disp, mf, hr, vr, 91)


# ---- guided cost-volume aggregation ------------------------------------------------------------------

def bilateral_windows(guide, hr, vr, sigma_s=2.0, sigma_r=0.2):
    """What computeAggregationWindows (guided_cost_filtering.h:26-100) returns for a Gaussian spatial weight and a Gaussian
    weight on the guide difference: 0 outside the image."""
    H, W = guide.shape
    win = np.zeros((H, W, 2 * vr + 1, 2 * hr + 1), np.float32)
    for di in range(-vr, vr + 1):
        for dj in range(-hr, hr + 1):
            ws = np.float32(np.exp(-(di * di + dj * dj) / (2 * sigma_s ** 2)))
            i0, i1, j0, j1 = max(0, -di), min(H, H - di), max(0, -dj), min(W, W - dj)
            diff = guide[i0 + di:i1 + di, j0 + dj:j1 + dj] - guide[i0:i1, j0:j1]
            win[i0:i1, j0:j1, di + vr, dj + hr] = ws * np.exp(-(diff * diff) / np.float32(2 * sigma_r ** 2)).astype(np.float32)
    return win


def test_aggregation_port_against_the_definition_and_shape_rules():
    rng = np.random.default_rng(2)
    cv = rng.uniform(0, 4, (9, 12, 5)).astype(np.float32)
    win = bilateral_windows(synth.random_image(9, 12, 3), 2, 1)
    got = port.variable_cost_volume_aggregation(cv, win)
    want = np.zeros(cv.shape, np.float64)
    for i in range(9):
        for j in range(12):
            num, den = np.zeros(5), 0.0
            for di in range(-1, 2):
                for dj in range(-2, 3):
                    if 0 <= i + di < 9 and 0 <= j + dj < 12:
                        num += float(win[i, j, di + 1, dj + 2]) * cv[i + di, j + dj].astype(np.float64)
                        den += float(win[i, j, di + 1, dj + 2])
            want[i, j] = num / den
    assert np.allclose(got, want, rtol=0, atol=2e-6)
    # uniform weights on a constant volume: the constant comes back
    assert np.allclose(port.variable_cost_volume_aggregation(np.full((5, 6, 3), 2.5, np.float32), np.ones((5, 6, 3, 3), np.float32)), 2.5)
    # guided_cost_filtering.h:108-130
    for bad in (np.ones((8, 12, 3, 3)), np.ones((9, 11, 3, 3)), np.ones((9, 12, 2, 3)), np.ones((9, 12, 3, 4)), np.ones((9, 12, 0, 3))):
        assert port.variable_cost_volume_aggregation(cv, bad.astype(np.float32)) is None


@pytest.mark.skipif(O.ref is None, reason="compiled reference (oracle/_ref) not present on this machine")
def test_aggregation_port_equals_compiled_reference():
    rng = np.random.default_rng(1)
    for (H, W, D, n0, n1) in ((9, 13, 5, 3, 3), (12, 17, 7, 5, 5), (8, 20, 4, 7, 7), (10, 11, 6, 3, 7), (7, 9, 3, 9, 1), (6, 6, 2, 1, 1)):
        cv = rng.uniform(-1, 1, (H, W, D)).astype(np.float32)
        win = rng.uniform(0.01, 1, (H, W, n0, n1)).astype(np.float32)
        assert same(O.ref.variable_cost_volume_aggregation(cv, win), port.variable_cost_volume_aggregation(cv, win))
    assert O.ref.variable_cost_volume_aggregation(cv, win[:5]) is None
    assert O.ref.variable_cost_volume_aggregation(cv, np.ones((6, 6, 2, 3), np.float32)) is None


def test_aggregation_golden():
    cv = G2[f"cv2d_{O.SSD}_1"][:, :, :, 2]          # an (H, W, dH) slice of the golden volume used as a 1-D volume
    win = bilateral_windows(synth.random_image(int(G2["H"]), int(G2["W"]), int(G2["seed_r"])), 2, 2)
    assert same(port.variable_cost_volume_aggregation(cv, win), G2["aggregated_ssd_5x5"])


# ---- meanFilter2D / sigmaFilter (correlation_base.h:881-1077) ----------------------------------------------------------

@pytest.mark.parametrize("hr,vr,eh,ev", [(0, 0, 0, 0), (1, 1, 1, 1), (5, 1, 7, 7), (1, 5, 7, 7), (5, 5, 7, 7)])
def test_ref_testMeanFilter(hr, vr, eh, ev):
    # test/unittests/testCorrelationFilters.cpp:82-136, same rows: the last interior pixel against the plain window mean
    rng = np.random.default_rng(100 + hr * 7 + vr)
    h, w = 2 * vr + 1 + ev, 2 * hr + 1 + eh
    img = rng.uniform(-1, 1, (h, w)).astype(np.float32)
    want = img[ev:, eh:].sum(dtype=np.float32) / ((2 * hr + 1) * (2 * vr + 1))
    for chk in (O.port, O.ref):
        if chk is None:
            continue
        avg = chk.mean_filter_2d(img, hr, vr)
        assert abs(want - avg[vr + ev, hr + eh]) < 1e-3
        assert np.isnan(avg[:vr]).all() and np.isnan(avg[:, :hr]).all()   # outside the interior nothing is defined / written


@pytest.mark.parametrize("hr,vr", [(1, 1), (5, 1), (1, 5), (5, 5)])
def test_ref_testSigmaFilter(hr, vr):
    # testCorrelationFilters.cpp:172-221
    rng = np.random.default_rng(200 + hr * 7 + vr)
    img = rng.uniform(-1, 1, (2 * vr + 1, 2 * hr + 1)).astype(np.float32)
    mean = img.sum(dtype=np.float32) / img.size
    want = np.sqrt(((img - mean) ** 2).sum(dtype=np.float32))
    for chk in (O.port, O.ref):
        if chk is None:
            continue
        sigma = chk.sigma_filter(chk.mean_filter_2d(img, hr, vr), img, hr, vr)
        assert abs(want - sigma[vr, hr]) < 1e-3


@pytest.mark.skipif(O.ref is None, reason="compiled reference (oracle/_ref) not present on this machine")
@pytest.mark.parametrize("shape", [(40, 57), (33, 41, 3), (21, 30, 1)])
@pytest.mark.parametrize("dtype", [np.float32, np.uint8])
@pytest.mark.parametrize("hr,vr", [(1, 1), (3, 2), (0, 4), (6, 0)])
def test_filters_port_equals_compiled_reference(shape, dtype, hr, vr):
    rng = np.random.default_rng(7)
    img = (rng.random(shape) * 255).astype(np.uint8) if dtype == np.uint8 else rng.uniform(-1, 1, shape).astype(np.float32)
    a, b = O.port.mean_filter_2d(img, hr, vr), O.ref.mean_filter_2d(img, hr, vr)
    assert np.array_equal(a, b, equal_nan=True)          # running sums in the reference's order: bit for bit
    sa, sb = O.port.sigma_filter(b, img, hr, vr), O.ref.sigma_filter(b, img, hr, vr)
    assert np.array_equal(np.isnan(sa), np.isnan(sb))
    if dtype == np.float32 and len(shape) == 3:
        # the one instantiation g++ contracts into fused multiply-adds (see oracle/stevi_oracle.c): a few ulp
        inner = np.isfinite(sb)
        assert np.abs(sa[inner] - sb[inner]).max() <= 4e-7 * np.abs(sb[inner]).max()
    else:
        assert np.array_equal(sa, sb, equal_nan=True)
    assert O.port.mean_filter_2d(img[:2 * vr], hr, vr) is None if vr else True   # window taller than the image


GF = np.load(os.path.join(os.path.dirname(__file__), "golden", "reference_outputs_filters.npz"))


def filter_images():
    h, w = int(GF["H"]), int(GF["W"])
    return {"f32": synth.random_image(h, w, 301), "u8": synth.stereo_pair(h, w, 302)[0],
            "f32c3": np.stack([synth.random_image(h, w, 303 + k) for k in range(3)], axis=2),
            "u8c2": np.stack([synth.stereo_pair(h, w, 306 + k)[0] for k in range(2)], axis=2)}


@pytest.mark.parametrize("name", ["f32", "u8", "f32c3", "u8c2"])
@pytest.mark.parametrize("hr,vr", [(1, 1), (3, 2), (0, 4), (5, 0)])
def test_golden_filters(name, hr, vr):
    img = filter_images()[name]
    m = GF[f"mean_{name}_{hr}_{vr}"]
    assert same(port.mean_filter_2d(img, hr, vr), m)
    s, want = port.sigma_filter(m, img, hr, vr), GF[f"sigma_{name}_{hr}_{vr}"]
    if name == "f32c3":   # contracted by g++ in the compiled reference (oracle/stevi_oracle.c)
        assert np.array_equal(np.isnan(s), np.isnan(want)) and np.nanmax(np.abs(s - want)) <= 4e-7 * np.nanmax(want)
    else:
        assert same(s, want)


needs_ref_ibr = pytest.mark.skipif(O.ref_ibr is None, reason="oracle/_ref/libstevi_ref_ibr.so was not built")

# ---- feature-domain refinement (correlation/image_based_refinement.h) ----------------------------------------------------------
# The reference's code needs Eigen's FullPivHouseholderQR, which is not in this image: oracle/_ref/libstevi_ref_ibr.so is the
# reference's unmodified header with that one routine served by oracle/small_qr.h. What is pinned here: the stand-in solver
# against numpy's least squares; the reference's own known-answer tests (testCorrelationFilters.cpp:590-1108) on the compiled
# reference and on the C restatement; the restatement against the compiled reference on whole images.

def test_small_qr_solves_least_squares():
    import ctypes
    rng = np.random.default_rng(0)
    for m, n in ((25, 1), (25, 2), (49, 3), (81, 8), (9, 9), (5, 3)):
        A = rng.standard_normal((m, n)).astype(np.float32)
        b = rng.standard_normal(m).astype(np.float32)
        x = np.zeros(n, np.float32)
        Af = np.asfortranarray(A)
        O.port.lib.orc_qr_lstsq(m, n, Af.ctypes.data_as(ctypes.c_void_p), b.ctypes.data_as(ctypes.c_void_p), x.ctypes.data_as(ctypes.c_void_p))
        want = np.linalg.lstsq(A.astype(np.float64), b.astype(np.float64), rcond=None)[0]
        assert np.abs(x - want).max() < 2e-5 * max(1.0, np.abs(want).max())
    # rank-deficient: a basic solution (free unknowns 0) with the least-squares residual
    A = rng.standard_normal((30, 3)).astype(np.float32)
    A[:, 2] = A[:, 0] + A[:, 1]
    b = rng.standard_normal(30).astype(np.float32)
    x = np.zeros(3, np.float32)
    Af = np.asfortranarray(A)
    O.port.lib.orc_qr_lstsq(30, 3, Af.ctypes.data_as(ctypes.c_void_p), b.ctypes.data_as(ctypes.c_void_p), x.ctypes.data_as(ctypes.c_void_p))
    want = np.linalg.lstsq(A.astype(np.float64), b.astype(np.float64), rcond=1e-5)[0]   # rank 2 at float precision
    assert (x == 0).sum() == 1
    assert abs(np.linalg.norm(A @ x - b) - np.linalg.norm(A @ want - b)) < 1e-4


def _ibr_checkers():
    return [O.port] + ([O.ref_ibr] if O.ref_ibr is not None else [])


@pytest.mark.parametrize("mf", [O.NCC, O.ZNCC, O.SSD, O.ZSSD, O.SAD, O.ZSAD])
def test_barycentric_refinement_known_answers(mf):
    """testBarycentricNccRefinement / SsdRefinement / SadRefinement (testCorrelationFilters.cpp:590-836)."""
    rng = np.random.default_rng(5)
    for c_radius, adj in ((1, -0.3), (1, 0.3), (3, -0.25), (3, 0.25), (5, -0.3), (5, 0.3)):
        f = (2 * c_radius + 1) ** 2
        tgt = rng.uniform(-1, 1, (1, 3, f)).astype(np.float32)
        src = np.zeros((1, 3, f), np.float32)
        src[0, 0] = (1 + adj) * tgt[0, 1] - adj * tgt[0, 0] if adj < 0 else (1 - adj) * tgt[0, 1] + adj * tgt[0, 2]
        idx = np.array([[1, 0, 0]], np.int32)
        for chk in _ibr_checkers():
            out = chk.refine_barycentric(mf, tgt, src, idx)
            assert abs(out[0, 0] - (1 + adj)) < 1e-4, (chk.path, c_radius, adj, out)


@pytest.mark.parametrize("mf", [O.NCC, O.ZNCC, O.SSD, O.ZSSD])
def test_barycentric_symmetric_refinement_known_answers(mf):
    """testBarycentricSymmetric*Refinement (testCorrelationFilters.cpp:838-1108)."""
    rng = np.random.default_rng(6)
    for w in ((0.2, 0.7, 0.1), (0.05, 0.6, 0.35), (0.3, 0.4, 0.3)):
        tgt = rng.uniform(-1, 1, (1, 5, 49)).astype(np.float32)
        src = np.zeros((1, 5, 49), np.float32)
        src[0, 1] = w[0] * tgt[0, 1] + w[1] * tgt[0, 2] + w[2] * tgt[0, 3]
        idx = np.array([[0, 1, 0, 0, 0]], np.int32)
        for chk in _ibr_checkers():
            out = chk.refine_barycentric_symmetric(mf, 1, tgt, src, idx, 4)
            assert abs(out[0, 1] - (1 + w[2] - w[0])) < 1e-4


def _ibr_scene(mf, seed=3):
    rng = np.random.default_rng(seed)
    base = rng.uniform(-1, 1, (26, 80)).astype(np.float32)
    k = np.array([1, 2, 1], np.float32) / 4
    base = np.apply_along_axis(lambda r: np.convolve(r, k, mode="same"), 1, base).astype(np.float32)
    left, right = base[:, 4:68].copy(), (0.6 * base[:, 8:72] + 0.4 * base[:, 9:73]).astype(np.float32)
    fl = O.port.features_for_matchfunc(mf, O.port.unfold(left, 2, 1))
    fr = O.port.features_for_matchfunc(mf, O.port.unfold(right, 2, 1))
    cv = O.port.aggregate_cost(mf, fl, fr, 0, 9, O.R2L)
    return left, right, fl, fr, cv, O.port.wta(cv, O.strategy_of(mf))


def _mostly_equal(a, b, tol=1e-5, frac=0.998):
    assert np.array_equal(np.isnan(a), np.isnan(b))
    ok = np.isfinite(b)
    assert (np.abs(a[ok] - b[ok]) <= tol * np.maximum(1, np.abs(b[ok]))).mean() >= frac


@needs_ref_ibr
@pytest.mark.parametrize("mf", [O.NCC, O.ZNCC, O.SSD, O.ZSSD, O.SAD, O.ZSAD])
def test_port_refinements_equal_the_compiled_reference(mf):
    _, _, fl, fr, cv, idx = _ibr_scene(mf)
    inner = (slice(1, -1), slice(4, -14))   # SAD / ZSAD: zero-padded cells make 0/0 break points, sorted as NaN by the reference
    for d in (O.R2L, O.L2R):
        a, b = (fl, fr) if d == O.R2L else (fr, fl)
        i = idx if d == O.R2L else np.maximum(idx - 6, 0)
        p, r = O.port.refine_barycentric(mf, a, b, i, d), O.ref_ibr.refine_barycentric(mf, a, b, i, d)
        _mostly_equal(p[inner], r[inner]) if mf in (O.SAD, O.ZSAD) else _mostly_equal(p, r)
    _mostly_equal(O.port.refine_cost_symmetric(mf, fl, fr, idx, cv), O.ref_ibr.refine_cost_symmetric(mf, fl, fr, idx, cv))
    if mf in (O.SAD, O.ZSAD):
        return
    for radius in (1, 2):
        _mostly_equal(O.port.refine_barycentric_symmetric(mf, radius, fl, fr, idx, 10), O.ref_ibr.refine_barycentric_symmetric(mf, radius, fl, fr, idx, 10))
    left, right = _ibr_scene(mf)[:2]
    right = np.roll(right, 1, axis=0)
    fr2 = O.port.features_for_matchfunc(mf, O.port.unfold(right, 2, 1))
    search = (-2, 2, 0, 7)
    idx2 = O.port.index2d_to_disp(O.port.wta_2d(O.port.cost_volume_2d(mf, left, right, 2, 1, search), O.strategy_of(mf)), search)
    for cont in (0, 1, 2):
        for sym in (0, 1):
            _mostly_equal(O.port.refine_barycentric_2d(mf, cont, sym, fl, fr2, idx2, search),
                          O.ref_ibr.refine_barycentric_2d(mf, cont, sym, fl, fr2, idx2, search), tol=1e-4, frac=0.99)


# ---- on-demand cost volumes and PatchMatch (on_demand_cost_volume.h, patchmatch.h) -----------------------------------------------

needs_ref_pm = pytest.mark.skipif(O.ref_ibr is None or not O.ref_ibr.has("patch_match"), reason="oracle/_ref/libstevi_ref_ibr.so was not built")


@needs_ref_pm
@pytest.mark.parametrize("mf", [O.ZNCC, O.SSD, O.ZSAD])
def test_port_on_demand_and_patch_match_equal_the_compiled_reference(mf):
    """The C restatement against the reference's own classes: the on-demand truncated volume (incl. cells without a value, a
    search range that does not start at 0) and the deterministic part of PatchMatch (no random search, given initial map; 1-D and
    2-D searches, all four scan directions) bit for bit."""
    L, R, _ = synth.stereo_pair(40, 96, 3, d_bg=3, d_sq=8, d_max=0, dtype=np.float32)
    fs, ft = port.unfold(R, 2, 1), port.unfold(L, 2, 1)
    rng = np.random.default_rng(int(mf))
    for search in ((0, 11), (-2, 9), (3, 12)):
        d = rng.integers(search[0] - 1, search[1] + 2, fs.shape[:2]).astype(np.int32)
        for radius in (1, 2):
            assert same(port.on_demand_truncated_cost_volume(mf, fs, ft, search, d, radius), O.ref_ibr.on_demand_truncated_cost_volume(mf, fs, ft, search, d, radius))
    init = rng.integers(-1, 14, fs.shape[:2] + (1,)).astype(np.int32)
    for n_iter in (1, 2, 3, 4, 6):
        assert same(port.patch_match(mf, fs, ft, (0, 11), n_iter, 0, init), O.ref_ibr.patch_match(mf, fs, ft, (0, 11), n_iter, 0, init))
    ft2 = port.unfold(np.roll(L, 1, axis=0), 2, 1)
    init2 = np.stack([rng.integers(-2, 3, fs.shape[:2]), rng.integers(0, 10, fs.shape[:2])], axis=2).astype(np.int32)
    d2 = np.stack([rng.integers(-3, 4, fs.shape[:2]), rng.integers(-1, 11, fs.shape[:2])], axis=2).astype(np.int32)
    search2 = (-2, 2, 0, 9)
    assert same(port.on_demand_truncated_cost_volume(mf, fs, ft2, search2, d2, 1), O.ref_ibr.on_demand_truncated_cost_volume(mf, fs, ft2, search2, d2, 1))
    assert same(port.patch_match(mf, fs, ft2, search2, 5, 0, init2), O.ref_ibr.patch_match(mf, fs, ft2, search2, 5, 0, init2))


def test_on_demand_cells_equal_the_full_cost_volume():
    """testOnDemandCostVolume.cpp:62-112: every cell the on-demand volume gives equals the cell of the full cost volume."""
    L, R, _ = synth.stereo_pair(24, 64, 5, d_bg=2, d_sq=6, d_max=0, dtype=np.float32)
    fs, ft = port.unfold(R, 1, 1), port.unfold(L, 1, 1)
    for mf in (O.ZNCC, O.SSD, O.SAD):
        full = port.feature_volume_to_cost_volume(mf, ft, fs, 10, O.R2L)
        for d in range(10):
            disp = np.full(fs.shape[:2], d, np.int32)
            cells = port.on_demand_truncated_cost_volume(mf, fs, ft, (0, 9), disp, 0)[..., 0]
            ok = np.arange(64)[None, :] + d < 64
            assert np.array_equal(cells[np.broadcast_to(ok, cells.shape)], full[..., d][np.broadcast_to(ok, cells.shape)])


def test_golden_refinement_patch_match_on_demand():
    """The C restatement against outputs of the reference's own code stored in tests/golden/reference_outputs_refinement.npz
    (generator: tests/golden/make_golden.py refinement): runs where neither /root/reference nor oracle/_ref exist."""
    G3 = np.load(os.path.join(os.path.dirname(__file__), "golden", "reference_outputs_refinement.npz"))
    left, right = G3["left"], G3["right"]
    fl_raw, fr_raw = port.unfold(left, 2, 1), port.unfold(right, 2, 1)
    inner = (slice(1, -1), slice(4, -14))
    for name in ("NCC", "ZNCC", "SSD", "ZSSD", "SAD", "ZSAD"):
        mf = getattr(O, name)
        fl, fr = port.features_for_matchfunc(mf, fl_raw), port.features_for_matchfunc(mf, fr_raw)
        cv = port.aggregate_cost(mf, fl, fr, 0, 9, O.R2L)
        idx = port.wta(cv, O.strategy_of(mf))
        assert np.array_equal(idx, G3[f"idx_{name}"])
        b = port.refine_barycentric(mf, fl, fr, idx)
        _mostly_equal(b[inner], G3[f"bary_{name}"][inner]) if name in ("SAD", "ZSAD") else _mostly_equal(b, G3[f"bary_{name}"])
        _mostly_equal(port.refine_cost_symmetric(mf, fl, fr, idx, cv), G3[f"costsym_{name}"])
        if name not in ("SAD", "ZSAD"):
            for r in (1, 2):
                _mostly_equal(port.refine_barycentric_symmetric(mf, r, fl, fr, idx, 10), G3[f"barysym{r}_{name}"])
    for name in ("ZNCC", "SSD", "ZSAD"):
        mf = getattr(O, name)
        for n_iter in (1, 2, 4):
            assert same(port.patch_match(mf, fr_raw, fl_raw, (0, 9), n_iter, 0, G3["pm_init"]), G3[f"pm_{name}_{n_iter}"])
        assert same(port.on_demand_truncated_cost_volume(mf, fr_raw, fl_raw, (-2, 9), G3["od_disp"], 1), G3[f"od_{name}"])


def test_patch_refinement_port_matches_compiled_reference(oracle):
    """refineDisp2dCostPatchInterpolation (cost_based_refinement.h:371-433): the port's closed-form normal equations against the reference's
    refineCostPatch (:71-126) compiled over the Eigen stand-in (float Gauss-Jordan inverses). Same kept/dropped pixels, deltas within 1e-4."""
    if oracle.ref_ibr is None:
        pytest.skip("oracle/_ref/libstevi_ref_ibr.so not built")
    rng = np.random.default_rng(12)
    t = rng.uniform(0.5, 2.0, (30, 90, 5, 3)).astype(np.float32)
    t[:, :, 2, 1] = 0.1
    t[::7, ::5, 2, 1] = 3.0             # centre is not the extremum: deltas beyond one pixel are dropped
    raw = rng.integers(-5, 6, (30, 90, 2)).astype(np.int32)
    for kernel in (oracle.PARABOLA, oracle.GAUSSIAN):
        a = oracle.port.refine_2d_patch(t, raw, kernel)
        b = oracle.ref_ibr.refine_2d_patch(t, raw, kernel)
        assert np.array_equal((a != raw).all(axis=2), (b != raw).all(axis=2))
        assert np.max(np.abs(a - b)) <= 1e-4
    for shape in ((3, 3, 4, 3), (3, 3, 3, 1)):
        assert oracle.ref_ibr.refine_2d_patch(np.zeros(shape, np.float32), np.zeros((3, 3, 2), np.int32), oracle.PARABOLA) is None
