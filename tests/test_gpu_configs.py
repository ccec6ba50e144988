"""The five BASELINE.json configurations on the GPU against the CPU oracle at (or near) full size, and one check per
SURVEY.md §8(a) row against the COMPILED REFERENCE itself (oracle.ref = the unmodified LibStevi headers, oracle/_ref) when
it is present — it is on the GPU box: oracle/_ref travels with the snapshot.

* C1  640x480, ZNCC 5x5, D=64: unfoldBasedCostVolume -> extractSelectedIndex<Score> -> selectedIndexToDisp, the calls of
      test/benchmarks/benchmarkCrossCorrelationAlgorithms.cpp:93-95, full frame against oracle.ref.
* C2  census 9x7 + SGM<8> + WTA, 1920x1080, D=256: the WHOLE frame — cost volume, SGM volume and disparity map bit for bit
      (the oracle port's SGM is O(D) per pixel and OpenMP: seconds); plus a 48-row full-width band through the real recurrence
      (Score strategy; CORRECTED Cost).
* C3  ZNCC 9x9 + cost refinement, 1920 wide, D=192: 96-row band, volume 1e-4, winners, refined disparities, NaN pattern.
* C4  coarse-to-fine ZNCC, depth 3, on a 512x768 crop: disp_estimate equal on >= 99.9 % of the pixels.
* C5  1280x720, D=128 census+SGM: the full frame, bit for bit.
"""
import numpy as np
import pytest

from libstevi_b200 import synth

pytestmark = pytest.mark.gpu


def _need_ref(oracle):
    if oracle.ref is None:
        pytest.skip("oracle/_ref (compiled reference) was not built")
    return oracle.ref


# ---- C1 -----------------------------------------------------------------------------------------------------------------

def test_c1_full_frame_against_the_compiled_reference(api, ctx, oracle):
    ref = _need_ref(oracle)
    H, W, D = 480, 640, 64
    L, R, gt = synth.stereo_pair(H, W, 1, d_bg=8, d_sq=32, d_max=0, dtype=np.float32)
    cv = api.unfoldBasedCostVolume(api.ZNCC, L, R, 2, 2, D, layout="c", ctx=ctx)
    want = ref.cost_volume(oracle.ZNCC, L, R, 2, 2, D)
    assert np.array_equal(np.isnan(cv), np.isnan(want))
    assert np.nanmax(np.abs(cv - want)) <= 1e-4          # normalised scores, absolute
    idx = api.extractSelectedIndex(want, api.SCORE, ctx=ctx)
    assert np.array_equal(idx, ref.wta(want, oracle.SCORE))   # same volume in: the winners are exact (Q5)
    disp = api.selectedIndexToDisp(idx, ctx=ctx)
    assert np.array_equal(disp, ref.index_to_disp(idx))
    # winners of OUR volume differ from the reference's only where its best and runner-up are closer than the tolerance
    mine = api.extractSelectedIndex(cv, api.SCORE, ctx=ctx)
    srt = np.sort(np.nan_to_num(want, nan=-2.0), axis=2)
    clear = (srt[:, :, -1] - srt[:, :, -2]) > 2e-4
    assert np.array_equal(mine[clear], idx[clear])
    inner = (slice(8, -8), slice(8, W - D))
    assert (mine[inner] == gt[inner]).mean() > 0.95


# ---- C2 / C5 --------------------------------------------------------------------------------------------------------------

def _census_sgm_case(api, ctx, oracle, L, R, D):
    cv, sg, disp = api.censusSgmPipeline(L, R, 4, 3, D, 10, 120, 100, 8, return_volumes=True, ctx=ctx)
    ocv = oracle.port.cost_volume(oracle.CENSUS, L, R, 4, 3, D)
    assert np.array_equal(cv, ocv)
    del cv
    osg = oracle.port.sgm(ocv, 8, oracle.COST, 10, 120, 100)
    del ocv
    assert np.array_equal(sg, osg)
    del sg
    odisp = oracle.port.index_to_disp(oracle.port.wta(osg, oracle.COST))
    assert np.array_equal(disp, odisp)
    # the disparity-only call (no volume in memory) gives the same map
    assert np.array_equal(api.censusSgmPipeline(L, R, 4, 3, D, 10, 120, 100, 8, ctx=ctx), odisp)
    return odisp


def test_c2_full_frame_cv_sgm_wta_against_the_oracle(api, ctx, oracle):
    H, W, D = 1080, 1920, 256
    L, R, gt = synth.stereo_pair(H, W, 2, d_bg=8, d_sq=32, d_max=200)
    disp = _census_sgm_case(api, ctx, oracle, L, R, D)   # the whole frame: 531 M cells of CV and of SGM volume, bit for bit
    inner = (slice(16, -16), slice(16, W - D))
    assert (np.abs(disp[inner] - gt[inner]) <= 1).mean() > 0.5


def test_c2_general_recurrence_band_against_the_oracle(api, ctx, oracle):
    # the same band through the real recurrence (Score strategy on the census costs, and CORRECTED Cost): sgm_sweep.cu at full width
    H, W, D = 1080, 1920, 256
    L, R, _ = synth.stereo_pair(H, W, 2, d_bg=8, d_sq=32, d_max=200)
    i0, n = 500, 48
    cv = api.unfoldBasedCostVolume(api.CENSUS, np.ascontiguousarray(L[i0:i0 + n]), np.ascontiguousarray(R[i0:i0 + n]), 4, 3, D, ctx=ctx)
    got = api.sgmCostVolume(cv, 10, 120, Pout=100, nDirections=8, strategy=api.SCORE, ctx=ctx)
    assert np.array_equal(got, oracle.port.sgm(np.ascontiguousarray(cv), 8, oracle.SCORE, 10, 120, 100))
    got = api.sgmCostVolume(cv, 10, 120, Pout=100, nDirections=8, strategy=api.COST, semantics=api.CORRECTED, ctx=ctx)
    assert np.array_equal(got, oracle.port.sgm_corrected(np.ascontiguousarray(cv), 8, oracle.COST, 10, 120, 100))


@pytest.mark.parametrize("seed", [1000, 1001])
def test_c5_full_frame_against_the_oracle(api, ctx, oracle, seed):
    H, W, D = 720, 1280, 128
    L, R, gt = synth.stereo_pair(H, W, seed, d_bg=8, d_sq=32, d_max=100)
    disp = _census_sgm_case(api, ctx, oracle, L, R, D)
    inner = (slice(16, -16), slice(16, W - D))
    assert (np.abs(disp[inner] - gt[inner]) <= 1).mean() > 0.5


# ---- C3 -----------------------------------------------------------------------------------------------------------------

def test_c3_band_volume_winners_refinement_against_the_oracle(api, ctx, oracle):
    H, W, D, hr, vr = 1080, 1920, 192, 4, 4
    L, R, _ = synth.stereo_pair(H, W, 3, d_bg=8, d_sq=32, d_max=120, dtype=np.float32)
    i0, n = 470, 96
    Lb, Rb = np.ascontiguousarray(L[i0:i0 + n]), np.ascontiguousarray(R[i0:i0 + n])
    Lb[40:52, 300:420] = 0.25   # a flat patch: NaN scores (Q6) and NaN refined disparities
    cv, idx, tcv, refined = api.localMatchingPipeline(api.ZNCC, Lb, Rb, hr, vr, D, api.PARABOLA, return_all=True, layout="c", ctx=ctx)
    ocv = oracle.port.cost_volume(oracle.ZNCC, Lb, Rb, hr, vr, D)
    assert np.array_equal(np.isnan(cv), np.isnan(ocv))
    assert np.isnan(ocv).any()
    assert np.nanmax(np.abs(cv - ocv)) <= 1e-4
    oidx = oracle.port.wta(ocv, oracle.SCORE)
    srt = np.sort(np.nan_to_num(ocv, nan=-2.0), axis=2)
    clear = (srt[:, :, -1] - srt[:, :, -2]) > 2e-4
    assert np.array_equal(idx[clear], oidx[clear])
    # the refinement stage on the oracle's own volume and winners: NaN pattern identical, values to 1e-3 px where the parabola is
    # well conditioned
    t2 = api.truncatedCostVolume(ocv, oidx, hr, vr, 1, ctx=ctx)
    otcv = oracle.port.truncated_cost_volume(ocv, oidx, hr, vr, 1)
    assert np.array_equal(t2, otcv, equal_nan=True)
    r2 = api.refineDispCostInterpolation(otcv, oidx, api.PARABOLA, ctx=ctx)
    oref = oracle.port.refine(otcv, oidx, oracle.PARABOLA)
    assert np.array_equal(np.isnan(r2), np.isnan(oref))
    curv = np.abs(otcv[:, :, 2] - 2 * otcv[:, :, 1] + otcv[:, :, 0])
    ok = np.isfinite(oref) & (curv > 1e-3)
    assert np.abs(r2[ok] - oref[ok]).max() <= 1e-3
    # ... and end to end where the winners agree
    same = (idx == oidx) & ok
    assert np.array_equal(np.isnan(refined[idx == oidx]), np.isnan(oref[idx == oidx]))
    assert np.abs(refined[same] - oref[same]).max() <= 1e-3


# ---- C4 -----------------------------------------------------------------------------------------------------------------

def test_c4_crop_disp_estimate_against_the_oracle(api, ctx, oracle):
    H, W, D, r, depth = 512, 768, 128, 3, 3
    L, R, gt = synth.stereo_pair(H, W, 4, d_bg=12, d_sq=40, d_max=0, dtype=np.float32)
    tcv, disp = api.hiearchicalTruncatedCostVolume(api.ZNCC, depth, L, R, r, r, D, 2, ctx=ctx)
    otcv, odisp = oracle.port.hierarchical(oracle.ZNCC, depth, L, R, r, r, D, 2)
    assert (disp == odisp).mean() >= 0.999
    same = disp == odisp
    both = np.isfinite(tcv) & np.isfinite(otcv) & same[:, :, None]
    assert np.abs(tcv[both] - otcv[both]).max() <= 1e-4
    side = min(H, W) // 3
    i0, j0 = (H - side) // 2, (W - side) // 2
    inner = (slice(i0 + 16, i0 + side - 16), slice(j0 + 16, j0 + side - 16))
    assert (disp[inner] == gt[inner]).mean() >= 0.95


# ---- one check per §8(a) row against the compiled reference ---------------------------------------------------------------------

def test_rows_against_the_compiled_reference(api, ctx, oracle):
    ref = _need_ref(oracle)
    rng = np.random.default_rng(11)
    img = rng.uniform(0, 1, (37, 52)).astype(np.float32)
    img8 = rng.integers(0, 256, (37, 52)).astype(np.uint8)
    L, R, _ = synth.stereo_pair(40, 72, 5, d_bg=3, d_sq=9, d_max=12)
    Lf, Rf = L.astype(np.float32) / np.float32(255), R.astype(np.float32) / np.float32(255)
    # a1 unfold, a3 census, a6 feature volume
    assert np.array_equal(api.unfold(2, 1, img, ctx=ctx), ref.unfold(img, 2, 1))
    assert np.array_equal(api.unfold(3, 2, img8, ctx=ctx), ref.unfold(img8, 3, 2))
    assert np.array_equal(api.censusTransform2D(img8, 4, 3, ctx=ctx), ref.census(img8, 4, 3))
    fv = api.getFeatureVolumeForMatchFunc(api.ZNCC, img, 2, 2, ctx=ctx)
    assert np.nanmax(np.abs(fv - ref.feature_volume(oracle.ZNCC, img, 2, 2))) <= 1e-5
    # a4/a5/a7 cost volumes: census bit-exact; float functions to tolerance
    D = 20
    cv = api.unfoldBasedCostVolume(api.CENSUS, L, R, 4, 3, D, layout="c", ctx=ctx)
    rcv = ref.cost_volume(oracle.CENSUS, L, R, 4, 3, D)
    assert np.array_equal(cv, rcv)
    for mf, omf in ((api.ZNCC, oracle.ZNCC), (api.SSD, oracle.SSD), (api.SAD, oracle.SAD), (api.NCC, oracle.NCC)):
        got = api.unfoldBasedCostVolume(mf, Lf, Rf, 2, 2, D, layout="c", ctx=ctx)
        want = ref.cost_volume(omf, Lf, Rf, 2, 2, D)
        assert np.nanmax(np.abs(got - want)) <= 1e-4 * max(1.0, float(np.nanmax(np.abs(want))))
    # u8 images through the reference's own u8 instantiation (not float-converted ones)
    for mf, omf in ((api.SSD, oracle.SSD), (api.SAD, oracle.SAD), (api.CC, oracle.CC)):
        got = api.unfoldBasedCostVolume(mf, L, R, 2, 2, D, layout="c", ctx=ctx)
        want = ref.cost_volume(omf, L, R, 2, 2, D)
        assert np.nanmax(np.abs(got - want)) <= 1e-5 * max(1.0, float(np.nanmax(np.abs(want))))
    # a8 SGM (compat collapse and the real recurrence), a9 WTA, a10, a11
    for ndir in (4, 8):
        assert np.array_equal(api.sgmCostVolume(cv, 10, 120, Pout=100, nDirections=ndir, ctx=ctx), ref.sgm(rcv, ndir, oracle.COST, 10, 120, 100))
        assert np.array_equal(api.sgmCostVolume(cv, 3, 20, Pout=50, nDirections=ndir, strategy=api.SCORE, ctx=ctx),
                              ref.sgm(rcv, ndir, oracle.SCORE, 3, 20, 50))
    zcv = ref.cost_volume(oracle.ZNCC, Lf, Rf, 2, 2, D)
    idx = api.extractSelectedIndex(zcv, api.SCORE, ctx=ctx)
    ridx = ref.wta(zcv, oracle.SCORE)
    assert np.array_equal(idx, ridx)
    assert np.array_equal(api.extractSelectedIndex(rcv, api.COST, ctx=ctx), ref.wta(rcv, oracle.COST))
    assert np.array_equal(api.selectedCost(zcv, ridx, ctx=ctx), ref.selected_cost(zcv, ridx))
    t = api.truncatedCostVolume(zcv, ridx, 2, 2, 1, ctx=ctx)
    rt = ref.truncated_cost_volume(zcv, ridx, 2, 2, 1)
    assert np.array_equal(t, rt, equal_nan=True)
    for k, ok in ((api.PARABOLA, oracle.PARABOLA), (api.EQUIANGULAR, oracle.EQUIANGULAR)):
        a, b = api.refineDispCostInterpolation(rt, ridx, k, ctx=ctx), ref.refine(rt, ridx, ok)
        assert np.array_equal(np.isnan(a), np.isnan(b))
        fin = np.isfinite(b)
        assert np.abs(a[fin] - b[fin]).max() <= 1e-5
    # a12 pooling, a13 guided volume / coarse-to-fine driver
    assert np.allclose(api.averagePoolingDownsample(img, ctx=ctx), ref.avgpool2(img), atol=1e-6)
    tcv, disp = api.hiearchicalTruncatedCostVolume(api.ZNCC, 2, Lf, Rf, 2, 2, D, 2, ctx=ctx)
    rtcv, rdisp = ref.hierarchical(oracle.ZNCC, 2, Lf, Rf, 2, 2, D, 2)
    assert (disp == rdisp).mean() >= 0.995
