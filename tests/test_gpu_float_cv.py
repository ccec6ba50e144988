"""GPU parity of the float cost volumes (CC/NCC/ZCC/ZNCC/SSD/ZSSD/SAD/ZSAD) against the CPU oracle.

Tolerances: the correlations are compared absolutely (hard gate 1e-4 on values in [-1,1] for the
normalised ones; the un-normalised ones scaled by the largest magnitude), the costs relatively to the largest cost.
The box-sum path reorders the summation, the direct path (`exact_sums`) keeps the reference's channel order.
"""
import numpy as np
import pytest

from libstevi_b200 import synth

pytestmark = pytest.mark.gpu

MFS = ["CC", "NCC", "ZCC", "ZNCC", "SSD", "ZSSD", "SAD", "ZSAD"]


def images(h, wl, wr, seed):
    return synth.random_image(h, wl, seed), synth.random_image(h, wr, seed + 100)


def check(api, oracle, name, got, want, exact):
    assert got.shape == want.shape and got.dtype == np.float32
    assert np.array_equal(np.isnan(got), np.isnan(want))
    scale = 1.0 if name in ("NCC", "ZNCC") else max(1.0, float(np.nanmax(np.abs(want))))
    err = float(np.nanmax(np.abs(got - want))) / scale
    assert err <= (2e-6 if exact else 1e-4 if name in ("NCC", "ZNCC") else 1e-5), (name, err)


@pytest.mark.parametrize("name", MFS)
@pytest.mark.parametrize("hr,vr", [(2, 2), (4, 4), (3, 1), (1, 3)])
@pytest.mark.parametrize("direction", [0, 1])
def test_float_cost_volume_box(api, ctx, oracle, name, hr, vr, direction):
    L, R = images(41, 150, 150, 5)
    mf = getattr(api, name)
    got = api.unfoldBasedCostVolume(mf, L, R, hr, vr, 23, dDir=direction, ctx=ctx)
    want = oracle.port.cost_volume(getattr(oracle, name), L, R, hr, vr, 23, direction)
    check(api, oracle, name, got, want, exact=False)


@pytest.mark.parametrize("name", MFS)
@pytest.mark.parametrize("hr,vr", [(2, 2), (5, 3), (0, 2), (7, 0)])
def test_float_cost_volume_direct(api, ctx, oracle, name, hr, vr):
    L, R = images(23, 70, 70, 7)
    mf = getattr(api, name)
    for direction in (0, 1):
        got = api.unfoldBasedCostVolume(mf, L, R, hr, vr, 11, dDir=direction, exact_sums=True, ctx=ctx)
        want = oracle.port.cost_volume(getattr(oracle, name), L, R, hr, vr, 11, direction)
        check(api, oracle, name, got, want, exact=True)


@pytest.mark.parametrize("layout", ["ref", "c", "f"])
@pytest.mark.parametrize("w,D", [(128, 64), (131, 67), (260, 5), (9, 3)])
def test_float_cost_volume_layouts_and_ragged(api, ctx, oracle, layout, w, D):
    L, R = images(19, w, w, 11)
    for name in ("ZNCC", "SSD", "SAD"):
        got = api.unfoldBasedCostVolume(getattr(api, name), L, R, 3, 2, D, layout=layout, ctx=ctx)
        want = oracle.port.cost_volume(getattr(oracle, name), L, R, 3, 2, D)
        check(api, oracle, name, got, want, exact=False)


def test_float_cost_volume_different_widths_and_heights(api, ctx, oracle):
    L, _ = images(20, 90, 90, 1)
    _, R = images(20, 64, 64, 2)
    for direction in (0, 1):
        for name in ("ZNCC", "ZSSD"):
            got = api.unfoldBasedCostVolume(getattr(api, name), L, R, 2, 2, 30, dDir=direction, ctx=ctx)
            want = oracle.port.cost_volume(getattr(oracle, name), L, R, 2, 2, 30, direction)
            check(api, oracle, name, got, want, exact=False)
    L2, _ = images(21, 90, 90, 1)
    assert api.unfoldBasedCostVolume(api.ZNCC, L2, R, 2, 2, 8, ctx=ctx) is None  # cross_correlations.h:751-759


def test_float_cost_volume_many_row_bands(api, ctx, oracle):
    # several bands of the running vertical sums, D not a multiple of the 64-wide disparity block
    L, R = images(150, 136, 136, 21)
    for name in ("ZNCC", "ZSSD", "SAD"):
        got = api.unfoldBasedCostVolume(getattr(api, name), L, R, 4, 4, 70, ctx=ctx)
        want = oracle.port.cost_volume(getattr(oracle, name), L, R, 4, 4, 70)
        check(api, oracle, name, got, want, exact=False)


def test_zncc_flat_windows_give_nan_like_the_reference(api, ctx, oracle):
    L, R = images(30, 80, 80, 3)
    L[10:22, 20:40] = 0.25  # flat area larger than the window: norm == 0 -> NaN features (Q6)
    R[5:9, 50:60] = 0.5     # smaller than the window: finite
    for exact in (False, True):
        got = api.unfoldBasedCostVolume(api.ZNCC, L, R, 2, 2, 16, exact_sums=exact, ctx=ctx)
        want = oracle.port.cost_volume(oracle.ZNCC, L, R, 2, 2, 16)
        assert np.isnan(want).any()
        check(api, oracle, "ZNCC", got, want, exact)


def test_u8_images_for_unnormalised_functions(api, ctx, oracle):
    L, R, _ = synth.stereo_pair(33, 140, 9, d_bg=3, d_sq=9)
    for name in ("SSD", "SAD", "CC"):
        got = api.unfoldBasedCostVolume(getattr(api, name), L, R, 3, 3, 20, ctx=ctx)
        want = oracle.port.cost_volume(getattr(oracle, name), L.astype(np.float32), R.astype(np.float32), 3, 3, 20)
        assert np.array_equal(got, want)  # integer-valued sums below 2^24 are exact in any order
    with pytest.raises(api.SteviError):
        api.unfoldBasedCostVolume(api.ZNCC, L, R, 3, 3, 20, ctx=ctx)


def test_zncc_wta_recovers_ground_truth(api, ctx):
    L, R, gt = synth.stereo_pair(96, 200, 4, d_bg=5, d_sq=17, dtype=np.float32)
    cv = api.unfoldBasedCostVolume(api.ZNCC, L, R, 2, 2, 32, ctx=ctx)
    idx = api.extractSelectedIndex(cv, api.SCORE, ctx=ctx)
    inner = (slice(6, -6), slice(6, -40))
    assert (idx[inner] == gt[inner]).mean() > 0.95  # occluded columns next to the square cannot match
