"""GPU parity of the materialised feature volumes: unfold (2-D / 3-D images, padding,
orientation), the UnFoldCompressor unfold, censusFeatures and getFeatureVolumeForMatchFunc against the CPU oracle."""
import numpy as np
import pytest

from libstevi_b200 import synth

pytestmark = pytest.mark.gpu

MASK = np.array([[1, 1, 2, 2, 2],
                 [1, 3, 3, 3, 2],
                 [0, 3, 7, 3, 0],
                 [5, 3, 3, 3, 4],
                 [5, 5, 0, 4, 4]], np.int32)


def images():
    img2 = synth.random_image(23, 37, 21)
    img3 = np.stack([synth.random_image(23, 37, 22 + k) for k in range(3)], -1)
    return img2, img3


@pytest.mark.parametrize("orientation", [0, 1, 2, 3])
def test_unfold_orientations(api, ctx, oracle, orientation):
    for img in images():
        got = api.unfold(2, 1, img, orientation=orientation, ctx=ctx)
        assert np.array_equal(got, oracle.port.unfold_general(img, 2, 1, None, orientation))


@pytest.mark.parametrize("pad", [(0, 0, 0, 0), (3, 1, 0, 2), (4, 4, 4, 4)])
def test_unfold_padding(api, ctx, oracle, pad):
    for img in images():
        got = api.unfold(1, 2, img, padding=pad, ctx=ctx)
        want = oracle.port.unfold_general(img, 1, 2, pad, 0)
        assert got.shape == want.shape and np.array_equal(got, want)


def test_unfold_u8_and_known_answer(api, ctx, oracle):
    img = synth.random_image(9, 12, 5, dtype=np.uint8)
    got = api.unfold(2, 1, img, ctx=ctx)
    assert got.dtype == np.uint8 and np.array_equal(got, oracle.port.unfold(img, 2, 1))
    assert np.array_equal(api.unfold(2, 1, img, T_O=np.float32, ctx=ctx), oracle.port.unfold(img, 2, 1).astype(np.float32))
    # testCorrelationFilters.cpp:384-445: the centre pixel's features are the image in window order
    small = synth.random_image(3, 5, 9)
    fv = api.unfold(2, 1, small, ctx=ctx)
    assert fv.shape == (3, 5, 15) and np.array_equal(fv[1, 2], small.ravel())


@pytest.mark.parametrize("pad", [None, (0, 0, 0, 0), (1, 2, 3, 0)])
def test_unfold_compressed(api, ctx, oracle, pad):
    for img in images():
        for mask in (MASK, MASK[:3, 1:4]):
            got = api.unfoldCompressed(mask, img, padding=pad, ctx=ctx)
            want = oracle.port.unfold_compressed(mask, img, pad)
            assert got.shape == want.shape and np.allclose(got, want, atol=1e-6)
    u8 = synth.random_image(13, 17, 2, dtype=np.uint8)
    got = api.unfoldCompressed(MASK, u8, ctx=ctx)
    assert np.allclose(got, oracle.port.unfold_compressed(MASK, u8.astype(np.float32)), atol=1e-4)


def test_unfold_compressor_known_answer(api, ctx):
    # testCorrelationFilters.cpp:514-588 idea: every super-pixel feature is the mean of its pixels
    img = synth.random_image(5, 5, 3)
    got = api.unfoldCompressed(MASK, img, padding=(0, 0, 0, 0), ctx=ctx)
    assert got.shape == (1, 1, 6)
    for f, label in enumerate(sorted(set(MASK[MASK > 0].tolist()))):
        assert abs(got[0, 0, f] - img[MASK == label].mean()) < 1e-6


@pytest.mark.parametrize("hr,vr", [(4, 3), (3, 3), (2, 2), (1, 1)])
def test_census_features(api, ctx, oracle, hr, vr):
    img = synth.random_image(19, 27, 8)
    fv = oracle.port.unfold_general(img, hr, vr)
    got = api.censusFeatures(fv, ctx=ctx)
    assert np.array_equal(got, oracle.port.census_features(fv))
    assert np.array_equal(got, api.censusTransform2D(img, hr, vr, ctx=ctx))
    u8 = synth.random_image(19, 27, 8, dtype=np.uint8)
    assert np.array_equal(api.censusFeatures(oracle.port.unfold(u8, hr, vr), ctx=ctx), oracle.port.census(u8, hr, vr))


@pytest.mark.parametrize("name", ["CC", "NCC", "ZCC", "ZNCC", "SSD", "ZSSD", "SAD", "ZSAD"])
def test_feature_volume_for_matchfunc(api, ctx, oracle, name):
    img = synth.random_image(17, 29, 12)
    img[4:10, 5:12] = 0.5  # a flat area: NaN features for the normalised functions (Q6)
    got = api.getFeatureVolumeForMatchFunc(getattr(api, name), img, 2, 1, ctx=ctx)
    want = oracle.port.feature_volume(getattr(oracle, name), img, 2, 1)
    assert np.array_equal(np.isnan(got), np.isnan(want))
    assert np.nanmax(np.abs(got - want)) <= 1e-6
