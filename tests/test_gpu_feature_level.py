"""Feature-level entry points (caller-supplied float feature volumes): getFeatureVolumeForMatchFunc on
a ready-made volume, aggregateCost incl. its searchOffset<1> overload, featureVolume2CostVolume, the UnFoldCompressor overload of
unfoldBasedCostVolume and multi-channel (nImDim = 3) images — against the oracle port and, when present, the compiled reference
(cross_correlations.h:194-308, :645-791). Tolerances of DESIGN.md §2: the sums run in the reference's channel order with
separately rounded operations (2e-6 relative); normalised features 1e-5."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

FUNCS = ["CC", "NCC", "ZCC", "ZNCC", "SSD", "ZSSD", "SAD", "ZSAD"]


def checkers(oracle):
    return [oracle.port] + ([oracle.ref] if oracle.ref is not None else [])


def close(a, b, rel=2e-6):
    assert a.shape == b.shape
    assert np.array_equal(np.isnan(a), np.isnan(b))
    scale = max(1.0, float(np.nanmax(np.abs(b)))) if np.isfinite(b).any() else 1.0
    assert np.nanmax(np.abs(a - b)) <= rel * scale if np.isfinite(b).any() else True


@pytest.mark.parametrize("name", FUNCS)
def test_features_for_matchfunc(api, ctx, oracle, name):
    rng = np.random.default_rng(3)
    fv = rng.uniform(-1, 1, (13, 29, 17)).astype(np.float32)
    fv[4, 5, :] = 0.5   # a constant vector: NaN for the normalised zero-mean functions (Q6)
    got = api.getFeatureVolumeForMatchFuncFromFeatures(getattr(api, name), fv, ctx=ctx)
    for chk in checkers(oracle):
        close(got, chk.features_for_matchfunc(getattr(oracle, name), fv), 1e-5)


@pytest.mark.parametrize("name", FUNCS)
@pytest.mark.parametrize("ddir", ["r2l", "l2r"])
def test_aggregate_cost_and_search_offsets(api, ctx, oracle, name, ddir):
    rng = np.random.default_rng(5)
    fl = rng.uniform(-1, 1, (9, 75, 21)).astype(np.float32)
    fr = rng.uniform(-1, 1, (9, 70, 21)).astype(np.float32)   # widths may differ, only heights must agree (:209-211)
    d_api, d_o = (api.RIGHT_TO_LEFT, oracle.R2L) if ddir == "r2l" else (api.LEFT_TO_RIGHT, oracle.L2R)
    for rng_ in ((0, 39), (-3, 5), (4, 4), (-40, -1)):
        got = api.aggregateCost(getattr(api, name), fl, fr, searchRange=rng_, dDir=d_api, layout="c", ctx=ctx)
        for chk in checkers(oracle):
            close(got, chk.aggregate_cost(getattr(oracle, name), fl, fr, rng_[0], rng_[1], d_o))
    got = api.aggregateCost(getattr(api, name), fl, fr, 12, dDir=d_api, ctx=ctx)   # the disp_width overload, reference layout
    close(np.ascontiguousarray(got), oracle.port.aggregate_cost(getattr(oracle, name), fl, fr, 0, 11, d_o))
    assert api.aggregateCost(getattr(api, name), fl[:5], fr, 8, dDir=d_api, ctx=ctx) is None   # heights differ: empty array


@pytest.mark.parametrize("name", FUNCS)
def test_feature_volume_to_cost_volume(api, ctx, oracle, name):
    rng = np.random.default_rng(7)
    fl = rng.uniform(0, 1, (11, 48, 25)).astype(np.float32)
    fr = rng.uniform(0, 1, (11, 48, 25)).astype(np.float32)
    got = api.featureVolume2CostVolume(getattr(api, name), fl, fr, 16, layout="c", ctx=ctx)
    for chk in checkers(oracle):
        close(got, chk.feature_volume_to_cost_volume(getattr(oracle, name), fl, fr, 16), 1e-5)


def test_feature_volume_to_cost_volume_many_features_and_torch(api, ctx, oracle):
    torch = pytest.importorskip("torch")
    rng = np.random.default_rng(8)
    fl = rng.uniform(0, 1, (6, 70, 147)).astype(np.float32)   # 7x7 windows x 3 channels (benchmarkStereoMatchingModels.cpp:417-418)
    fr = rng.uniform(0, 1, (6, 70, 147)).astype(np.float32)
    want = oracle.port.feature_volume_to_cost_volume(oracle.ZNCC, fl, fr, 40)
    close(api.featureVolume2CostVolume(api.ZNCC, fl, fr, 40, layout="c", ctx=ctx), want, 1e-5)
    tctx = api.Context(0, torch_stream=True)
    got = api.featureVolume2CostVolume(api.ZNCC, torch.from_numpy(fl).cuda(), torch.from_numpy(fr).cuda(), 40, ctx=tctx)
    assert got.stride(1) == 1   # the reference's layout {W*D, 1, W}
    close(got.cpu().numpy(), want, 1e-5)
    tctx.close()


def test_census_on_feature_volumes(api, ctx, oracle):
    rng = np.random.default_rng(9)
    fl = rng.integers(0, 256, (10, 40, 65)).astype(np.float32)
    fr = rng.integers(0, 256, (10, 40, 65)).astype(np.float32)
    got = api.featureVolume2CostVolume(api.CENSUS, fl, fr, 9, layout="c", ctx=ctx)
    cl, cr = oracle.port.census_features(fl), oracle.port.census_features(fr)
    want = api.hammingCostVolume(cl, cr, 9, layout="c", ctx=ctx)
    assert got.dtype == np.uint16 and np.array_equal(got, want)


GRPIX_LIKE = np.array([[1, 1, 2, 2, 3], [1, 4, 4, 5, 3], [6, 4, 0, 5, 7], [6, 8, 8, 9, 7], [10, 10, 11, 9, 9]], np.int32)


@pytest.mark.parametrize("name", ["ZNCC", "SSD", "ZSAD", "NCC"])
@pytest.mark.parametrize("channels", [0, 3])
def test_compressor_overload(api, ctx, oracle, name, channels):
    rng = np.random.default_rng(10)
    shape = (14, 60) + ((channels,) if channels else ())
    L, R = rng.uniform(0, 1, shape).astype(np.float32), rng.uniform(0, 1, shape).astype(np.float32)
    got = api.unfoldBasedCostVolumeCompressed(getattr(api, name), L, R, GRPIX_LIKE, 12, layout="c", ctx=ctx)
    for chk in checkers(oracle):
        close(got, chk.cost_volume_compressed(getattr(oracle, name), GRPIX_LIKE, L, R, 12), 1e-5)


@pytest.mark.parametrize("name", ["ZNCC", "SAD", "ZSSD", "CC"])
def test_multichannel_images(api, ctx, oracle, name):
    rng = np.random.default_rng(11)
    L, R = rng.uniform(0, 1, (12, 50, 3)).astype(np.float32), rng.uniform(0, 1, (12, 50, 3)).astype(np.float32)
    got = api.unfoldBasedCostVolumeMultiChannel(getattr(api, name), L, R, 2, 1, 10, layout="c", ctx=ctx)
    for chk in checkers(oracle):
        close(got, chk.cost_volume_multichannel(getattr(oracle, name), L, R, 2, 1, 10), 1e-5)
    assert api.unfoldBasedCostVolumeMultiChannel(getattr(api, name), L, R[:, :, :2], 2, 1, 10, ctx=ctx) is None   # channel counts differ
