"""GPU parity of the census -> Hamming cost volume -> SGM -> WTA path against the CPU oracle (bit-exact).

Everything goes through the C ABI (libstevi_b200.api is a ctypes veneer over it). Sizes are kept where the oracle's
O(D^2)-free restatement finishes in seconds; the full-size configurations are covered through properties in
test_gpu_fullsize.py.
"""
import numpy as np
import pytest

from libstevi_b200 import synth

pytestmark = pytest.mark.gpu


def pair(h, w, seed, dtype=np.uint8, d_max=12):
    return synth.stereo_pair(h, w, seed, d_bg=2, d_sq=7, d_max=d_max, dtype=dtype)


@pytest.mark.parametrize("hr,vr", [(4, 3), (3, 3), (2, 2), (5, 1), (1, 1), (6, 5)])
@pytest.mark.parametrize("dtype", [np.uint8, np.float32])
def test_census_transform(api, ctx, oracle, hr, vr, dtype):
    L, _, _ = pair(45, 70, 3, dtype)
    got = api.censusTransform2D(L, hr, vr, ctx=ctx)
    want = oracle.port.census(L, hr, vr)
    assert got.shape == want.shape
    assert np.array_equal(got, want)


def test_census_transform_strided_input(api, ctx, oracle):
    L, _, _ = pair(40, 66, 5)
    Lt = np.asfortranarray(L)  # first-index-fastest image
    got = api.censusTransform2D(Lt, 4, 3, ctx=ctx)
    assert np.array_equal(got, oracle.port.census(L, 4, 3))


@pytest.mark.parametrize("layout", ["ref", "c", "f"])
@pytest.mark.parametrize("direction", [0, 1])
@pytest.mark.parametrize("hr,vr,D", [(4, 3, 16), (3, 3, 21), (6, 5, 9), (2, 2, 8), (4, 3, 1)])
def test_census_cost_volume(api, ctx, oracle, layout, direction, hr, vr, D):
    L, R, _ = pair(37, 88, 11)
    got = api.unfoldBasedCostVolume(api.CENSUS, L, R, hr, vr, D, dDir=direction, layout=layout, ctx=ctx)
    want = oracle.port.cost_volume(oracle.CENSUS, L, R, hr, vr, D, direction)
    assert got.dtype == np.uint16 and got.shape == want.shape
    assert np.array_equal(got, want)


@pytest.mark.parametrize("w", [1024, 1032, 1283, 2056])
def test_census_cost_volume_wide_rows(api, ctx, oracle, w):
    # more than one 1024-column block per row, vector and scalar store paths, ragged last thread
    L, R, _ = pair(9, w, 13)
    got = api.unfoldBasedCostVolume(api.CENSUS, L, R, 4, 3, 40, ctx=ctx)
    assert np.array_equal(got, oracle.port.cost_volume(oracle.CENSUS, L, R, 4, 3, 40))


def test_census_cost_volume_different_widths(api, ctx, oracle):
    L, _, _ = pair(20, 90, 1)
    _, R, _ = pair(20, 64, 2)
    for direction in (0, 1):
        got = api.unfoldBasedCostVolume(api.CENSUS, L, R, 4, 3, 12, dDir=direction, ctx=ctx)
        want = oracle.port.cost_volume(oracle.CENSUS, L, R, 4, 3, 12, direction)
        assert got.shape == want.shape and np.array_equal(got, want)


def test_cost_volume_height_mismatch_is_empty(api, ctx):
    L, _, _ = pair(20, 64, 1)
    _, R, _ = pair(21, 64, 2)
    assert api.unfoldBasedCostVolume(api.CENSUS, L, R, 4, 3, 12, ctx=ctx) is None  # cross_correlations.h:751


@pytest.mark.parametrize("tcv", [np.uint16, np.int32, np.float32])
def test_census_cost_volume_types(api, ctx, oracle, tcv):
    L, R, _ = pair(24, 72, 4, np.float32)
    got = api.unfoldBasedCostVolume(api.CENSUS, L, R, 3, 3, 10, TCV=tcv, ctx=ctx)
    want = oracle.port.cost_volume(oracle.CENSUS, L, R, 3, 3, 10)
    assert got.dtype == tcv and np.array_equal(got.astype(np.int64), want.astype(np.int64))


def test_hamming_q2_rounding_and_many_words(api, ctx, oracle):
    # raw 32-bit features exercise the float rounding of target words (Q2), incl. words >= 0xFFFFFF80
    rng = np.random.default_rng(5)
    for nw in (1, 2, 3, 4, 6):
        fl = rng.integers(0, 2**32, size=(6, 50, nw), dtype=np.uint64).astype(np.uint32)
        fr = rng.integers(0, 2**32, size=(6, 50, nw), dtype=np.uint64).astype(np.uint32)
        fl[0, :8, 0] = np.uint32(0xFFFFFF80) + np.arange(8, dtype=np.uint32) * 16
        got = api.hammingCostVolume(fl, fr, 11, ctx=ctx)
        rounded = np.vectorize(lambda v: oracle.port.lib.orc_round_through_float(int(v)), otypes=[np.uint32])(fl)
        want = np.zeros((6, 50, 11), np.uint16)
        for d in range(11):
            t = np.zeros_like(fr)
            t[:, :50 - d] = rounded[:, d:]
            want[:, :, d] = np.bitwise_count(fr ^ t).sum(axis=2)
        assert np.array_equal(got, want), nw


@pytest.mark.parametrize("ndir", [4, 8, 16])
@pytest.mark.parametrize("layout", ["ref", "c", "f"])
def test_sgm_cost_compat(api, ctx, oracle, ndir, layout):
    L, R, _ = pair(41, 72, 21)
    cv = api.unfoldBasedCostVolume(api.CENSUS, L, R, 4, 3, 24, layout=layout, ctx=ctx)
    got = api.sgmCostVolume(cv, 10, 120, Pout=100, nDirections=ndir, strategy=api.COST, ctx=ctx)
    want = oracle.port.sgm(np.ascontiguousarray(cv), ndir, oracle.COST, 10, 120, 100)
    assert got.dtype == np.float32
    assert np.array_equal(got, want)


@pytest.mark.parametrize("margins", [(0, 0, 0, 0), (3, 2, 5, 1), (10, 10, 10, 10)])
def test_sgm_margins(api, ctx, oracle, margins):
    L, R, _ = pair(33, 64, 8)
    cv = api.unfoldBasedCostVolume(api.CENSUS, L, R, 4, 3, 16, ctx=ctx)
    got = api.sgmCostVolume(cv, 10, 120, margins=margins, Pout=100, ctx=ctx)
    want = oracle.port.sgm(np.ascontiguousarray(cv), 8, oracle.COST, 10, 120, 100, margins)
    assert np.array_equal(got, want)


def test_sgm_random_u16_volume(api, ctx, oracle):
    # large costs: the a-priori exactness bound fails, the measured one holds
    rng = np.random.default_rng(9)
    cv = rng.integers(0, 3000, size=(37, 53, 19)).astype(np.uint16)
    got = api.sgmCostVolume(cv, 1, 2, Pout=100, ctx=ctx)
    assert np.array_equal(got, oracle.port.sgm(cv, 8, oracle.COST, 1, 2, 100))


@pytest.mark.parametrize("strategy", [0, 1])
@pytest.mark.parametrize("dtype", [np.uint16, np.float32, np.int32])
@pytest.mark.parametrize("layout", ["ref", "c", "f"])
def test_wta(api, ctx, oracle, strategy, dtype, layout):
    rng = np.random.default_rng(3)
    v = rng.integers(0, 12, size=(19, 70, 40)).astype(dtype)  # many ties: the LAST extremum must win (Q5)
    order = {"ref": (0, 2, 1), "c": (0, 1, 2), "f": (2, 1, 0)}[layout]
    mem = np.ascontiguousarray(v.transpose(order))
    vol = mem.transpose(tuple(order.index(k) for k in range(3)))
    assert np.array_equal(vol, v)
    got = api.extractSelectedIndex(vol, strategy, ctx=ctx)
    if dtype == np.int32:
        want = oracle.port.wta(v.astype(np.float32), strategy)
    else:
        want = oracle.port.wta(v, strategy)
    assert np.array_equal(got, want)


def test_wta_nan_rules(api, ctx, oracle):
    rng = np.random.default_rng(4)
    v = rng.standard_normal((8, 33, 64)).astype(np.float32)
    v[0, :, 0] = np.nan          # NaN at d = 0 sticks
    v[1, :, 5] = np.nan          # NaN elsewhere never wins
    v[2, :, :] = np.nan
    v[3, :, :] = np.inf
    for layout_order in ((0, 1, 2), (0, 2, 1)):
        mem = np.ascontiguousarray(v.transpose(layout_order))
        vol = mem.transpose(tuple(layout_order.index(k) for k in range(3)))
        for strategy in (0, 1):
            assert np.array_equal(api.extractSelectedIndex(vol, strategy, ctx=ctx), oracle.port.wta(v, strategy))


def test_index_to_disp(api, ctx, oracle):
    idx = np.arange(35, dtype=np.int32).reshape(5, 7)
    for direction in (0, 1):
        assert np.array_equal(api.selectedIndexToDisp(idx, 3, direction, ctx=ctx), oracle.port.index_to_disp(idx, direction, 3))


@pytest.mark.parametrize("h,w,D", [(48, 64, 16), (60, 136, 40)])
@pytest.mark.parametrize("ndir", [4, 8])
def test_pipeline_matches_stages_and_oracle(api, ctx, oracle, h, w, D, ndir):
    L, R, gt = pair(h, w, 31, d_max=0)
    cv, sg, disp = api.censusSgmPipeline(L, R, 4, 3, D, 10, 120, Pout=100, nDirections=ndir, return_volumes=True, ctx=ctx)
    ocv = oracle.port.cost_volume(oracle.CENSUS, L, R, 4, 3, D)
    osg = oracle.port.sgm(ocv, ndir, oracle.COST, 10, 120, 100)
    odisp = oracle.port.index_to_disp(oracle.port.wta(osg, oracle.COST))
    assert np.array_equal(cv, ocv)
    assert np.array_equal(sg, osg)
    assert np.array_equal(disp, odisp)
    # disparity-only call gives the same map
    assert np.array_equal(api.censusSgmPipeline(L, R, 4, 3, D, 10, 120, Pout=100, nDirections=ndir, ctx=ctx), odisp)


@pytest.mark.parametrize("h,w,D,hr,vr,ndir,direction,margins,pout", [
    (37, 83, 19, 4, 3, 8, 1, (0, 0, 0, 0), 100),     # ragged width and disparity count
    (37, 83, 19, 4, 3, 8, 0, (0, 0, 0, 0), 100),     # left-to-right
    (40, 1100, 70, 2, 2, 16, 1, (0, 0, 0, 0), 35),   # more than one 1024-column block, 16 directions (Q8-free subset)
    (52, 96, 33, 3, 1, 4, 1, (3, 2, 5, 1), 100),     # margins
    (30, 64, 8, 5, 4, 8, 1, (0, 0, 0, 0), -7),       # 11x9 window (3 census words, one written), negative Pout
    (20, 40, 64, 1, 1, 8, 1, (0, 0, 0, 0), 100),     # disparity range wider than the image
])
def test_disparity_only_pipeline_equals_materialised(api, ctx, h, w, D, hr, vr, ndir, direction, margins, pout):
    """The disparity-only call keeps no volume in memory (costs recomputed from the census words by the g-map pass and by
    the aggregation + WTA pass); its map is the one of the call that materialises both volumes."""
    L, R, _ = pair(h, w, 5 + h, d_max=0)
    kw = dict(Pout=pout, nDirections=ndir, margins=margins, dDir=direction, ctx=ctx)
    _, sg, want = api.censusSgmPipeline(L, R, hr, vr, D, 10, 120, return_volumes=True, **kw)
    got = api.censusSgmPipeline(L, R, hr, vr, D, 10, 120, **kw)
    assert np.array_equal(got, want)
    # and it is the last row-wise minimum of the aggregated volume (Q5), folded into a disparity
    idx = D - 1 - np.argmin(sg[:, :, ::-1], axis=2)
    assert np.array_equal(got, idx if direction == 1 else -idx)


def test_pipeline_device_operands(api, ctx, oracle):
    torch = pytest.importorskip("torch")
    L, R, _ = pair(64, 128, 77)
    dl, dr = torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda()
    tctx = api.Context(0, torch_stream=True)
    disp = api.censusSgmPipeline(dl, dr, 4, 3, 32, 10, 120, ctx=tctx)
    torch.cuda.synchronize()
    ocv = oracle.port.cost_volume(oracle.CENSUS, L, R, 4, 3, 32)
    want = oracle.port.wta(oracle.port.sgm(ocv, 8, oracle.COST, 10, 120, 100), oracle.COST)
    assert np.array_equal(disp.cpu().numpy(), want)
    tctx.close()
