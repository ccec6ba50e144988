"""On-demand cost evaluation and PatchMatch (correlation/on_demand_cost_volume.h, correlation/patchmatch.h)
against the reference's unmodified code (oracle/_ref/libstevi_ref_ibr.so).

* truncatedCostVolume(disp, radius) of the on-demand volumes: the same cells, equal to the reference's to 2e-6 of the largest magnitude
  (sums in its channel order), cells without a value equal to its default, and equal to a gather from the full cost volume
  (the reference's own test, test/unittests/testOnDemandCostVolume.cpp:62-172).
* PatchMatch with nRandomSearch = 0 from a given initial map is deterministic in the reference (its random draws are seeded from
  std::random_device): equal disparity maps, 1-D and 2-D searches, all four scan directions, the early end on an iteration
  without a change.
* With the random search both sides are randomised: what is compared is the agreement with the exhaustive winner."""
import numpy as np
import pytest

from libstevi_b200 import synth

pytestmark = pytest.mark.gpu

FUNCS = ["ZNCC", "SSD", "ZSAD"]


def need_ref(oracle):
    if oracle.ref_ibr is None or not oracle.ref_ibr.has("patch_match"):
        pytest.skip("oracle/_ref/libstevi_ref_ibr.so was not built")
    return oracle.ref_ibr


def scene(oracle, h=40, w=96, seed=3, hr=2, vr=1):
    L, R, gt = synth.stereo_pair(h, w, seed, d_bg=3, d_sq=8, d_max=0, dtype=np.float32)
    return oracle.port.unfold(R, hr, vr), oracle.port.unfold(L, hr, vr), L, R, gt   # source = right image, target = left


@pytest.mark.parametrize("name", FUNCS)
@pytest.mark.parametrize("search", [(0, 11), (-2, 9), (3, 12)])
@pytest.mark.parametrize("radius", [1, 2])
def test_on_demand_truncated_cost_volume_1d(api, ctx, oracle, name, search, radius):
    ref = need_ref(oracle)
    fs, ft, *_ = scene(oracle)
    rng = np.random.default_rng(5)
    disp = rng.integers(search[0] - 1, search[1] + 2, fs.shape[:2]).astype(np.int32)
    got = api.onDemandTruncatedCostVolume(getattr(api, name), fs, ft, search, disp, radius, ctx=ctx)
    want = ref.on_demand_truncated_cost_volume(getattr(oracle, name), fs, ft, search, disp, radius)
    missing = np.abs(want) > 1e30 if name != "ZNCC" else want == np.float32(1.17549435e-38)
    assert np.array_equal(got[missing], want[missing])
    assert np.abs(got[~missing] - want[~missing]).max() <= 2e-6 * max(1.0, np.abs(want[~missing]).max())
    if search[0] == 0:   # equal to the cells of the full cost volume (testOnDemandCostVolume.cpp:62-112)
        mf = getattr(oracle, name)
        cv = oracle.port.aggregate_cost(mf, ft, fs, 0, search[1], oracle.R2L) if False else None
        full = oracle.port.feature_volume_to_cost_volume(mf, ft, fs, search[1] + 1, oracle.R2L)
        jj = np.arange(fs.shape[1])[None, :]
        for k in range(2 * radius + 1):
            d = disp + k - radius
            ok = (d >= 0) & (d <= search[1]) & (jj + d < fs.shape[1])
            cells = np.take_along_axis(full, np.clip(d, 0, search[1])[..., None], axis=2)[..., 0]
            assert np.abs(got[..., k][ok] - cells[ok]).max() <= 2e-6 * max(1.0, np.abs(cells[ok]).max())
            assert np.array_equal(missing[..., k], ~ok)


@pytest.mark.parametrize("name", ["ZNCC", "SSD"])
def test_on_demand_truncated_cost_volume_2d(api, ctx, oracle, name):
    ref = need_ref(oracle)
    fs, ft, *_ = scene(oracle, 30, 64)
    search = (-2, 2, 0, 9)
    rng = np.random.default_rng(6)
    disp = np.stack([rng.integers(-3, 4, fs.shape[:2]), rng.integers(-1, 11, fs.shape[:2])], axis=2).astype(np.int32)
    got = api.onDemandTruncatedCostVolume(getattr(api, name), fs, ft, search, disp, 1, ctx=ctx)
    want = ref.on_demand_truncated_cost_volume(getattr(oracle, name), fs, ft, search, disp, 1)
    missing = np.abs(want) > 1e30 if name != "ZNCC" else want == np.float32(1.17549435e-38)
    assert np.array_equal(got[missing], want[missing])
    assert np.abs(got[~missing] - want[~missing]).max() <= 2e-6 * max(1.0, np.abs(want[~missing]).max())


@pytest.mark.parametrize("name", FUNCS)
@pytest.mark.parametrize("n_iter", [1, 2, 3, 4, 7])
def test_patch_match_propagation_equals_the_reference_1d(api, ctx, oracle, name, n_iter):
    ref = need_ref(oracle)
    fs, ft, *_ = scene(oracle)
    rng = np.random.default_rng(11 + n_iter)
    init = rng.integers(-1, 14, fs.shape[:2] + (1,)).astype(np.int32)   # some start outside the range (0, 11): no cell there
    got = api.patchMatch(getattr(api, name), fs, ft, (0, 11), n_iter, 0, initializer=init, ctx=ctx)
    want = ref.patch_match(getattr(oracle, name), fs, ft, (0, 11), n_iter, 0, init)
    assert np.array_equal(want, oracle.port.patch_match(getattr(oracle, name), fs, ft, (0, 11), n_iter, 0, init))   # the C restatement
    assert (got == want).mean() >= 0.999   # equal up to ties decided by the last bit of a sum
    assert np.array_equal(got, want) or name == "ZNCC"


@pytest.mark.parametrize("name", ["ZNCC", "SSD"])
def test_patch_match_propagation_equals_the_reference_2d(api, ctx, oracle, name):
    ref = need_ref(oracle)
    fs, ft, L, R, _ = scene(oracle, 32, 72)
    ft2 = oracle.port.unfold(np.roll(L, 1, axis=0), 2, 1)
    search = (-2, 2, 0, 9)
    rng = np.random.default_rng(13)
    init = np.stack([rng.integers(-2, 3, fs.shape[:2]), rng.integers(0, 10, fs.shape[:2])], axis=2).astype(np.int32)
    for n_iter in (1, 4, 6):
        got = api.patchMatch(getattr(api, name), fs, ft2, search, n_iter, 0, initializer=init, ctx=ctx)
        want = ref.patch_match(getattr(oracle, name), fs, ft2, search, n_iter, 0, init)
        assert (got == want).all(axis=2).mean() >= 0.999


def test_patch_match_ends_on_an_iteration_without_change(api, ctx, oracle):
    ref = need_ref(oracle)
    fs, ft, *_ = scene(oracle)
    init = np.full(fs.shape[:2] + (1,), 5, np.int32)   # a constant map: propagation changes nothing, the run ends after one iteration
    got = api.patchMatch(api.SSD, fs, ft, (0, 11), 9, 0, initializer=init, ctx=ctx)
    want = ref.patch_match(oracle.SSD, fs, ft, (0, 11), 9, 0, init)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("name", ["ZNCC", "ZSAD"])
def test_patch_match_with_random_search_finds_the_exhaustive_winner(api, ctx, oracle, name):
    ref = need_ref(oracle)
    L, R, gt = synth.stereo_pair(96, 160, 1, d_bg=4, d_sq=12, d_max=0, dtype=np.float32)
    fs, ft = oracle.port.unfold(R, 2, 2), oracle.port.unfold(L, 2, 2)
    mf = getattr(oracle, name)
    best = oracle.port.wta(oracle.port.cost_volume(mf, L, R, 2, 2, 16), oracle.strategy_of(mf))
    inner = (slice(4, -4), slice(4, -20))
    got = api.patchMatch(getattr(api, name), fs, ft, (0, 15), 5, 4, seed=7, ctx=ctx)[..., 0]
    theirs = ref.patch_match(mf, fs, ft, (0, 15), 5, 4)[..., 0]
    a, b = float((got[inner] == best[inner]).mean()), float((theirs[inner] == best[inner]).mean())
    assert a >= 0.97 and a >= b - 0.01, (a, b)
    assert got.min() >= 0 and got.max() <= 15
    again = api.patchMatch(getattr(api, name), fs, ft, (0, 15), 5, 4, seed=7, ctx=ctx)[..., 0]
    assert np.array_equal(got, again)   # the counter-based generator makes a run repeatable
    assert not np.array_equal(got, api.patchMatch(getattr(api, name), fs, ft, (0, 15), 1, 4, seed=8, ctx=ctx)[..., 0])


def test_shape_rules(api, ctx):
    a, b = np.zeros((5, 8, 9), np.float32), np.zeros((6, 8, 9), np.float32)
    assert api.patchMatch(api.SSD, a, b, (0, 3), ctx=ctx) is None                      # 1-D search: heights differ, empty array
    assert api.patchMatch(api.SSD, a, np.zeros((5, 8, 4), np.float32), (0, 3), ctx=ctx) is None
    assert api.patchMatch(api.SSD, a, b, (-1, 1, 0, 3), 1, 0, ctx=ctx).shape == (5, 8, 2)
    with pytest.raises(api.SteviError):
        api.patchMatch(api.CENSUS, a, a, (0, 3), ctx=ctx)


def test_full_size_patch_match_properties(api, ctx, oracle):
    """1080p, 5x5 windows, device-resident: PatchMatch ends within the search range, agrees with the exhaustive winner on almost
    all pixels, never ends on a worse cell than it started from, and is repeatable."""
    torch = pytest.importorskip("torch")
    L, R, gt = synth.stereo_pair(1080, 1920, 9, d_bg=8, d_sq=32, d_max=60, dtype=np.float32)
    tctx = api.Context(0, torch_stream=True)
    dl, dr = torch.from_numpy(L).cuda(), torch.from_numpy(R).cuda()
    fl, fr = api.unfold(2, 2, dl, ctx=tctx), api.unfold(2, 2, dr, ctx=tctx)
    init = torch.randint(0, 128, (1080, 1920, 1), dtype=torch.int32, device="cuda")
    out = api.patchMatch(api.ZNCC, fr, fl, (0, 127), 5, 4, initializer=init, seed=3, ctx=tctx)
    again = api.patchMatch(api.ZNCC, fr, fl, (0, 127), 5, 4, initializer=init, seed=3, ctx=tctx)
    torch.cuda.synchronize()
    assert torch.equal(out, again)
    d = out[..., 0].cpu().numpy()
    assert d.min() >= 0 and d.max() <= 127
    best = api.extractSelectedIndex(api.unfoldBasedCostVolume(api.ZNCC, dl, dr, 2, 2, 128, ctx=tctx), api.SCORE, ctx=tctx).cpu().numpy()
    inner = (slice(8, -8), slice(8, 1920 - 136))
    assert (d[inner] == best[inner]).mean() > 0.97
    # the cost of the final disparity is at least the cost of the initial one wherever the initial one named a cell
    c0 = api.onDemandTruncatedCostVolume(api.ZNCC, fr, fl, (0, 127), init, 0, ctx=tctx)[..., 0].cpu().numpy()
    c1 = api.onDemandTruncatedCostVolume(api.ZNCC, fr, fl, (0, 127), out, 0, ctx=tctx)[..., 0].cpu().numpy()
    named = c0 != np.float32(1.17549435e-38)
    assert (c1[named] >= c0[named]).all()
    tctx.close()
