"""GPU parity of the general SGM recurrence (Score strategy, float volumes) against the CPU oracle.

Bit-exact for the 4- and 8-direction sets (same float operation order as the reference per path, directions and start
loops in the reference's order). 16 directions: lines of one direction overlap (Q8), the GPU adds their contributions
with float atomics => bit-exact for integer-valued data, 1e-4 otherwise.
"""
import numpy as np
import pytest

from libstevi_b200 import synth

pytestmark = pytest.mark.gpu


def census_cv(api, ctx, h=37, w=70, D=24, seed=5, layout="ref"):
    L, R, _ = synth.stereo_pair(h, w, seed, d_bg=2, d_sq=7, d_max=10)
    return api.unfoldBasedCostVolume(api.CENSUS, L, R, 4, 3, D, layout=layout, ctx=ctx)


def zncc_like_volume(h, w, D, seed):
    rng = np.random.default_rng(seed)
    return rng.uniform(-1, 1, size=(h, w, D)).astype(np.float32)


@pytest.mark.parametrize("ndir", [4, 8, 16])
@pytest.mark.parametrize("layout", ["ref", "c"])
def test_score_mode_u16(api, ctx, oracle, ndir, layout):
    cv = census_cv(api, ctx, layout=layout)
    got = api.sgmCostVolume(cv, 3, 20, Pout=50, nDirections=ndir, strategy=api.SCORE, ctx=ctx)
    want = oracle.port.sgm(np.ascontiguousarray(cv), ndir, oracle.SCORE, 3, 20, 50)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("ndir", [4, 8])
@pytest.mark.parametrize("D", [7, 32, 40, 100, 200])
def test_score_mode_f32_bit_exact(api, ctx, oracle, ndir, D):
    v = zncc_like_volume(21, 45, D, D)
    got = api.sgmCostVolume(v, 0.001, 0.01, Pout=100, nDirections=ndir, strategy=api.SCORE, ctx=ctx)
    want = oracle.port.sgm(v, ndir, oracle.SCORE, 0.001, 0.01, 100)
    assert np.array_equal(got, want)


def test_score_mode_f32_16_directions(api, ctx, oracle):
    v = zncc_like_volume(21, 45, 33, 3)
    got = api.sgmCostVolume(v, 0.001, 0.01, Pout=100, nDirections=16, strategy=api.SCORE, ctx=ctx)
    want = oracle.port.sgm(v, 16, oracle.SCORE, 0.001, 0.01, 100)
    assert np.allclose(got, want, atol=1e-4)


@pytest.mark.parametrize("margins", [(0, 0, 0, 0), (2, 3, 1, 4)])
@pytest.mark.parametrize("P", [(0.5, 2.0), (2.0, 0.5)])  # P2 < P1 exercises the exclusion of the 3-window
def test_score_margins_and_penalties(api, ctx, oracle, margins, P):
    v = zncc_like_volume(25, 38, 19, 11)
    got = api.sgmCostVolume(v, P[0], P[1], margins=margins, Pout=3.5, nDirections=8, strategy=api.SCORE, ctx=ctx)
    want = oracle.port.sgm(v, 8, oracle.SCORE, P[0], P[1], 3.5, margins)
    assert np.array_equal(got, want)


def test_cost_mode_float_volume_compat(api, ctx, oracle):
    # float costs: no integer collapse; Q3 evaluated per path in the reference's operation order
    v = np.abs(zncc_like_volume(23, 41, 17, 2)) * 10
    for ndir in (4, 8):
        got = api.sgmCostVolume(v, 0.5, 2.0, Pout=100, nDirections=ndir, strategy=api.COST, ctx=ctx)
        assert np.array_equal(got, oracle.port.sgm(v, ndir, oracle.COST, 0.5, 2.0, 100))


def test_non_finite_scores(api, ctx, oracle):
    v = zncc_like_volume(12, 30, 20, 4)
    v[3, 5, :] = np.nan       # a flat ZNCC patch gives NaN scores (Q6)
    v[6, 7, 3] = np.inf
    v[8, :, 0] = -np.inf
    got = api.sgmCostVolume(v, 0.1, 0.4, Pout=10, nDirections=8, strategy=api.SCORE, ctx=ctx)
    want = oracle.port.sgm(v, 8, oracle.SCORE, 0.1, 0.4, 10)
    assert np.array_equal(got, want, equal_nan=True)


def test_large_u16_costs_fall_back_to_the_recurrence(api, ctx, oracle):
    # costs so large that the collapsed integer evaluation cannot be proven exact: the general kernel takes over
    rng = np.random.default_rng(1)
    cv = rng.integers(60000, 65535, size=(20, 300, 9)).astype(np.uint16)
    cv[:, ::2, :] //= 50
    got = api.sgmCostVolume(cv, 1, 2, Pout=100, ctx=ctx)
    assert np.array_equal(got, oracle.port.sgm(cv, 8, oracle.COST, 1, 2, 100))


# ---- volumes with contiguous columns (the reference's cost-volume layout {W*D, 1, W}): sgm_sweep.cu ---------------------------
# horizontal paths by k_sgm_sweep_h, vertical and diagonal paths by the sheared-strip sweeps of k_sgm_sweep_v; additions in the
# reference's order => bit-identical for the 4- and 8-direction sets

def ref_layout(v):
    """same values, memory order [H][D][W]"""
    return np.ascontiguousarray(v.transpose(0, 2, 1)).transpose(0, 2, 1)


@pytest.mark.parametrize("ndir", [4, 8])
@pytest.mark.parametrize("shape", [(21, 45, 7), (21, 44, 32), (33, 48, 40), (19, 70, 100), (40, 37, 200), (25, 64, 256),
                                   (18, 36, 300), (12, 40, 512), (130, 50, 20), (9, 200, 16), (1, 30, 9), (30, 1, 9), (5, 3, 2)])
def test_sweep_score_f32_bit_exact(api, ctx, oracle, ndir, shape):
    h, w, D = shape
    v = zncc_like_volume(h, w, D, h + w + D)
    got = api.sgmCostVolume(ref_layout(v), 0.001, 0.01, Pout=100, nDirections=ndir, strategy=api.SCORE, ctx=ctx)
    want = oracle.port.sgm(v, ndir, oracle.SCORE, 0.001, 0.01, 100)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("ndir", [4, 8])
def test_sweep_against_the_compiled_reference(api, ctx, oracle, ndir):
    if oracle.ref is None:
        pytest.skip("oracle/_ref was not built")
    v = zncc_like_volume(26, 75, 48, 77)
    got = api.sgmCostVolume(ref_layout(v), 0.002, 0.02, Pout=5, nDirections=ndir, strategy=api.SCORE, ctx=ctx)
    assert np.array_equal(got, oracle.ref.sgm(v, ndir, oracle.SCORE, 0.002, 0.02, 5))
    c = (np.abs(v) * 10).astype(np.float32)
    got = api.sgmCostVolume(ref_layout(c), 0.5, 2.0, Pout=100, nDirections=ndir, strategy=api.COST, ctx=ctx)
    assert np.array_equal(got, oracle.ref.sgm(c, ndir, oracle.COST, 0.5, 2.0, 100))


@pytest.mark.parametrize("margins", [(2, 3, 1, 4), (0, 5, 0, 0), (17, 0, 3, 0)])
@pytest.mark.parametrize("ndir", [4, 8])
def test_sweep_margins(api, ctx, oracle, margins, ndir):
    v = zncc_like_volume(45, 58, 19, 11)
    got = api.sgmCostVolume(ref_layout(v), 0.5, 2.0, margins=margins, Pout=3.5, nDirections=ndir, strategy=api.SCORE, ctx=ctx)
    want = oracle.port.sgm(v, ndir, oracle.SCORE, 0.5, 2.0, 3.5, margins)
    assert np.array_equal(got, want)


def test_sweep_u16_score_and_large_costs(api, ctx, oracle):
    cv = census_cv(api, ctx, h=60, w=90, D=24, layout="ref")
    for ndir in (4, 8):
        got = api.sgmCostVolume(cv, 3, 20, Pout=50, nDirections=ndir, strategy=api.SCORE, ctx=ctx)
        assert np.array_equal(got, oracle.port.sgm(np.ascontiguousarray(cv), ndir, oracle.SCORE, 3, 20, 50))
    rng = np.random.default_rng(1)
    big = rng.integers(60000, 65535, size=(20, 300, 9)).astype(np.uint16)
    big[:, ::2, :] //= 50
    got = api.sgmCostVolume(ref_layout(big), 1, 2, Pout=100, ctx=ctx)   # too large for the collapsed integer evaluation
    assert np.array_equal(got, oracle.port.sgm(big, 8, oracle.COST, 1, 2, 100))


def test_sweep_cost_mode_float_volume_compat(api, ctx, oracle):
    v = np.abs(zncc_like_volume(43, 81, 17, 2)) * 10
    for ndir in (4, 8):
        got = api.sgmCostVolume(ref_layout(v), 0.5, 2.0, Pout=100, nDirections=ndir, strategy=api.COST, ctx=ctx)
        assert np.array_equal(got, oracle.port.sgm(v, ndir, oracle.COST, 0.5, 2.0, 100))


def test_sweep_non_finite_scores(api, ctx, oracle):
    v = zncc_like_volume(32, 50, 20, 4)
    v[3, 5, :] = np.nan
    v[6, 7, 3] = np.inf
    v[8, :, 0] = -np.inf
    v[20, 17:40, :] = np.nan
    got = api.sgmCostVolume(ref_layout(v), 0.1, 0.4, Pout=10, nDirections=8, strategy=api.SCORE, ctx=ctx)
    want = oracle.port.sgm(v, 8, oracle.SCORE, 0.1, 0.4, 10)
    assert np.array_equal(got, want, equal_nan=True)


@pytest.mark.parametrize("shape", [(24, 72, 32), (17, 40, 64), (20, 136, 128), (12, 320, 32), (9, 264, 256)])
def test_sweep_tma_tiles(api, ctx, oracle, shape):
    """D == 32 * DH, W % 8 == 0, float costs, no left / right margin: the horizontal tiles travel by tensor-map bulk copies
    (k_sgm_sweep_h_tma); Pout columns, non-finite scores, every strategy / semantics, row margins."""
    h, w, D = shape
    v = zncc_like_volume(h, w, D, 3 * h + w + D)
    v[2, 5, :] = np.nan
    v[4, w // 2, D // 3] = np.inf
    v[5, :, 0] = -np.inf
    for ndir in (4, 8):
        got = api.sgmCostVolume(ref_layout(v), 0.02, 0.3, Pout=1.5, nDirections=ndir, strategy=api.SCORE, ctx=ctx)
        assert np.array_equal(got, oracle.port.sgm(v, ndir, oracle.SCORE, 0.02, 0.3, 1.5), equal_nan=True)
    c = (np.abs(zncc_like_volume(h, w, D, h)) * 10).astype(np.float32)
    got = api.sgmCostVolume(ref_layout(c), 0.5, 2.0, Pout=100, nDirections=8, strategy=api.COST, ctx=ctx)
    assert np.array_equal(got, oracle.port.sgm(c, 8, oracle.COST, 0.5, 2.0, 100))
    got = api.sgmCostVolume(ref_layout(c), 0.5, 2.0, Pout=100, nDirections=8, strategy=api.COST, semantics=api.CORRECTED, ctx=ctx)
    want = oracle.port.sgm_corrected(c, 8, oracle.COST, 0.5, 2.0, 100)
    assert np.allclose(got, want, rtol=0, atol=1e-4 * max(1.0, float(np.abs(want).max())))
    mg = (0, 2, 0, 1)   # (left, top, right, bottom): rows only
    got = api.sgmCostVolume(ref_layout(v), 0.5, 2.0, margins=mg, Pout=3.5, nDirections=8, strategy=api.SCORE, ctx=ctx)
    assert np.array_equal(got, oracle.port.sgm(v, 8, oracle.SCORE, 0.5, 2.0, 3.5, mg), equal_nan=True)


def test_sweep_column_slices_and_unaligned_rows(api, ctx, oracle):
    # views whose rows are not 16-byte aligned take the scalar column tiles of the horizontal kernel
    v = zncc_like_volume(20, 64, 24, 9)
    base = ref_layout(v)
    view = base[:, 3:50, :]
    got = api.sgmCostVolume(view, 0.001, 0.01, Pout=100, nDirections=8, strategy=api.SCORE, ctx=ctx)
    assert np.array_equal(got, oracle.port.sgm(np.ascontiguousarray(view), 8, oracle.SCORE, 0.001, 0.01, 100))


# ---- STEVI_CORRECTED: textbook recurrence, every direction live (the product's own definition, checked against the oracle's
# restatement of it: exact for integer-valued volumes, where every order of additions gives the same sums; 1e-4 for float ones,
# whose additions the sweeps do in sweep order instead of the reference's direction order)

@pytest.mark.parametrize("layout", ["ref", "c"])
@pytest.mark.parametrize("ndir", [4, 8])
@pytest.mark.parametrize("strategy", ["cost", "score"])
def test_corrected_semantics_integer_volume(api, ctx, oracle, layout, ndir, strategy):
    cv = census_cv(api, ctx, h=50, w=75, D=40, layout=layout)
    strat_a, strat_o = (api.COST, oracle.COST) if strategy == "cost" else (api.SCORE, oracle.SCORE)
    got = api.sgmCostVolume(cv, 3, 20, Pout=50, nDirections=ndir, strategy=strat_a, semantics=api.CORRECTED, ctx=ctx)
    want = oracle.port.sgm_corrected(np.ascontiguousarray(cv), ndir, strat_o, 3, 20, 50)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("margins", [(0, 0, 0, 0), (3, 1, 2, 5)])
def test_corrected_semantics_float_volume(api, ctx, oracle, margins):
    v = np.abs(zncc_like_volume(41, 67, 23, 8)) * 10
    for vol in (v, ref_layout(v)):
        got = api.sgmCostVolume(vol, 0.5, 2.0, margins=margins, Pout=7, nDirections=8, strategy=api.COST, semantics=api.CORRECTED, ctx=ctx)
        want = oracle.port.sgm_corrected(v, 8, oracle.COST, 0.5, 2.0, 7, margins)
        assert np.allclose(got, want, rtol=0, atol=1e-4)


@pytest.mark.parametrize("D", [513, 700])
def test_more_than_512_disparities(api, ctx, oracle, D):
    # the reference has no limit on D: beyond the register-resident kernels a block per line takes over
    v = zncc_like_volume(6, 20, D, D)
    for vol in (v, ref_layout(v)):
        got = api.sgmCostVolume(vol, 0.001, 0.01, Pout=100, nDirections=8, strategy=api.SCORE, ctx=ctx)
        assert np.array_equal(got, oracle.port.sgm(v, 8, oracle.SCORE, 0.001, 0.01, 100))
    c = (np.abs(v) * 100).astype(np.uint16)
    c[:, :, 0] = 65000   # too large for the collapsed integer evaluation
    got = api.sgmCostVolume(c, 1, 2, Pout=100, ctx=ctx)
    assert np.array_equal(got, oracle.port.sgm(c, 8, oracle.COST, 1, 2, 100))
