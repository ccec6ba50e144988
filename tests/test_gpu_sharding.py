"""Row-band partitioning on the GPU: the rows a rank computes from its band (plus halos, in global coordinates) are
bit-identical to the same rows of the whole-frame call — for local matching with refinement and for the coarse-to-fine
driver at every depth. All bands are run on one GPU here; bench.py --workload c4 --gpus N runs them one per rank."""
import numpy as np
import pytest

from libstevi_b200 import sharding as S
from libstevi_b200 import synth

pytestmark = pytest.mark.gpu


def eq(a, b):
    return a.shape == b.shape and np.array_equal(a, b, equal_nan=True)


@pytest.mark.parametrize("world", [2, 3, 8])
def test_local_matching_bands_equal_whole_frame(api, ctx, world):
    L, R, _ = synth.stereo_pair(70, 96, 5, d_bg=3, d_sq=9, dtype=np.float32)
    _, idx, _, refined = api.localMatchingPipeline(api.ZNCC, L, R, 3, 2, 20, api.PARABOLA, return_all=True, ctx=ctx)
    for g0, g1 in S.row_bands(70, world):
        bi, br = S.local_matching_rows(api, api.ZNCC, L, R, 3, 2, 20, g0, g1, api.PARABOLA, ctx=ctx)
        assert eq(bi, idx[g0:g1]) and eq(br, refined[g0:g1])


@pytest.mark.parametrize("depth", [1, 2, 3])
@pytest.mark.parametrize("height,world", [(96, 2), (101, 4), (64, 8)])
def test_hierarchical_bands_equal_whole_frame(api, ctx, depth, height, world):
    L, R, _ = synth.stereo_pair(height, 120, 9, d_bg=4, d_sq=11, dtype=np.float32)
    tcv, disp = api.hiearchicalTruncatedCostVolume(api.ZNCC, depth, L, R, 2, 2, 24, 2, ctx=ctx)
    for g0, g1 in S.row_bands(height, world, 2):
        if g1 <= g0:
            continue
        bt, bd = S.hierarchical_rows(api, api.ZNCC, depth, L, R, 2, 2, 24, g0, g1, 2, ctx=ctx)
        assert eq(bd, disp[g0:g1]) and eq(bt, tcv[g0:g1])
        # the same from a band buffer that only holds the rows the plan asks for
        plan = S.band_plan(height, depth, [2] * (depth + 1), g0, g1)[-1]
        bt2, bd2 = S.hierarchical_rows(api, api.ZNCC, depth, L[plan.img_lo:plan.img_hi], R[plan.img_lo:plan.img_hi], 2, 2, 24,
                                       g0, g1, 2, band_row0=plan.img_lo, full_height=height, ctx=ctx)
        assert eq(bd2, bd) and eq(bt2, bt)


def test_guided_band_argument_checks(api, ctx):
    L, R, _ = synth.stereo_pair(40, 50, 1, dtype=np.float32)
    guide = np.zeros((20, 25), np.int32)
    with pytest.raises(api.SteviError):  # no halo on the inner (top) edge
        api.computeGuidedCVBand(api.ZNCC, L[10:30], R[10:30], 2, 2, guide, 10, 40, 10, 10, 0, 20, ctx=ctx)
    with pytest.raises(api.SteviError):  # guide band does not reach the rows the upsampling touches
        api.computeGuidedCVBand(api.ZNCC, L[8:32], R[8:32], 2, 2, guide[:3], 8, 40, 10, 20, 0, 20, ctx=ctx)
