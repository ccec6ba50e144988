"""Deferred host completion (stevi_ctx_set_deferred_host) and HostPairStream: pairs left in flight from page-locked host
memory give the same bytes as the same calls made one at a time."""
import numpy as np
import pytest

from libstevi_b200 import synth

pytestmark = pytest.mark.gpu


def pinned(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).pin_memory().numpy()


def test_census_sgm_pairs_in_flight(api, ctx):
    pairs = [synth.stereo_pair(60, 112, 30 + k, d_bg=3, d_sq=7 + k) for k in range(7)]
    want = [api.censusSgmPipeline(L, R, 4, 3, 24, 10.0, 120.0, ctx=ctx) for L, R, _ in pairs]
    hs = api.HostPairStream(depth=3)
    held = [(pinned(L), pinned(R)) for L, R, _ in pairs]
    got = [pinned(np.full((60, 112), -7, np.int32)) for _ in pairs]
    for (l, r), out in zip(held, got):
        api.censusSgmPipeline(l, r, 4, 3, 24, 10.0, 120.0, out=out, ctx=hs.acquire())
    hs.drain()
    for w, g in zip(want, got):
        assert np.array_equal(w, g)
    assert hs.launches > 0
    hs.close()


def test_local_matching_pairs_in_flight(api, ctx):
    pairs = [synth.stereo_pair(48, 100, 50 + k, d_bg=2, d_sq=6, dtype=np.float32) for k in range(5)]
    want = [api.localMatchingPipeline(api.ZNCC, L, R, 3, 3, 16, api.PARABOLA, ctx=ctx) for L, R, _ in pairs]
    hs = api.HostPairStream(depth=2)
    held = [(pinned(L), pinned(R)) for L, R, _ in pairs]
    got = [pinned(np.zeros((48, 100), np.float32)) for _ in pairs]
    for (l, r), out in zip(held, got):
        api.localMatchingPipeline(api.ZNCC, l, r, 3, 3, 16, api.PARABOLA, out=out, ctx=hs.acquire())
    hs.drain()
    for w, g in zip(want, got):
        assert np.array_equal(w, g, equal_nan=True)
    hs.close()


def test_deferred_flag_off_by_default_and_reversible(api):
    c = api.Context(0)
    L, R, _ = synth.stereo_pair(40, 64, 3)
    a = api.censusSgmPipeline(L, R, 2, 2, 8, 5.0, 50.0, ctx=c)          # synchronous: result in place on return
    c.set_deferred_host(True)
    out = pinned(np.zeros((40, 64), np.int32))
    api.censusSgmPipeline(pinned(L), pinned(R), 2, 2, 8, 5.0, 50.0, out=out, ctx=c)
    c.synchronize()
    assert np.array_equal(a, out)
    c.set_deferred_host(False)
    assert np.array_equal(a, api.censusSgmPipeline(L, R, 2, 2, 8, 5.0, 50.0, ctx=c))
    c.close()


# ---- narrow wire: index / disparity maps returned to host memory cross PCIe as u8 / u16 and are widened by the host ----------

@pytest.mark.parametrize("D", [40, 256, 300])
@pytest.mark.parametrize("ddir", ["r2l", "l2r"])
def test_narrow_wire_gives_the_same_disparities(api, D, ddir):
    from libstevi_b200 import synth
    L, R, _ = synth.stereo_pair(60, 400, 9, d_bg=5, d_sq=20, d_max=60)
    direction = api.RIGHT_TO_LEFT if ddir == "r2l" else api.LEFT_TO_RIGHT
    a, b = api.Context(0), api.Context(0)
    b.set_narrow_wire(False)
    try:
        da = api.censusSgmPipeline(L, R, 4, 3, D, 10, 120, 100, 8, dDir=direction, ctx=a)
        db = api.censusSgmPipeline(L, R, 4, 3, D, 10, 120, 100, 8, dDir=direction, ctx=b)
        assert da.dtype == np.int32 and np.array_equal(da, db)
        assert (da.min() < 0) == (ddir == "l2r")
        # a strided host result (not a dense map) takes the plain int32 copy and still agrees
        wide = np.full((60, 2 * da.shape[1]), -7, np.int32)
        api.censusSgmPipeline(L, R, 4, 3, D, 10, 120, 100, 8, dDir=direction, out=wide[:, ::2], ctx=a)
        assert np.array_equal(wide[:, ::2], db) and (wide[:, 1::2] == -7).all()
        # extractSelectedIndex of a host volume
        v = np.random.default_rng(1).integers(0, 50, (30, 40, D)).astype(np.uint16)
        assert np.array_equal(api.extractSelectedIndex(v, api.COST, ctx=a), api.extractSelectedIndex(v, api.COST, ctx=b))
    finally:
        a.close(); b.close()


def test_narrow_wire_with_pairs_in_flight(api):
    from libstevi_b200 import synth
    hs = api.HostPairStream(3, 0)
    pairs = [synth.stereo_pair(48, 200, 30 + k, d_bg=3, d_sq=11, d_max=20)[:2] for k in range(7)]
    outs = [np.empty((48, 200), np.int32) for _ in pairs]
    for (L, R), o in zip(pairs, outs):
        api.censusSgmPipeline(L, R, 4, 3, 32, 10, 120, 100, 8, out=o, ctx=hs.acquire())
    hs.drain()
    c = api.Context(0)
    c.set_narrow_wire(False)
    for (L, R), o in zip(pairs, outs):
        assert np.array_equal(o, api.censusSgmPipeline(L, R, 4, 3, 32, 10, 120, 100, 8, ctx=c))
    c.close(); hs.close()


# ---- the batch entry of the C ABI: pairs over several contexts (and devices), one pair in flight per context ---------------

@pytest.mark.parametrize("n_ctx", [1, 3])
def test_batch_entry_equals_single_calls(api, n_ctx):
    from libstevi_b200 import synth
    pairs = [synth.stereo_pair(40, 180, 50 + k, d_bg=3, d_sq=9, d_max=16)[:2] for k in range(7)]
    outs = [np.full((40, 180), -1, np.int32) for _ in pairs]
    ctxs = [api.Context(0) for _ in range(n_ctx)]
    api.batchCensusSgm(ctxs, [(L, R, o) for (L, R), o in zip(pairs, outs)], 4, 3, 24, 10, 120, 100, 8)
    one = api.Context(0)
    for (L, R), o in zip(pairs, outs):
        assert np.array_equal(o, api.censusSgmPipeline(L, R, 4, 3, 24, 10, 120, 100, 8, ctx=one))
    # u8 results and an empty batch
    o8 = [np.empty((40, 180), np.uint8) for _ in pairs]
    api.batchCensusSgm(ctxs, [(L, R, o) for (L, R), o in zip(pairs, o8)], 4, 3, 24, 10, 120, 100, 8)
    assert all(np.array_equal(a.astype(np.int32), b) for a, b in zip(o8, outs))
    api.batchCensusSgm(ctxs, [], 4, 3, 24, 10, 120, 100, 8)
    # a failing pair stops the batch with the library's status
    with pytest.raises(api.SteviError):
        api.batchCensusSgm(ctxs, [(pairs[0][0], pairs[0][1][:30], outs[0])], 4, 3, 24, 10, 120, 100, 8)
    for c in ctxs + [one]:
        c.close()


def test_batch_entry_over_all_devices(api):
    import ctypes
    n_dev = api.lib.stevi_device_count()
    if n_dev < 2:
        pytest.skip("one device")
    from libstevi_b200 import synth
    pairs = [synth.stereo_pair(48, 160, 70 + k, d_bg=3, d_sq=9, d_max=16)[:2] for k in range(4 * n_dev + 1)]
    outs = [np.empty((48, 160), np.int32) for _ in pairs]
    ctxs = [api.Context(d) for d in range(n_dev) for _ in range(2)]
    api.batchCensusSgm(ctxs, [(L, R, o) for (L, R), o in zip(pairs, outs)], 4, 3, 24, 10, 120, 100, 8)
    one = api.Context(0)
    for (L, R), o in zip(pairs, outs):
        assert np.array_equal(o, api.censusSgmPipeline(L, R, 4, 3, 24, 10, 120, 100, 8, ctx=one))
    for c in ctxs + [one]:
        c.close()
