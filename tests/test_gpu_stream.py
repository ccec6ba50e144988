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
