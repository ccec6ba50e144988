"""stevi_census_cost_volume_noalloc (include/stevi_b200/c_api.h, csrc/noalloc.cu): the census cost volume with its scratch
memory — the census words of both images — in the library's static __device__ array instead of the context's workspace. The
entry point has no counterpart in the reference; it exists for callers that cannot call cudaMalloc. Checked through the C ABI
against stevi_cost_volume(STEVI_MF_CENSUS) on the same device operands, bit for bit (integer work: the bar of every census test
here), on calls made one after the other so that the shared array holds the previous call's contents each time.

Collected last on purpose (a failure here hides nothing behind it)."""
import ctypes

import numpy as np
import pytest

from libstevi_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.fixture()
def tctx(api):
    # bound to torch's current stream: the tensors below are produced and read on that stream
    c = api.Context(0, torch_stream=True)
    yield c
    c.close()


def _dev(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _regular(api, c, L, R, hr, vr, D, direction, out, semantics=None):
    l, r, o = api.describe(L), api.describe(R), api.describe(out)
    c.check(api.lib.stevi_cost_volume(c.handle, api.CENSUS, ctypes.byref(l), ctypes.byref(r), hr, vr, D, direction,
                                      api.COMPAT if semantics is None else semantics, ctypes.byref(o)))
    return out


def _noalloc(api, c, L, R, hr, vr, D, direction, out, semantics=None):
    l, r, o = api.describe(L), api.describe(R), api.describe(out)
    return api.lib.stevi_census_cost_volume_noalloc(c.handle, ctypes.byref(l), ctypes.byref(r), hr, vr, D, direction,
                                                    api.COMPAT if semantics is None else semantics, ctypes.byref(o))


def _volume(torch, H, W, D, dtype, layout, fill):
    # "c": (H,W,D) dense; "ref": the reference's own layout, strides {W*D, 1, W} (cross_correlations.h:220).
    # u16 volumes are made as int16 storage and re-viewed (torch fills few unsigned types)
    base = {"uint16": torch.int16, "int32": torch.int32, "float32": torch.float32}[dtype]
    v = torch.full((H, W, D) if layout == "c" else (H, D, W), fill, dtype=base, device="cuda")
    if layout != "c":
        v = v.permute(0, 2, 1)
    return v.view(torch.uint16) if dtype == "uint16" else v


def _host(t):
    import torch
    return (t.view(torch.int16) if t.dtype == torch.uint16 else t).cpu().numpy()


CASES = [
    # H, W_l, W_r, hr, vr, D, image dtype, volume dtype, layout, direction
    (720, 1280, 1280, 4, 3, 32, np.uint8, "uint16", "ref", "r2l"),     # 14.7 MB of census words: most of the array
    (61, 113, 97, 3, 2, 20, np.float32, "float32", "c", "l2r"),         # ragged sizes, float images, the other direction
    (720, 1280, 1280, 4, 3, 32, np.uint8, "uint16", "ref", "r2l"),     # the first case again, over the second one's leftovers
    (200, 320, 320, 5, 5, 48, np.uint8, "int32", "c", "r2l"),          # 11x11: four words per pixel
]


def test_calls_in_a_row_equal_the_regular_entry_point(api, tctx):
    import torch
    before_ws = tctx.workspace_bytes
    fresh = api.Context(0, torch_stream=True)   # only ever sees the no-allocation entry point
    try:
        for n, (H, Wl, Wr, hr, vr, D, idt, vdt, layout, direction) in enumerate(CASES):
            L, R, _ = synth.stereo_pair(H, max(Wl, Wr), 40 + n, d_bg=3, d_sq=9, d_max=0, dtype=idt)
            L, R = _dev(L[:, :Wl]), _dev(R[:, :Wr])
            dcode = api.RIGHT_TO_LEFT if direction == "r2l" else api.LEFT_TO_RIGHT
            Ws = Wr if direction == "r2l" else Wl
            want = _host(_regular(api, tctx, L, R, hr, vr, D, dcode, _volume(torch, H, Ws, D, vdt, layout, 7)))
            assert want.max() > 0
            alloc0 = torch.cuda.memory_allocated()
            for rep in range(2):   # twice in a row on the same operands, then on to the next case
                got = _volume(torch, H, Ws, D, vdt, layout, 9 + rep) if rep == 0 else got
                if rep:
                    (got.view(torch.int16) if vdt == "uint16" else got).fill_(11)
                torch.cuda.synchronize()
                free_a = torch.cuda.mem_get_info()[0]
                fresh.check(_noalloc(api, fresh, L, R, hr, vr, D, dcode, got))
                torch.cuda.synchronize()
                free_b = torch.cuda.mem_get_info()[0]
                # the first call of a configuration may load kernels; the second must leave the device's memory alone
                assert rep == 0 or free_b >= free_a, (n, free_a, free_b)
                assert np.array_equal(_host(got), want), (n, rep)
            del got
            assert torch.cuda.memory_allocated() == alloc0
        assert fresh.workspace_bytes == 0          # the context's arena was never touched
    finally:
        fresh.close()
    assert tctx.workspace_bytes >= before_ws


def test_corrected_semantics_and_python_mirror(api, tctx):
    import torch
    L, R, _ = synth.stereo_pair(90, 160, 77, d_bg=2, d_sq=6)
    L, R = _dev(L), _dev(R)
    for sem in (api.COMPAT, api.CORRECTED):
        want = api.unfoldBasedCostVolume(api.CENSUS, L, R, 4, 3, 24, semantics=sem, layout="c", ctx=tctx)
        out = torch.zeros((90, 160, 24), dtype=torch.int16, device="cuda").view(torch.uint16)
        got = api.censusCostVolumeNoAlloc(L, R, 4, 3, 24, out, semantics=sem, ctx=tctx)
        assert got is out
        torch.cuda.synchronize()
        assert np.array_equal(_host(got), _host(want))


def test_sizes_beyond_the_static_array_are_refused(api, tctx):
    import torch
    H, W, D = 1080, 1920, 16
    L = _dev(synth.stereo_pair(H, W, 5)[0])
    out = torch.zeros((H, W, D), dtype=torch.int16, device="cuda").view(torch.uint16)
    rc = _noalloc(api, tctx, L, L, 4, 3, D, api.RIGHT_TO_LEFT, out)   # 2 x 1080 x 1920 x 2 words x 4 B = 33 MB > 16 MiB
    assert rc == api.E_UNSUPPORTED
    msg = api.lib.stevi_last_error(tctx.handle).decode()
    assert "STEVI_NOALLOC_SCRATCH_BYTES" in msg and str(api.NOALLOC_SCRATCH_BYTES) in msg
    # one word per pixel (3x3 .. 5x5 windows) brings the same pair under the limit: 2 x 8.3 MB. (CORRECTED: the reference's
    # census leaves a partly filled word out, Q1, so COMPAT 5x5 costs are all zero.)
    want = _host(_regular(api, tctx, L, L, 2, 2, D, api.RIGHT_TO_LEFT,
                          torch.zeros((H, W, D), dtype=torch.int16, device="cuda").view(torch.uint16), api.CORRECTED))
    assert want.max() > 0
    out2 = torch.zeros((H, W, D), dtype=torch.int16, device="cuda").view(torch.uint16)
    assert _noalloc(api, tctx, L, L, 2, 2, D, api.RIGHT_TO_LEFT, out2, api.CORRECTED) == api.OK
    torch.cuda.synchronize()
    assert np.array_equal(_host(out2), want)


def test_host_operands_and_timing_are_refused(api, tctx):
    import torch
    L = np.zeros((32, 48), np.uint8)
    Ld = torch.zeros((32, 48), dtype=torch.uint8, device="cuda")
    out = torch.zeros((32, 48, 8), dtype=torch.int16, device="cuda").view(torch.uint16)
    assert _noalloc(api, tctx, L, Ld, 1, 1, 8, api.RIGHT_TO_LEFT, out) == api.E_ARG
    assert "device memory only" in api.lib.stevi_last_error(tctx.handle).decode()
    assert _noalloc(api, tctx, Ld, Ld, 1, 1, 8, api.RIGHT_TO_LEFT, np.zeros((32, 48, 8), np.uint16)) == api.E_ARG
    tctx.set_timing(True)
    try:
        assert _noalloc(api, tctx, Ld, Ld, 1, 1, 8, api.RIGHT_TO_LEFT, out) == api.E_UNSUPPORTED
    finally:
        tctx.set_timing(False)
    assert _noalloc(api, tctx, Ld, Ld, 1, 1, 8, api.RIGHT_TO_LEFT, out) == api.OK
    torch.cuda.synchronize()
