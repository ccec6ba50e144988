"""meanFilter2D / sigmaFilter (correlation_base.h:881-1077) on the GPU against the CPU oracle: the interior the reference
defines bit for bit (the running sums are walked in its order), the border 0; against the compiled reference when present
(the sigma of (H,W,C) float images within 4e-7 relative: g++ contracts that one instantiation into fused multiply-adds);
the reference's own unit tests (testCorrelationFilters.cpp:82-136, :172-221) restated; the committed golden outputs."""
import os

import numpy as np
import pytest

from libstevi_b200 import synth

pytestmark = pytest.mark.gpu

GF = np.load(os.path.join(os.path.dirname(__file__), "golden", "reference_outputs_filters.npz"))


def image(shape, dtype, seed):
    h, w = shape[:2]
    c = shape[2] if len(shape) == 3 else None
    planes = [synth.random_image(h, w, seed + k, dtype=dtype) for k in range(c or 1)]
    return np.stack(planes, axis=2) if c else planes[0]


def interior_equal(got, want, hr, vr):
    h, w = want.shape
    inner = (slice(vr, h - vr), slice(hr, w - hr))
    assert np.array_equal(got[inner], want[inner]), np.nanmax(np.abs(got[inner] - want[inner]))
    border = np.ones((h, w), bool)
    border[inner] = False
    assert (got[border] == 0).all()


@pytest.mark.parametrize("shape", [(40, 57), (67, 130), (33, 41, 3), (21, 30, 1), (35, 70, 40)])
@pytest.mark.parametrize("dtype", [np.float32, np.uint8])
@pytest.mark.parametrize("hr,vr", [(0, 0), (1, 1), (3, 2), (0, 4), (6, 0), (9, 9)])
def test_filters_equal_the_oracle(api, ctx, oracle, shape, dtype, hr, vr):
    img = image(shape, dtype, 11)
    mean = api.meanFilter2D(hr, vr, img, ctx=ctx)
    want = oracle.port.mean_filter_2d(img, hr, vr)
    interior_equal(mean, want, hr, vr)
    sigma = api.sigmaFilter(hr, vr, mean, img, ctx=ctx)
    interior_equal(sigma, oracle.port.sigma_filter(mean, img, hr, vr), hr, vr)
    if oracle.ref is not None:
        interior_equal(mean, oracle.ref.mean_filter_2d(img, hr, vr), hr, vr)
        rs = oracle.ref.sigma_filter(mean, img, hr, vr)
        inner = (slice(vr, shape[0] - vr), slice(hr, shape[1] - hr))
        if dtype == np.float32 and len(shape) == 3:
            assert np.abs(sigma[inner] - rs[inner]).max() <= 4e-7 * np.abs(rs[inner]).max()
        else:
            assert np.array_equal(sigma[inner], rs[inner])


def test_full_frame_and_strided_operands(api, ctx, oracle):
    import torch
    H, W = 1080, 1920
    img = synth.random_image(H, W, 5)
    want = oracle.port.mean_filter_2d(img, 4, 4)
    t = torch.from_numpy(img).cuda()
    tc = api.Context(0, torch_stream=True)
    try:
        mean = api.meanFilter2D(4, 4, t, ctx=tc)
        sigma = api.sigmaFilter(4, 4, mean, t, ctx=tc)
        interior_equal(mean.cpu().numpy(), want, 4, 4)
        interior_equal(sigma.cpu().numpy(), oracle.port.sigma_filter(want, img, 4, 4), 4, 4)
        # a column-major image and a channel-first multi-channel image: any strides are taken as they are
        tf = torch.from_numpy(np.asfortranarray(img[:300, :500])).cuda()
        sub = np.ascontiguousarray(img[:300, :500])
        interior_equal(api.meanFilter2D(2, 3, tf, ctx=tc).cpu().numpy(), oracle.port.mean_filter_2d(sub, 2, 3), 2, 3)
        chw = torch.from_numpy(image((3, 64, 90), np.float32, 21)).cuda()    # planes first in memory, viewed (H,W,C)
        hwc = chw.permute(1, 2, 0)
        hwc_np = np.ascontiguousarray(hwc.cpu().numpy())
        m = api.meanFilter2D(3, 1, hwc, ctx=tc)
        interior_equal(m.cpu().numpy(), oracle.port.mean_filter_2d(hwc_np, 3, 1), 3, 1)
        interior_equal(api.sigmaFilter(3, 1, m, hwc, ctx=tc).cpu().numpy(), oracle.port.sigma_filter(m.cpu().numpy(), hwc_np, 3, 1), 3, 1)
    finally:
        tc.close()


@pytest.mark.parametrize("hr,vr,eh,ev", [(0, 0, 0, 0), (1, 1, 1, 1), (5, 1, 7, 7), (1, 5, 7, 7), (5, 5, 7, 7)])
def test_reference_unit_tests(api, ctx, hr, vr, eh, ev):
    # testCorrelationFilters.cpp:82-136: the last interior pixel against the plain window mean; :172-221 the sigma of one window
    h, w = 2 * vr + 1 + ev, 2 * hr + 1 + eh
    img = synth.random_image(h, w, 100 + 7 * hr + vr)
    avg = api.meanFilter2D(hr, vr, img, ctx=ctx)
    assert abs(img[ev:, eh:].sum(dtype=np.float32) / ((2 * hr + 1) * (2 * vr + 1)) - avg[vr + ev, hr + eh]) < 1e-3
    one = np.ascontiguousarray(img[:2 * vr + 1, :2 * hr + 1])
    m = api.meanFilter2D(hr, vr, one, ctx=ctx)
    s = api.sigmaFilter(hr, vr, m, one, ctx=ctx)
    assert abs(np.sqrt(((one - one.mean(dtype=np.float32)) ** 2).sum(dtype=np.float32)) - s[vr, hr]) < 1e-3


@pytest.mark.parametrize("name", ["f32", "u8", "f32c3", "u8c2"])
@pytest.mark.parametrize("hr,vr", [(1, 1), (3, 2), (0, 4), (5, 0)])
def test_golden(api, ctx, name, hr, vr):
    h, w = int(GF["H"]), int(GF["W"])
    img = {"f32": lambda: synth.random_image(h, w, 301), "u8": lambda: synth.stereo_pair(h, w, 302)[0],
           "f32c3": lambda: np.stack([synth.random_image(h, w, 303 + k) for k in range(3)], axis=2),
           "u8c2": lambda: np.stack([synth.stereo_pair(h, w, 306 + k)[0] for k in range(2)], axis=2)}[name]()
    want = GF[f"mean_{name}_{hr}_{vr}"]
    inner = (slice(vr, h - vr), slice(hr, w - hr))
    mean = api.meanFilter2D(hr, vr, img, ctx=ctx)
    assert np.array_equal(mean[inner], want[inner])
    sigma, ws = api.sigmaFilter(hr, vr, mean, img, ctx=ctx), GF[f"sigma_{name}_{hr}_{vr}"]
    if name == "f32c3":
        assert np.abs(sigma[inner] - ws[inner]).max() <= 4e-7 * ws[inner].max()
    else:
        assert np.array_equal(sigma[inner], ws[inner])


def test_windows_that_do_not_fit_are_refused(api, ctx):
    img = synth.random_image(8, 9, 1)
    assert api.meanFilter2D(5, 1, img, ctx=ctx) is None      # 11 columns needed
    assert api.meanFilter2D(1, 4, img, ctx=ctx) is None      # 9 rows needed
    assert api.meanFilter2D(4, 3, img, ctx=ctx) is not None  # 9 x 7: one interior column, two interior rows
    assert api.sigmaFilter(5, 1, np.zeros((8, 9), np.float32), img, ctx=ctx) is None
