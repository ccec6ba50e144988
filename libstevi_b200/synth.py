"""Deterministic synthetic rectified stereo pairs (random texture, known disparity field).

Same idea as the reference's ``generateParallaxSquareImage`` (test/test_correlation_utils.h:324-391): a random
right image, a background shifted by ``d_bg``, a centred square shifted by ``d_sq`` and, optionally, a horizontal
disparity ramp so that many disparities occur. The generator is counter-based (SplitMix64 on
``(seed, i*W + j)``), so it does not depend on any library's RNG. Calling convention ``f(L, R)`` with
``dispDirection::RightToLeft`` gives expected index == ground truth.
"""
import numpy as np

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(seed, n, stream=0):
    """n 64-bit values: SplitMix64 finaliser applied to (seed, stream, counter)."""
    with np.errstate(over="ignore"):
        z = np.arange(n, dtype=np.uint64)
        z = z + np.uint64(seed) * np.uint64(0x9E3779B97F4A7C15) + np.uint64(stream) * np.uint64(0xD1B54A32D192ED03)
        z = z + np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return z


def stereo_pair(h, w, seed, d_bg=8, d_sq=32, d_max=0, dtype=np.uint8):
    """Returns (left, right, gt) with gt(i,j) the disparity of right pixel (i,j): left(i, j+gt) == right(i,j)."""
    right = (splitmix64(seed, h * w, 0) >> np.uint64(56)).astype(np.uint8).reshape(h, w)
    left = (splitmix64(seed, h * w, 1) >> np.uint64(56)).astype(np.uint8).reshape(h, w)  # unfilled pixels: noise
    jj = np.arange(w)
    gt = np.full((h, w), d_bg, np.int32)
    if d_max > 0:
        gt += (jj * d_max // (2 * w)).astype(np.int32)[None, :]
    side = min(h, w) // 3
    i0, j0 = (h - side) // 2, (w - side) // 2
    sq = np.zeros((h, w), bool)
    sq[i0:i0 + side, j0:j0 + side] = True
    ii = np.arange(h)[:, None] * np.ones((1, w), np.int64)
    for mask, g in ((~sq, gt), (sq, np.full((h, w), d_sq, np.int32))):  # background first, then the square
        tj = jj[None, :] + g
        ok = mask & (tj < w)
        left[ii[ok], tj[ok]] = right[ok]
    gt = np.where(sq, d_sq, gt).astype(np.int32)
    if dtype == np.uint8:
        return left, right, gt
    return (left.astype(np.float32) / np.float32(255.0)), (right.astype(np.float32) / np.float32(255.0)), gt


def random_image(h, w, seed, dtype=np.float32, stream=7):
    """Uniform noise image: u8 in 0..255 or float32 in [-1, 1) (test/unittests/testCorrelationFilters.cpp:76-79)."""
    bits = splitmix64(seed, h * w, stream)
    if dtype == np.uint8:
        return (bits >> np.uint64(56)).astype(np.uint8).reshape(h, w)
    u = (bits >> np.uint64(40)).astype(np.float32) / np.float32(1 << 24)
    return (u * np.float32(2.0) - np.float32(1.0)).reshape(h, w)
