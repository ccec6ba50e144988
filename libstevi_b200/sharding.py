"""Host-side partitioning of the dense-stereo path over the GPUs of one box.

* independent pairs (BASELINE configs[4]): pair k -> rank k mod G, no data-path collective;
* one large frame (configs[2], [3]): contiguous ROW BANDS. Every stage of local matching is row-local given a
  v_radius halo of input rows; the coarse-to-fine driver needs, per level, the rows its bilinear guide upsampling and
  its 2x2 pooling reach, all indexed in GLOBAL coordinates (hierarchical.h:116,132; downsampling.h:99) — `band_plan`
  derives those row ranges and `hierarchical_rows` runs them. Bands are gathered with one all_gather (NCCL on the GPUs,
  gloo in the CPU tests); nothing else crosses ranks.

Only plans, slicing and collectives live here; all arithmetic is in the CUDA library behind `libstevi_b200.api`.
"""
from collections import namedtuple

import numpy as np

LevelPlan = namedtuple("LevelPlan", "height out_lo out_hi img_lo img_hi")

# The float cost-volume kernel keeps running window sums per block of 32 image rows anchored at global multiples of 32
# (costvol_box.cuh: BOX_BAND_ROWS, stevi_ctx_set_row_origin). A band reproduces the whole frame's rounding bit for bit
# when it also holds the rows from the start of the 32-row block its first output row falls into.
KERNEL_BAND_ROWS = 32


def cv_rows(g0, g1, v_radius, height):
    """Input rows [lo, hi) from which rows [g0, g1) of a float cost volume come out bit-identical to the whole frame's."""
    lo, hi = halo_rows(g0, g1, v_radius, height)
    return min(lo, max(0, g0 // KERNEL_BAND_ROWS * KERNEL_BAND_ROWS - v_radius)), hi


def shard_pairs(n_pairs, world, rank):
    """Indices of the pairs rank `rank` owns: k with k mod world == rank."""
    return list(range(rank, n_pairs, world))


def row_bands(height, world, align=1):
    """[(g0, g1)] per rank: contiguous, near-equal row bands covering 0..height, starts aligned to `align` rows."""
    edges = [min(height, ((height * r // world) + align - 1) // align * align) for r in range(world)] + [height]
    return [(edges[r], max(edges[r], edges[r + 1])) for r in range(world)]


def halo_rows(g0, g1, halo, height):
    """Input rows [lo, hi) a band of output rows [g0, g1) needs for windows of +-halo rows."""
    return max(0, g0 - halo), min(height, g1 + halo)


def pyramid_heights(height, depth):
    """Heights of levels 0 (coarsest) .. depth (the input): ceil-halving like averagePoolingDownsample(2)."""
    hs = [height]
    for _ in range(depth):
        hs.append((hs[-1] + 1) // 2)
    return hs[::-1]


def band_plan(height, depth, v_radii, g0, g1):
    """Row ranges per pyramid level needed to produce rows [g0, g1) of the finest level.
    out_lo/out_hi: rows of that level's result (guide for the next finer level) that are needed;
    img_lo/img_hi: image rows of that level that must be present (own windows + what the coarser levels pool from)."""
    hs = pyramid_heights(height, depth)
    out = [None] * (depth + 1)
    out[depth] = (g0, g1)
    for k in range(depth, 0, -1):  # guide rows level k reads from level k-1 (same bound as the library's argument check)
        a, b = out[k]
        hg, h = hs[k - 1], hs[k]
        v_lo = max(0, a * (hg - 1) // (h - 1) - 1)
        v_hi = min(hg - 1, (b - 1) * (hg - 1) // (h - 1) + 2)
        out[k - 1] = (v_lo, v_hi + 1)
    img = [None] * (depth + 1)
    for k in range(depth + 1):  # coarsest first: a level must hold the pre-image of every coarser level's band
        lo, hi = halo_rows(out[k][0], out[k][1], v_radii[k], hs[k])
        if k == 0:  # the full cost volume of the coarsest level: see cv_rows
            lo, hi = cv_rows(out[0][0], out[0][1], v_radii[0], hs[0])
        if k > 0:
            lo, hi = min(lo, 2 * img[k - 1][0]), max(hi, min(2 * img[k - 1][1], hs[k]))
        img[k] = (lo - lo % 2 if k > 0 else lo, hi)  # even start: 2x2 pooling windows are anchored at global row 0
    # the pooled band of level k starts at img[k+1].lo / 2, which may lie above img[k].lo: widen coarse bands accordingly
    for k in range(depth - 1, -1, -1):
        img[k] = (img[k + 1][0] // 2, (img[k + 1][1] + 1) // 2)
    return [LevelPlan(hs[k], out[k][0], out[k][1], img[k][0], img[k][1]) for k in range(depth + 1)]


def hierarchical_rows(api, matchFunc, depth, img_l, img_r, h_radius, v_radius, disp_width, g0, g1, upscale_disp_radius=2,
                      dDir=1, band_row0=0, full_height=None, ctx=None):
    """Rows [g0, g1) of hiearchicalTruncatedCostVolume on a frame of `full_height` rows. img_l / img_r hold the frame's rows
    from global row `band_row0` on (the whole frame by default); they must cover band_plan(...)[-1].img_lo..img_hi.
    Returns (tcv, disp) for those rows, identical to the rows of the whole-frame call."""
    H = int(full_height if full_height is not None else img_l.shape[0])
    hr = [int(h_radius)] * (depth + 1) if np.isscalar(h_radius) else [int(v) for v in h_radius]
    vr = [int(v_radius)] * (depth + 1) if np.isscalar(v_radius) else [int(v) for v in v_radius]
    plan = band_plan(H, depth, vr, g0, g1)
    fine = plan[depth]
    if fine.img_lo < band_row0 or fine.img_hi > band_row0 + img_l.shape[0]:
        raise ValueError(f"image rows {band_row0}..{band_row0 + img_l.shape[0]} do not cover the needed {fine.img_lo}..{fine.img_hi}")
    ls = [None] * (depth + 1)
    rs = [None] * (depth + 1)
    ls[depth] = img_l[fine.img_lo - band_row0:fine.img_hi - band_row0]
    rs[depth] = img_r[fine.img_lo - band_row0:fine.img_hi - band_row0]
    for k in range(depth - 1, -1, -1):
        ls[k] = api.averagePoolingDownsample(ls[k + 1], 2, ctx=ctx)
        rs[k] = api.averagePoolingDownsample(rs[k + 1], 2, ctx=ctx)
        assert ls[k].shape[0] == plan[k].img_hi - plan[k].img_lo
    widths = [disp_width]
    for _ in range(depth):
        widths.append((widths[-1] + 1) // 2)
    widths = widths[::-1]
    p0 = plan[0]
    ctx = ctx if ctx is not None else api.default_context(img_l.device.index if hasattr(img_l, "data_ptr") else 0)
    ctx.set_row_origin(p0.img_lo)
    try:
        cv = api.unfoldBasedCostVolume(matchFunc, ls[0], rs[0], hr[0], vr[0], widths[0], dDir=dDir, ctx=ctx)
    finally:
        ctx.set_row_origin(0)
    idx = api.extractSelectedIndex(cv, api.strategy_of(matchFunc), ctx=ctx)
    guide = idx[p0.out_lo - p0.img_lo:p0.out_hi - p0.img_lo]
    tcv = None
    for k in range(1, depth + 1):
        p, q = plan[k], plan[k - 1]
        tcv, guide = api.computeGuidedCVBand(matchFunc, ls[k], rs[k], hr[k], vr[k], guide, p.img_lo, p.height, p.out_lo,
                                             p.out_hi - p.out_lo, q.out_lo, q.height, upscale_disp_radius, dDir, ctx=ctx)
    return tcv, guide


def local_matching_rows(api, matchFunc, img_l, img_r, h_radius, v_radius, disp_width, g0, g1, kernel=1, dDir=1, ctx=None):
    """Rows [g0, g1) of localMatchingPipeline (cost volume -> WTA -> refinement) from the band plus its v_radius halo.
    Returns (idx, refined) for those rows."""
    H = img_l.shape[0]
    lo, hi = cv_rows(g0, g1, v_radius, H)
    ctx = ctx if ctx is not None else api.default_context(img_l.device.index if hasattr(img_l, "data_ptr") else 0)
    ctx.set_row_origin(lo)  # anchors the cost-volume kernel's row blocks: the band's rows round like the whole frame's
    try:
        _, idx, _, refined = api.localMatchingPipeline(matchFunc, img_l[lo:hi], img_r[lo:hi], h_radius, v_radius, disp_width,
                                                       kernel, dDir, return_all=True, ctx=ctx)
    finally:
        ctx.set_row_origin(0)
    # the validity tests of truncatedCostVolume (`i < v_radius or i + v_radius >= H`, correlation_base.h:606) look at the
    # frame border: inside the frame the halo rows take their place, at the frame border band and frame coincide
    return idx[g0 - lo:g1 - lo], refined[g0 - lo:g1 - lo]


def gather_rows(local, bands, dist, group=None):
    """all_gather of per-rank row bands (torch tensors, first axis = rows of unequal count) into the full array."""
    import torch
    world = dist.get_world_size(group)
    rows = max(b[1] - b[0] for b in bands)
    if all(b[1] - b[0] == rows for b in bands):
        # equal bands: the ranks' rows land in place in one output tensor (no padding, no concatenation)
        out = torch.empty((world * rows,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        dist.all_gather_into_tensor(out, local.contiguous(), group=group)
        return out
    pad = torch.zeros((rows,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[:local.shape[0]] = local
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad, group=group)
    return torch.cat([parts[r][:bands[r][1] - bands[r][0]] for r in range(world)], 0)


def max_over_ranks(value, dist, device="cpu", group=None):
    """Max of a per-rank scalar (timings are reported as the slowest rank's)."""
    import torch
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())
