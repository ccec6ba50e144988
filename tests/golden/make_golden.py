"""Generates tests/golden/*.npz from the UNMODIFIED reference headers (oracle/_ref/libstevi_ref.so, built from
/root/reference by oracle/Makefile). Run in the build container only (the GPU box has no reference tree):

    python tests/golden/make_golden.py

Inputs are produced by libstevi_b200.synth (counter-based, library independent) from the seeds stored in each file,
so the fixtures hold the reference OUTPUTS plus the parameters needed to regenerate the inputs.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import oracle as O  # noqa: E402
from libstevi_b200 import synth  # noqa: E402


def main():
    ref = O.ref
    assert ref is not None, "build oracle/_ref first (make -C oracle ref)"
    H, W, D = 26, 44, 10
    L8, R8, gt = synth.stereo_pair(H, W, 101, d_bg=2, d_sq=5, d_max=6)
    Lf, Rf, _ = synth.stereo_pair(H, W, 102, d_bg=2, d_sq=5, d_max=6, dtype=np.float32)
    out = {"H": H, "W": W, "D": D, "seed_u8": 101, "seed_f32": 102}

    # census features and census cost volumes (u8 and f32 images, several windows, both directions)
    for hr, vr in ((4, 3), (3, 3), (2, 2), (6, 5)):
        out[f"census_u8_{hr}_{vr}"] = ref.census(L8, hr, vr)
        out[f"census_f32_{hr}_{vr}"] = ref.census(Lf, hr, vr)
        for direction in (0, 1):
            out[f"censuscv_u8_{hr}_{vr}_{direction}"] = ref.cost_volume(O.CENSUS, L8, R8, hr, vr, D, direction)
    cv = ref.cost_volume(O.CENSUS, L8, R8, 4, 3, D)
    # SGM on the census volume: all direction sets, both strategies, with and without margins
    for ndir in (4, 8, 16):
        for strategy in (0, 1):
            out[f"sgm_u16_{ndir}_{strategy}"] = ref.sgm(cv, ndir, strategy, 10, 120, 100)
    out["sgm_u16_8_0_margins"] = ref.sgm(cv, 8, 0, 10, 120, 100, (2, 1, 3, 2))
    out["wta_cost"] = ref.wta(out["sgm_u16_8_0"], 0)
    out["wta_cv_u16"] = ref.wta(cv, 0)
    # float cost volumes for every float matching function
    for mf in (O.CC, O.NCC, O.SSD, O.SAD, O.ZCC, O.ZNCC, O.ZSSD, O.ZSAD):
        for direction in (0, 1):
            out[f"cv_f32_{mf}_{direction}"] = ref.cost_volume(mf, Lf, Rf, 2, 2, D, direction)
    zn = out[f"cv_f32_{O.ZNCC}_1"]
    out["sgm_f32_8_1"] = ref.sgm(zn, 8, 1, 0.001, 0.01, 100)
    out["sgm_f32_8_0"] = ref.sgm(out[f"cv_f32_{O.SSD}_1"], 8, 0, 0.5, 2.0, 100)
    idx = ref.wta(zn, 1)
    out["wta_score"] = idx
    out["disp"] = ref.index_to_disp(idx, 1, 3)
    out["selected_cost"] = ref.selected_cost(zn, idx)
    for tdir in (0, 1, 2):
        out[f"tcv_{tdir}"] = ref.truncated_cost_volume(zn, idx, 2, 2, 1, 1, tdir)
    for kernel in (0, 1, 2):
        out[f"refined_{kernel}"] = ref.refine(out["tcv_0"], idx, kernel)
    out["avgpool"] = ref.avgpool2(Lf)
    out["avgpool_odd"] = ref.avgpool2(np.ascontiguousarray(Lf[:25, :43]))
    for depth in (1, 2):
        tcv, disp = ref.hierarchical(O.ZNCC, depth, Lf, Rf, 2, 2, D, 2)
        out[f"hier_tcv_{depth}"], out[f"hier_disp_{depth}"] = tcv, disp
    out["unfold_u8"] = ref.unfold(L8[:6, :7], 2, 1)
    out["features_zncc"] = ref.feature_volume(O.ZNCC, Lf[:8, :9], 1, 1)
    np.savez_compressed(os.path.join(HERE, "reference_outputs.npz"), **out)
    print("wrote", os.path.join(HERE, "reference_outputs.npz"), os.path.getsize(os.path.join(HERE, "reference_outputs.npz")), "bytes")


def main_2d():
    """2-D search: unfoldBased2dDisparityCostVolume, extractSelected2dIndex, selected2dIndexToDisp,
    truncatedBidirectionaCostVolume, refineDisp2dCostInterpolation."""
    ref = O.ref
    assert ref is not None, "build oracle/_ref first (make -C oracle ref)"
    H, W = 11, 15
    search = (-1, 2, -3, 3)
    Lf, Rf = synth.random_image(H, W, 201), synth.random_image(H, W, 202)
    out = {"H": H, "W": W, "seed_l": 201, "seed_r": 202, "search": np.asarray(search, np.int32)}
    for mf in (O.CC, O.NCC, O.SSD, O.SAD, O.ZCC, O.ZNCC, O.ZSSD, O.ZSAD):
        for direction in (0, 1):
            out[f"cv2d_{mf}_{direction}"] = ref.cost_volume_2d(mf, Lf, Rf, 2, 1, search, direction)
    for mf, strategy in ((O.ZNCC, 1), (O.SSD, 0)):
        idx = ref.wta_2d(out[f"cv2d_{mf}_1"], strategy)
        out[f"wta2d_{mf}"] = idx
        out[f"disp2d_{mf}"] = ref.index2d_to_disp(idx, search)
    # truncatedBidirectionaCostVolume + refineDisp2dCostInterpolation (all kernels, both isotropy hypotheses)
    for mf in (O.ZNCC, O.SSD):
        for r0, r1 in ((1, 1), (2, 1), (1, 0)):
            t = ref.truncated_cost_volume_2d(out[f"cv2d_{mf}_1"], out[f"wta2d_{mf}"], r0, r1)
            out[f"tcv2d_{mf}_{r0}_{r1}"] = t
            for kernel in (O.EQUIANGULAR, O.PARABOLA, O.GAUSSIAN):
                for aniso in (0, 1):
                    out[f"refined2d_{mf}_{r0}_{r1}_{kernel}_{aniso}"] = ref.refine_2d(t, out[f"disp2d_{mf}"], kernel, aniso)
    # variableCostVolumeAggregation on a slice of the SSD volume with bilateral weights (tests/test_oracle_pinning.bilateral_windows)
    sys.path.insert(0, os.path.dirname(HERE))
    from test_oracle_pinning import bilateral_windows
    out["aggregated_ssd_5x5"] = ref.variable_cost_volume_aggregation(np.ascontiguousarray(out[f"cv2d_{O.SSD}_1"][:, :, :, 2]),
                                                                      bilateral_windows(Rf, 2, 2))
    path = os.path.join(HERE, "reference_outputs_2d.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


def main_filters():
    """meanFilter2D / sigmaFilter (correlation_base.h:881-1077) on (H,W) and (H,W,C) u8|f32 images. Only the interior the
    reference writes is stored; the rest of each array is NaN."""
    ref = O.ref
    assert ref is not None, "build oracle/_ref first (make -C oracle ref)"
    H, W = 23, 31
    out = {"H": H, "W": W, "seed": 301}
    imgs = {"f32": synth.random_image(H, W, 301),
            "u8": synth.stereo_pair(H, W, 302)[0],
            "f32c3": np.stack([synth.random_image(H, W, 303 + k) for k in range(3)], axis=2),
            "u8c2": np.stack([synth.stereo_pair(H, W, 306 + k)[0] for k in range(2)], axis=2)}
    for name, img in imgs.items():
        for hr, vr in ((1, 1), (3, 2), (0, 4), (5, 0)):
            m = ref.mean_filter_2d(img, hr, vr)
            out[f"mean_{name}_{hr}_{vr}"] = m
            out[f"sigma_{name}_{hr}_{vr}"] = ref.sigma_filter(m, img, hr, vr)
    path = os.path.join(HERE, "reference_outputs_filters.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


def main_refinement():
    """Feature-domain refinement (image_based_refinement.h), on-demand truncated volumes and the deterministic part of PatchMatch
    (on_demand_cost_volume.h, patchmatch.h) from oracle/_ref/libstevi_ref_ibr.so: the reference's unmodified code; the Eigen calls
    inside the refinements run on the stand-in solver of oracle/small_qr.h (see oracle/shims/Eigen)."""
    ref = O.ref_ibr
    assert ref is not None, "build oracle/_ref first (make -C oracle ref)"
    H, W = 22, 56
    rng = np.random.default_rng(401)
    base = rng.uniform(-1, 1, (H, W + 16)).astype(np.float32)
    k = np.array([1, 2, 1], np.float32) / 4
    base = np.apply_along_axis(lambda r: np.convolve(r, k, mode="same"), 1, base).astype(np.float32)
    left, right = base[:, 4:4 + W].copy(), (0.6 * base[:, 8:8 + W] + 0.4 * base[:, 9:9 + W]).astype(np.float32)
    out = {"H": H, "W": W, "left": left, "right": right}
    fl_raw, fr_raw = O.port.unfold(left, 2, 1), O.port.unfold(right, 2, 1)
    for name in ("NCC", "ZNCC", "SSD", "ZSSD", "SAD", "ZSAD"):
        mf = getattr(O, name)
        fl, fr = O.port.features_for_matchfunc(mf, fl_raw), O.port.features_for_matchfunc(mf, fr_raw)
        cv = O.port.aggregate_cost(mf, fl, fr, 0, 9, O.R2L)
        idx = O.port.wta(cv, O.strategy_of(mf))
        out[f"idx_{name}"] = idx
        out[f"bary_{name}"] = ref.refine_barycentric(mf, fl, fr, idx)
        out[f"costsym_{name}"] = ref.refine_cost_symmetric(mf, fl, fr, idx, cv)
        if name not in ("SAD", "ZSAD"):
            out[f"barysym1_{name}"] = ref.refine_barycentric_symmetric(mf, 1, fl, fr, idx, 10)
            out[f"barysym2_{name}"] = ref.refine_barycentric_symmetric(mf, 2, fl, fr, idx, 10)
    init = rng.integers(-1, 12, (H, W, 1)).astype(np.int32)
    disp = rng.integers(-3, 12, (H, W)).astype(np.int32)
    out["pm_init"], out["od_disp"] = init, disp
    for name in ("ZNCC", "SSD", "ZSAD"):
        mf = getattr(O, name)
        for n_iter in (1, 2, 4):
            out[f"pm_{name}_{n_iter}"] = ref.patch_match(mf, fr_raw, fl_raw, (0, 9), n_iter, 0, init)
        out[f"od_{name}"] = ref.on_demand_truncated_cost_volume(mf, fr_raw, fl_raw, (-2, 9), disp, 1)
    path = os.path.join(HERE, "reference_outputs_refinement.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    if "refinement" in sys.argv[1:]:
        main_refinement()
    elif "2d" in sys.argv[1:]:
        main_2d()
    elif "filters" in sys.argv[1:]:
        main_filters()
    else:
        main()
        main_2d()
        main_filters()
        main_refinement()
