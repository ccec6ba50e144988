"""The C++ facade (include/StereoVision/...): same names, template parameters and defaults as the reference's headers,
bodies forwarding to the C ABI. CPU part: the demo translation unit compiles with plain g++ -std=c++17 and links against
the shared library (no nvcc, no torch). GPU part: the demo's results equal the Python mirror's on the same inputs."""
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "libstevi_b200")


def build_demo(tmp_path):
    exe = os.path.join(str(tmp_path), "facade_demo")
    cmd = ["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "facade_demo.cpp"), "-o", exe, "-L" + LIBDIR, "-lstevi_b200", "-Wl,-rpath," + LIBDIR]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    return exe


def test_facade_compiles_and_links(tmp_path):
    build_demo(tmp_path)


def test_facade_headers_carry_the_reference_names():
    inc = os.path.join(ROOT, "include", "StereoVision")
    text = "".join(open(os.path.join(dp, f)).read() for dp, _, fs in os.walk(inc) for f in fs)
    for name in ("unfoldBasedCostVolume", "sgmCostVolume", "extractSelectedIndex", "selectedIndexToDisp", "selectedCost",
                 "truncatedCostVolume", "extractInBoundDomain", "refineDispCostInterpolation", "hiearchicalTruncatedCostVolume",
                 "OffsetedCostVolume", "averagePoolingDownsample", "DownSampleWindows", "censusTransform2D", "censusFeatures",
                 "UnFoldCompressor", "channelFromCord", "dispExtractionStartegy", "truncatedCostVolumeDirection",
                 "matchingFunctions", "MatchingFunctionTraits", "hamming_cv_t", "census_data_t", "InterpolationKernel",
                 "sgmDirections", "PaddingMargins", "disp_t", "searchOffset", "unfoldBased2dDisparityCostVolume",
                 "extractSelected2dIndex", "selected2dIndexToDisp", "truncatedBidirectionaCostVolume", "refineDisp2dCostInterpolation",
                 "IsotropyHypothesis", "GrPix17R3Filter", "GrPix17R4Filter", "aggregateCost", "featureVolume2CostVolume",
                 "getFeatureVolumeForMatchFunc", "buildFeaturePyramid", "FeatureVolumePyramid", "meanFilter2D", "sigmaFilter",
                 "refineBarycentricDisp", "refineBarycentricSymmetricDisp", "refineBarycentric2dDisp", "refineBarycentricSymmetric2dDisp",
                 "refineCostSymmetricDisp", "refinedBarycentricDisp", "refinedBarycentricDispFeatureVol", "refinedCostSymmetricDisp",
                 "bidimensionalContiguity", "computeAggregationWindows", "variableCostVolumeAggregation", "readStevimg", "writeStevimg",
                 "patchMatch", "cachelessPatchMatch", "OnDemandStereoCostVolume", "OnDemandImageFlowVolume", "truncatedCostVolume"):
        assert name in text, name


_GRPIX17R3 = [[14, 14, 10, 10, 10, 16, 16], [14, 14, 6, 4, 7, 16, 16], [11, 6, 6, 4, 7, 7, 13], [11, 2, 2, 1, 3, 3, 13],
              [11, 8, 8, 5, 9, 9, 13], [15, 15, 8, 5, 9, 17, 17], [15, 15, 12, 12, 12, 17, 17]]


def test_shipped_compressor_masks_equal_the_reference(tmp_path):
    """GrPix17R3Filter / GrPix17R4Filter of the facade against the reference's own functions (oracle/_ref), unfold.h:476-691."""
    import ctypes
    import oracle as O
    if O.ref is None or not O.ref.has("grpix17_filter"):
        pytest.skip("oracle/_ref was not built")
    src = os.path.join(str(tmp_path), "masks.cpp")
    open(src, "w").write('#include <cstdio>\n#include <StereoVision/correlation/unfold.h>\nint main() { using namespace StereoVision::Correlation;\n'
                         'for (int r = 3; r <= 4; r++) { auto m = r == 3 ? CompressorGenerators::GrPix17R3Filter() : CompressorGenerators::GrPix17R4Filter();\n'
                         'for (int i = 0; i < 2 * r + 1; i++) for (int j = 0; j < 2 * r + 1; j++) std::printf("%d ", m.valueUnchecked(i, j)); std::printf("\\n"); } }\n')
    exe = os.path.join(str(tmp_path), "masks")
    r = subprocess.run(["g++", "-std=c++17", "-I" + os.path.join(ROOT, "include"), src, "-o", exe, "-L" + LIBDIR, "-lstevi_b200",
                        "-Wl,-rpath," + LIBDIR], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = subprocess.run([exe], capture_output=True, text=True).stdout.strip().splitlines()
    for radius, ln in zip((3, 4), lines):
        n = 2 * radius + 1
        want = np.zeros((n, n), np.int32)
        O.ref.lib.ref_grpix17_filter(radius, want.ctypes.data_as(ctypes.c_void_p))
        assert np.array_equal(np.array(ln.split(), np.int32).reshape(n, n), want)
    assert np.array_equal(np.array(_GRPIX17R3), np.array(lines[0].split(), int).reshape(7, 7))


def _mix(z):
    with np.errstate(over="ignore"):
        z = z + np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def _checksum(a):
    v = np.asarray(a, np.float64).ravel()  # C order = the demo's IndexBlock::moveToNextIndex order? see below
    k = np.arange(v.size)
    return float(np.sum(np.where(np.isfinite(v), v * (1 + k % 7), 0.0)))


@pytest.mark.gpu
def test_facade_results_equal_the_python_mirror(tmp_path, api, ctx):
    exe = build_demo(tmp_path)
    out = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    got = dict(ln.split() for ln in out.stdout.strip().splitlines())
    h, w, D = 40, 64, 16
    ii, jj = np.meshgrid(np.arange(h, dtype=np.uint64), np.arange(w, dtype=np.uint64), indexing="ij")
    r8 = (_mix(ii * np.uint64(w) + jj) >> np.uint64(56)).astype(np.uint8)
    l8 = np.zeros_like(r8)
    l8[:, 5:] = r8[:, :-5]
    l8[:, :5] = (_mix(np.uint64(1000003) + ii[:, :5] * np.uint64(w) + jj[:, :5]) >> np.uint64(56)).astype(np.uint8)
    lf, rf = l8.astype(np.float32) / np.float32(255), r8.astype(np.float32) / np.float32(255)
    cv, sg, disp = api.censusSgmPipeline(l8, r8, 4, 3, D, 10, 120, 100, 8, return_volumes=True, layout="c", ctx=ctx)
    # the demo walks arrays with IndexBlock::moveToNextIndex; the order only matters through the position weights, so
    # compare order-free sums as well
    assert float(got["mismatch_empty"]) == 1 and float(got["even_depth_empty"]) == 1 and float(got["even_side_empty"]) == 1
    assert float(got["search2d_match"]) > 0.95
    assert int(got["pyramid_levels"]) == 3 and got["pyramid_last_shape"] == "10x16x9" and int(got["pyramid_empty"]) == 1
    assert int(got["compressed_features"]) == 17 and float(got["compressed_match"]) > 0.9 and int(got["aggregate_offset_depth"]) == 8
    # the feature-level results against the Python mirror (order-free: both volumes are in the reference's layout)
    mask = np.array(_GRPIX17R3, np.int32)
    ccv = api.unfoldBasedCostVolumeCompressed(api.ZNCC, lf, rf, mask, D, ctx=ctx)
    l3, r3 = np.stack([lf, 1 - lf], axis=2).astype(np.float32), np.stack([rf, 1 - rf], axis=2).astype(np.float32)
    mcv = api.unfoldBasedCostVolumeMultiChannel(api.SAD, l3, r3, 1, 1, D, ctx=ctx)
    f2c = api.featureVolume2CostVolume(api.ZNCC, api.unfold(2, 2, lf, ctx=ctx), api.unfold(2, 2, rf, ctx=ctx), D, ctx=ctx)
    mean = api.meanFilter2D(2, 1, lf, ctx=ctx)
    assert int(got["mean_filter_empty"]) == 1
    for name, arr in (("compressed_cv", ccv), ("multichannel_cv", mcv), ("fv2cv", f2c), ("mean_filter", mean),
                      ("sigma_filter", api.sigmaFilter(2, 1, mean, lf, ctx=ctx))):
        a = np.asarray(arr)
        cands = {_checksum(a), _checksum(np.transpose(a))}
        assert any(abs(float(got[name]) - c) <= 2e-3 * max(1.0, abs(c)) for c in cands), (name, got[name], cands)
    zn = api.unfoldBasedCostVolume(api.ZNCC, lf, rf, 2, 2, D, layout="c", ctx=ctx)
    zi = api.extractSelectedIndex(zn, api.SCORE, ctx=ctx)
    # feature-domain refinement: the facade's results against the Python mirror's
    fvl, fvr = api.unfold(2, 2, lf, ctx=ctx), api.unfold(2, 2, rf, ctx=ctx)
    zero2 = np.zeros((h, w, 2), np.int32)
    for name, arr in (("barycentric", api.refinedBarycentricDisp(api.ZNCC, lf, rf, 2, 2, D, ctx=ctx)),
                      ("barycentric_symmetric", api.refineBarycentricSymmetricDisp(api.SSD, fvl, fvr, zi, D, ctx=ctx)),
                      ("barycentric_2d", api.refineBarycentric2dDisp(api.ZSSD, fvl, fvl, zero2, (-1, 1, -1, 1), api.ROOK, ctx=ctx))):
        a = np.asarray(arr)
        cands = {_checksum(a), _checksum(np.transpose(a))}
        assert any(abs(float(got[name]) - c) <= 1e-4 * max(1.0, abs(c)) for c in cands), (name, got[name], cands)
    assert int(got["barycentric_empty"]) == 1
    # guided aggregation: the caller's functors on the host, the aggregation on the GPU
    win = api.computeAggregationWindows(rf, 2, 1, lambda di, dj: np.exp(-np.float32(di * di + dj * dj) / np.float32(4)),
                                        lambda a, b: np.exp(-(a[..., 0] - b[..., 0]) ** 2 * np.float32(8)))
    f2c_c = np.ascontiguousarray(f2c)
    for name, arr, tol in (("aggregation_windows", win, 1e-4), ("aggregated", api.variableCostVolumeAggregation(f2c_c, win, ctx=ctx), 2e-3)):
        a = np.asarray(arr)
        cands = {_checksum(a), _checksum(np.transpose(a))}
        assert any(abs(float(got[name]) - c) <= tol * max(1.0, abs(c)) for c in cands), (name, got[name], cands)
    assert int(got["aggregated_empty"]) == 1
    # PatchMatch through the facade (seed 0) = the Python mirror's with the same seed; the on-demand cells around its result
    pm = api.patchMatch(api.ZNCC, fvr, fvl, (0, D - 1), 5, 4, seed=0, ctx=ctx)
    assert abs(float(got["patchmatch_match"]) - float((pm[4:h - 4, 4:w - 24, 0] == 5).mean())) < 1e-3 and float(got["patchmatch_match"]) > 0.95
    odt = np.asarray(api.onDemandTruncatedCostVolume(api.ZNCC, fvr, fvl, (0, D - 1), pm, 1, ctx=ctx))
    cands = {_checksum(odt), _checksum(np.transpose(odt))}
    assert any(abs(float(got["on_demand_tcv"]) - c) <= 1e-4 * max(1.0, abs(c)) for c in cands), (got["on_demand_tcv"], cands)
    assert int(got["patchmatch_empty"]) == 1
    assert abs(float(got["zncc_match"]) - float((zi[4:h - 4, 4:w - 24] == 5).mean())) < 1e-3
    assert float(got["zncc_match"]) > 0.95
    order = os.environ.get("STEVI_FACADE_ORDER", "")
    hier_disp = api.hiearchicalTruncatedCostVolume(api.ZNCC, 2, lf, rf, 2, 2, D, 2, ctx=ctx)[1]
    zfl, zfr = (api.getFeatureVolumeForMatchFuncFromFeatures(api.ZNCC, f, ctx=ctx) for f in (fvl, fvr))
    sums = {"census_cv": cv, "sgm": sg, "disp": disp, "zncc_idx": zi, "hier_disp": hier_disp,
            "guided_features_disp": api.computeGuidedCVFromFeatures(api.ZNCC, zfl, zfr, hier_disp, 2, ctx=ctx)[1]}
    for name, arr in sums.items():
        a = np.asarray(arr)
        cands = {_checksum(a), _checksum(np.transpose(a))}   # last-index-fastest or first-index-fastest walk
        assert any(abs(float(got[name]) - c) <= 1e-6 * max(1.0, abs(c)) for c in cands), (name, got[name], cands, order)
