// Reads like the bodies of the reference's benchmarks (test/benchmarks/benchmarkCrossCorrelationAlgorithms.cpp:93-95,
// :216-221, :289-293), written against the B200 facade headers. Prints one line per stage: checksum of the result.
// Built by tests/test_facade.py with g++ (no nvcc): the facade is plain C++17 over the C ABI.
#include <cstdio>
#include <cmath>
#include <StereoVision/correlation/cross_correlations.h>
#include <StereoVision/correlation/sgm.h>
#include <StereoVision/correlation/cost_based_refinement.h>
#include <StereoVision/correlation/hierarchical.h>
#include <StereoVision/correlation/image_based_refinement.h>
#include <StereoVision/correlation/guided_cost_filtering.h>
#include <StereoVision/correlation/patchmatch.h>

using namespace StereoVision;
using namespace StereoVision::Correlation;

static uint64_t mix(uint64_t z) { z += 0x9E3779B97F4A7C15ull; z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; return z ^ (z >> 31); }

template<class A> double checksum(A const& a) {
	double s = 0; auto sh = a.shape(); typename A::IndexBlock ix; ix.fill(0);
	if (a.empty()) return -1;
	size_t n = 1; for (auto v : sh) n *= v;
	for (size_t k = 0; k < n; k++) { double v = double(a.valueUnchecked(ix)); if (std::isfinite(v)) s += v * double(1 + k % 7); ix.moveToNextIndex(sh); }
	return s;
}

int main() {
	const int h = 40, w = 64, D = 16;
	Multidim::Array<uint8_t, 2> l8(h, w), r8(h, w);
	Multidim::Array<float, 2> lf(h, w), rf(h, w);
	for (int i = 0; i < h; i++) for (int j = 0; j < w; j++) {
		uint8_t v = uint8_t(mix(uint64_t(i) * w + j) >> 56);
		r8.at<Multidim::AccessCheck::Nocheck>(i, j) = v; rf.at<Multidim::AccessCheck::Nocheck>(i, j) = v / 255.f;
		int jt = j + 5;
		if (jt < w) { l8.at<Multidim::AccessCheck::Nocheck>(i, jt) = v; lf.at<Multidim::AccessCheck::Nocheck>(i, jt) = v / 255.f; }
	}
	for (int i = 0; i < h; i++) for (int j = 0; j < 5; j++) { uint8_t v = uint8_t(mix(1000003ull + uint64_t(i) * w + j) >> 56); l8.at<Multidim::AccessCheck::Nocheck>(i, j) = v; lf.at<Multidim::AccessCheck::Nocheck>(i, j) = v / 255.f; }

	auto cv = unfoldBasedCostVolume<matchingFunctions::CENSUS, uint8_t, uint8_t, 2, dispDirection::RightToLeft, hamming_cv_t>(l8, r8, 4, 3, D);
	auto sg = sgmCostVolume<8, dispExtractionStartegy::Cost>(cv, 10.f, 120.f, Margins(), 100.f);
	auto idx = extractSelectedIndex<dispExtractionStartegy::Cost>(sg);
	auto disp = selectedIndexToDisp<disp_t, dispDirection::RightToLeft>(idx, 0);
	std::printf("census_cv %.0f\nsgm %.0f\ndisp %.0f\n", checksum(cv), checksum(sg), checksum(disp));

	auto zn = unfoldBasedCostVolume<matchingFunctions::ZNCC>(lf, rf, 2, 2, D);
	auto zi = extractSelectedIndex<MatchingFunctionTraits<matchingFunctions::ZNCC>::extractionStrategy>(zn);
	auto tcv = truncatedCostVolume(zn, zi, 2, 2, 1);
	auto ref = refineDispCostInterpolation<InterpolationKernel::Parabola>(tcv, zi);
	int good = 0, n = 0;
	for (int i = 4; i < h - 4; i++) for (int j = 4; j < w - 24; j++) { n++; good += zi.valueUnchecked(i, j) == 5; }
	std::printf("zncc_idx %.0f\nzncc_match %.3f\nrefined %.3f\n", checksum(zi), double(good) / n, checksum(ref));

	// 2-D search chain, as test/testCorrelation2d.cpp:76-127 and the refinement callers use it
	Multidim::Array<float, 2> l2(h, w), r2(h, w);
	for (int i = 0; i < h; i++) for (int j = 0; j < w; j++) l2.at<Multidim::AccessCheck::Nocheck>(i, j) = 0.f;
	for (int i = 0; i < h; i++) for (int j = 0; j < w; j++) {
		r2.at<Multidim::AccessCheck::Nocheck>(i, j) = float(mix(77777ull + uint64_t(i) * w + j) >> 40) / float(1 << 24);
		int it = i + 1, jt = j - 2;   // the source pixel (i, j) of r2 is found at (i + 1, j - 2) in l2
		if (it >= 0 && it < h && jt >= 0 && jt < w) l2.at<Multidim::AccessCheck::Nocheck>(it, jt) = r2.valueUnchecked(i, j);
	}
	searchOffset<2> window(-2, 2, -3, 3);
	auto cv2 = unfoldBased2dDisparityCostVolume<matchingFunctions::ZNCC, float, float>(l2, r2, 2, 2, window);
	auto i2 = extractSelected2dIndex<dispExtractionStartegy::Score>(cv2);
	auto d2 = selected2dIndexToDisp(i2, window);
	auto t2 = truncatedBidirectionaCostVolume(cv2, i2, 1, 1);
	auto f2 = refineDisp2dCostInterpolation<InterpolationKernel::Parabola, IsotropyHypothesis::Anisotropic>(t2, d2);
	int good2 = 0, n2 = 0;
	for (int i = 5; i < h - 5; i++) for (int j = 6; j < w - 6; j++) { n2++; good2 += d2.valueUnchecked(i, j, 0) == 1 && d2.valueUnchecked(i, j, 1) == -2; }
	std::printf("search2d_match %.3f\nsearch2d_idx %.0f\nsearch2d_refined %.3f\n", double(good2) / n2, checksum(i2), checksum(f2));
	auto even2 = refineDisp2dCostInterpolation<InterpolationKernel::Parabola>(Multidim::Array<float, 4>(h, w, 3, 4), d2);
	std::printf("even_side_empty %d\n", int(even2.empty()));

	// feature-level forms: compressed unfold (the shipped 17-super-pixel mask), multi-channel images, aggregateCost on features
	UnFoldCompressor comp(CompressorGenerators::GrPix17R3Filter());
	auto ccv = unfoldBasedCostVolume<matchingFunctions::ZNCC, float, float>(lf, rf, comp, D);
	auto cidx = extractSelectedIndex<dispExtractionStartegy::Score>(ccv);
	int goodc = 0, nc = 0;
	for (int i = 5; i < h - 5; i++) for (int j = 5; j < w - 24; j++) { nc++; goodc += cidx.valueUnchecked(i, j) == 5; }
	std::printf("compressed_features %d\ncompressed_match %.3f\ncompressed_cv %.4f\n", comp.nFeatures(), double(goodc) / nc, checksum(ccv));
	Multidim::Array<float, 3> l3(h, w, 2), r3(h, w, 2);
	for (int i = 0; i < h; i++) for (int j = 0; j < w; j++) {
		l3.atUnchecked(i, j, 0) = lf.valueUnchecked(i, j); l3.atUnchecked(i, j, 1) = 1.f - lf.valueUnchecked(i, j);
		r3.atUnchecked(i, j, 0) = rf.valueUnchecked(i, j); r3.atUnchecked(i, j, 1) = 1.f - rf.valueUnchecked(i, j);
	}
	auto mcv = unfoldBasedCostVolume<matchingFunctions::SAD, float, float, 3>(l3, r3, 1, 1, D);
	auto fvl = unfold<float, float>(2, 2, lf), fvr = unfold<float, float>(2, 2, rf);
	auto f2c = featureVolume2CostVolume<matchingFunctions::ZNCC, float, float>(fvl, fvr, D);
	auto agg = aggregateCost<matchingFunctions::ZNCC, float, float>(getFeatureVolumeForMatchFunc<matchingFunctions::ZNCC>(fvl),
																		 getFeatureVolumeForMatchFunc<matchingFunctions::ZNCC>(fvr), searchOffset<1>(2, 9));
	std::printf("multichannel_cv %.3f\nfv2cv %.4f\naggregate_offset_depth %d\n", checksum(mcv), checksum(f2c), agg.shape()[2]);

	auto hier = hiearchicalTruncatedCostVolume<matchingFunctions::ZNCC, 2>(lf, rf, 2, 2, D, 2);
	std::printf("hier_disp %.0f\n", checksum(hier.disp_estimate));
	// computeGuidedCV in the reference's own form: processed feature volumes + the coarser level's estimate
	auto gcv = computeGuidedCV<matchingFunctions::ZNCC, float, float>(getFeatureVolumeForMatchFunc<matchingFunctions::ZNCC>(fvl),
																	  getFeatureVolumeForMatchFunc<matchingFunctions::ZNCC>(fvr), hier.disp_estimate, 2);
	std::printf("guided_features_disp %.0f\n", checksum(gcv.disp_estimate));
	auto pyr = buildFeaturePyramid<float>(lf, 1, 1, 3);
	auto psh = pyr.back().shape();
	std::printf("pyramid_levels %d\npyramid_last_shape %dx%dx%d\npyramid_empty %d\n", int(pyr.size()), psh[0], psh[1], psh[2],
				int(buildFeaturePyramid<float>(lf, 1, 1, 0).empty()));
	auto mean = meanFilter2D(2, 1, lf);
	auto sigma = sigmaFilter(2, 1, mean, lf);
	std::printf("mean_filter %.4f\nsigma_filter %.4f\nmean_filter_empty %d\n", checksum(mean), checksum(sigma),
				int(meanFilter2D(uint8_t(w), 1, lf).empty()));
	// feature-domain refinement (image_based_refinement.h): image-level and feature-level forms
	auto bary = refinedBarycentricDisp<matchingFunctions::ZNCC, float, float>(lf, rf, 2, 2, D);
	auto barysym = refineBarycentricSymmetricDisp<matchingFunctions::SSD>(fvl, fvr, zi, D);
	Multidim::Array<disp_t, 3> zero2(h, w, 2);
	for (int i = 0; i < h; i++) for (int j = 0; j < w; j++) { zero2.at<Multidim::AccessCheck::Nocheck>(i, j, 0) = 0; zero2.at<Multidim::AccessCheck::Nocheck>(i, j, 1) = 0; }
	auto bary2 = refineBarycentric2dDisp<matchingFunctions::ZSSD, Contiguity::Rook>(fvl, fvl, zero2, searchOffset<2>(-1, 1, -1, 1));
	std::printf("barycentric %.3f\nbarycentric_symmetric %.3f\nbarycentric_2d %.3f\nbarycentric_empty %d\n", checksum(bary), checksum(barysym), checksum(bary2),
				int(refinedBarycentricDisp<matchingFunctions::ZNCC, float, float>(lf, Multidim::Array<float, 2>(h + 1, w), 2, 2, D).empty()));
	// guided aggregation (guided_cost_filtering.h): bilateral weights from the caller's functors, aggregation on the GPU
	auto windows = computeAggregationWindows<float>(rf, 2, 1, [](int di, int dj) { return std::exp(-float(di * di + dj * dj) / 4.f); },
		[](Multidim::Array<float, 1> const& a, Multidim::Array<float, 1> const& b) { float d = a.valueUnchecked(0) - b.valueUnchecked(0); return std::exp(-d * d * 8.f); });
	auto aggregated = variableCostVolumeAggregation(f2c, windows);
	std::printf("aggregation_windows %.4f\naggregated %.3f\naggregated_empty %d\n", checksum(windows), checksum(aggregated),
				int(variableCostVolumeAggregation(f2c, Multidim::Array<float, 4>(h, w, 2, 3)).empty()));
	// PatchMatch over on-demand costs (patchmatch.h), what examples/stereo-match runs: source = right features, target = left
	auto pm = patchMatch<matchingFunctions::ZNCC, 1>(fvr, fvl, searchOffset<1>(0, D - 1), 5, 4);
	int pm_good = 0;
	for (int i = 4; i < h - 4; i++) for (int j = 4; j < w - 24; j++) pm_good += pm.valueUnchecked(i, j, 0) == 5;
	OnDemandStereoCostVolume<matchingFunctions::ZNCC> odcv(fvr, fvl, searchOffset<1>(0, D - 1));
	auto odt = odcv.truncatedCostVolume(pm, 1);
	std::printf("patchmatch_match %.3f\non_demand_tcv %.3f\npatchmatch_empty %d\n", double(pm_good) / ((h - 8) * (w - 28)), checksum(odt),
				int(patchMatch<matchingFunctions::ZNCC, 1>(fvr, Multidim::Array<float, 3>(h + 1, w, 25), searchOffset<1>(0, D - 1)).empty()));
	auto bad = unfoldBasedCostVolume<matchingFunctions::ZNCC>(lf, Multidim::Array<float, 2>(h + 1, w), 2, 2, D);
	std::printf("mismatch_empty %d\n", int(bad.empty()));
	auto even = refineDispCostInterpolation<InterpolationKernel::Parabola>(Multidim::Array<float, 3>(h, w, 4), zi);
	std::printf("even_depth_empty %d\n", int(even.empty()));
	return 0;
}
