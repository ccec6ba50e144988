// B200 facade of LibStevi's correlation/hierarchical.h: OffsetedCostVolume :33-37, FeatureVolumePyramid / buildFeaturePyramid
// :39-71, hiearchicalTruncatedCostVolume (sic) :235-314 (both overloads). The guided volume of each level is computed from the images directly (unfold and feature
// normalisation are fused into the kernel), so the feature-volume form of computeGuidedCV (:73-233) is offered on images.
#ifndef STEVI_B200_CORRELATION_HIERARCHICAL_H
#define STEVI_B200_CORRELATION_HIERARCHICAL_H

#include <vector>
#include "./cross_correlations.h"
#include "../interpolation/downsampling.h"

namespace StereoVision {
namespace Correlation {

template<typename TCV>
struct OffsetedCostVolume {
	Multidim::Array<TCV, 3> truncated_cost_volume;
	Multidim::Array<disp_t, 2> disp_estimate;
};

// hierarchical.h:39-71: the unfolded features of an image at nLevels resolutions, level k = unfold of the image after k
// 2x2 average poolings (level 0 = full resolution); an empty pyramid for nLevels <= 0
template<typename F_T>
using FeatureVolumePyramid = std::vector<Multidim::Array<F_T, 3>>;

template<typename F_T, int nImDim = 2>
FeatureVolumePyramid<F_T> buildFeaturePyramid(Multidim::Array<F_T, nImDim> const& img, uint8_t h_radius, uint8_t v_radius, int nLevels,
											  PaddingMargins const& padding = PaddingMargins(), UnfoldPatchOrientation orientation = Rotate0) {
	static_assert(nImDim == 2 && std::is_same_v<F_T, float>, "the feature pyramid is built from (H,W) float images on the B200 library");
	FeatureVolumePyramid<F_T> ret;
	if (nLevels <= 0) return ret;
	ret.reserve(nLevels);
	ret.emplace_back(unfold<F_T, F_T>(h_radius, v_radius, img, padding, orientation));
	Multidim::Array<F_T, 2> level = Interpolation::averagePoolingDownsample<F_T, F_T>(img, Interpolation::DownSampleWindows(2));
	for (int k = 1; k < nLevels; k++) {
		ret.emplace_back(unfold<F_T, F_T>(h_radius, v_radius, level, padding, orientation));
		if (k + 1 < nLevels) level = Interpolation::averagePoolingDownsample<F_T, F_T>(level, Interpolation::DownSampleWindows(2));
	}
	return ret;
}

template<matchingFunctions matchFunc, int depth, class T_L, class T_R, int nImDim = 2, dispDirection dDir = dispDirection::RightToLeft, class TCV = float>
OffsetedCostVolume<TCV> hiearchicalTruncatedCostVolume(Multidim::Array<T_L, nImDim> const& img_l, Multidim::Array<T_R, nImDim> const& img_r,
													   std::array<uint8_t, depth + 1> h_radiuses, std::array<uint8_t, depth + 1> v_radiuses,
													   disp_t disp_width, disp_t upscale_disp_radius = 2) {
	static_assert(depth > 0, "Minimum depth is 1");
	static_assert(nImDim == 2 && std::is_same_v<T_L, float> && std::is_same_v<T_R, float> && std::is_same_v<TCV, float>,
				  "the coarse-to-fine driver runs on float images on the B200 library");
	static_assert(MatchingFunctionTraits<matchFunc>::onB200 && !MatchingFunctionTraits<matchFunc>::isCensusBased, "matching function not covered");
	auto sh = img_l.shape();
	const int t = 2 * upscale_disp_radius + 1;
	OffsetedCostVolume<TCV> r{Multidim::Array<TCV, 3>({sh[0], sh[1], t}, {sh[1] * t, t, 1}), Multidim::Array<disp_t, 2>(sh[0], sh[1])};   // :107
	int32_t hr[depth + 1], vr[depth + 1];
	for (int k = 0; k <= depth; k++) { hr[k] = h_radiuses[k]; vr[k] = v_radiuses[k]; }
	stevi_array l = B200::describe(img_l), rr = B200::describe(img_r), tc = B200::describe(r.truncated_cost_volume), d = B200::describe(r.disp_estimate);
	if (!B200::ok(stevi_hierarchical(B200::threadContext(), int(matchFunc), depth, &l, &rr, hr, vr, 0, 0, disp_width, upscale_disp_radius, int(dDir), &tc, &d)))
		return OffsetedCostVolume<TCV>{Multidim::Array<TCV, 3>(), Multidim::Array<disp_t, 2>()};
	return r;
}

template<matchingFunctions matchFunc, int depth, class T_L, class T_R, int nImDim = 2, dispDirection dDir = dispDirection::RightToLeft, class TCV = float>
OffsetedCostVolume<TCV> hiearchicalTruncatedCostVolume(Multidim::Array<T_L, nImDim> const& img_l, Multidim::Array<T_R, nImDim> const& img_r,
													   uint8_t h_radius, uint8_t v_radius, disp_t disp_width, disp_t upscale_disp_radius = 2) {
	std::array<uint8_t, depth + 1> h_radiuses, v_radiuses;
	h_radiuses.fill(h_radius);
	v_radiuses.fill(v_radius);
	return hiearchicalTruncatedCostVolume<matchFunc, depth, T_L, T_R, nImDim, dDir, TCV>(img_l, img_r, h_radiuses, v_radiuses, disp_width, upscale_disp_radius);
}

// computeGuidedCV, :73-233: the caller's (processed) feature volumes and the coarser level's disparity estimate
template<matchingFunctions matchFunc, typename T_L, typename T_R, dispDirection dDir = dispDirection::RightToLeft, typename TCV = float>
OffsetedCostVolume<TCV> computeGuidedCV(Multidim::Array<T_L, 3> const& feature_vol_l, Multidim::Array<T_R, 3> const& feature_vol_r,
										Multidim::Array<disp_t, 2> disp_guide, disp_t upscale_disp_radius) {
	static_assert(std::is_same_v<T_L, float> && std::is_same_v<T_R, float> && std::is_same_v<TCV, float>, "float features and costs");
	auto sh = (dDir == dispDirection::RightToLeft ? feature_vol_r : feature_vol_l).shape();
	const int t = 2 * upscale_disp_radius + 1;
	OffsetedCostVolume<TCV> r{Multidim::Array<TCV, 3>({sh[0], sh[1], t}, {sh[1] * t, t, 1}), Multidim::Array<disp_t, 2>(sh[0], sh[1])};   // strides :107
	stevi_array l = B200::describe(feature_vol_l), rr = B200::describe(feature_vol_r), g = B200::describe(disp_guide),
				tc = B200::describe(r.truncated_cost_volume), d = B200::describe(r.disp_estimate);
	if (!B200::ok(stevi_guided_cost_volume_features(B200::threadContext(), int(matchFunc), &l, &rr, &g, upscale_disp_radius, int(dDir), &tc, &d)))
		return OffsetedCostVolume<TCV>{Multidim::Array<TCV, 3>(), Multidim::Array<disp_t, 2>()};
	return r;
}

// computeGuidedCV on images (unfold + features fused): guide = the coarser level's disp_estimate / selected index
template<matchingFunctions matchFunc, dispDirection dDir = dispDirection::RightToLeft>
OffsetedCostVolume<float> computeGuidedCV(Multidim::Array<float, 2> const& img_l, Multidim::Array<float, 2> const& img_r, uint8_t h_radius,
										  uint8_t v_radius, Multidim::Array<disp_t, 2> const& disp_guide, disp_t upscale_disp_radius = 2) {
	auto sh = (dDir == dispDirection::RightToLeft ? img_r : img_l).shape();
	const int t = 2 * upscale_disp_radius + 1;
	OffsetedCostVolume<float> r{Multidim::Array<float, 3>({sh[0], sh[1], t}, {sh[1] * t, t, 1}), Multidim::Array<disp_t, 2>(sh[0], sh[1])};
	stevi_array l = B200::describe(img_l), rr = B200::describe(img_r), g = B200::describe(disp_guide),
				tc = B200::describe(r.truncated_cost_volume), d = B200::describe(r.disp_estimate);
	if (!B200::ok(stevi_guided_cost_volume(B200::threadContext(), int(matchFunc), &l, &rr, h_radius, v_radius, &g, upscale_disp_radius, int(dDir), &tc, &d)))
		return OffsetedCostVolume<float>{Multidim::Array<float, 3>(), Multidim::Array<disp_t, 2>()};
	return r;
}

} // namespace Correlation
} // namespace StereoVision

#endif
