// B200 facade of LibStevi's correlation/cross_correlations.h: unfoldBasedCostVolume :740-765 (the fused unfold ->
// getFeatureVolumeForMatchFunc -> aggregateCost chain; 2-D and multi-channel images), its UnFoldCompressor overload :767-791,
// getFeatureVolumeForMatchFunc :645-722, aggregateCost :194-308 and featureVolume2CostVolume :724-738 on float feature volumes.
// Result layout like the reference's aggregateCost (:220): strides {W*D, 1, W}. 2-D search: unfoldBased2dDisparityCostVolume
// :794-822, result strides {W*dW*dH, 1, W*dW, W} (:340-341).
#ifndef STEVI_B200_CORRELATION_CROSS_CORRELATIONS_H
#define STEVI_B200_CORRELATION_CROSS_CORRELATIONS_H

#include "./matching_costs.h"
#include "./unfold.h"
#include "./census.h"

namespace StereoVision {
namespace Correlation {

template<matchingFunctions matchFunc, class T_L, class T_R, int nImDim = 2, dispDirection dDir = dispDirection::RightToLeft, typename TCV = float>
Multidim::Array<TCV, 3> unfoldBasedCostVolume(Multidim::Array<T_L, nImDim> const& img_l, Multidim::Array<T_R, nImDim> const& img_r,
											  uint8_t h_radius, uint8_t v_radius, disp_t disp_width) {
	static_assert(MatchingFunctionTraits<matchFunc>::onB200, "matching function not covered by the B200 library");
	static_assert(nImDim == 2 || nImDim == 3, "images are (H,W) or (H,W,C)");
	static_assert(std::is_same_v<T_L, T_R>, "both images must have the same element type");
	static_assert(MatchingFunctionTraits<matchFunc>::isCensusBased
				  ? (std::is_same_v<TCV, hamming_cv_t> || std::is_same_v<TCV, int32_t> || std::is_same_v<TCV, float>)
				  : std::is_same_v<TCV, float>, "cost volume type not covered");
	auto l_shape = img_l.shape();
	auto r_shape = img_r.shape();
	if (l_shape[0] != r_shape[0]) return Multidim::Array<TCV, 3>(0, 0, 0);   // :751-753
	if (nImDim == 3 && l_shape[nImDim - 1] != r_shape[nImDim - 1]) return Multidim::Array<TCV, 3>(0, 0, 0);   // :755-759
	const int h = l_shape[0], w = dDir == dispDirection::RightToLeft ? r_shape[1] : l_shape[1];
	Multidim::Array<TCV, 3> cv({h, w, disp_width}, {w * disp_width, 1, w});
	stevi_array l = B200::describe(img_l), r = B200::describe(img_r), o = B200::describe(cv);
	const int32_t rc = nImDim == 2
		? stevi_cost_volume(B200::threadContext(), int(matchFunc), &l, &r, h_radius, v_radius, disp_width, int(dDir), STEVI_COMPAT, &o)
		: stevi_cost_volume_multichannel(B200::threadContext(), int(matchFunc), &l, &r, h_radius, v_radius, disp_width, int(dDir), STEVI_COMPAT, &o);
	if (!B200::ok(rc)) return Multidim::Array<TCV, 3>(0, 0, 0);
	return cv;
}

// the UnFoldCompressor overload, :767-791
template<matchingFunctions matchFunc, class T_L, class T_R, int nImDim = 2, dispDirection dDir = dispDirection::RightToLeft, typename TCV = float>
Multidim::Array<TCV, 3> unfoldBasedCostVolume(Multidim::Array<T_L, nImDim> const& img_l, Multidim::Array<T_R, nImDim> const& img_r,
											  UnFoldCompressor const& compressor, disp_t disp_width) {
	static_assert(MatchingFunctionTraits<matchFunc>::onB200, "matching function not covered by the B200 library");
	static_assert(nImDim == 2 || nImDim == 3, "images are (H,W) or (H,W,C)");
	static_assert(std::is_same_v<T_L, T_R>, "both images must have the same element type");
	auto l_shape = img_l.shape();
	auto r_shape = img_r.shape();
	if (l_shape[0] != r_shape[0]) return Multidim::Array<TCV, 3>(0, 0, 0);
	if (nImDim == 3 && l_shape[nImDim - 1] != r_shape[nImDim - 1]) return Multidim::Array<TCV, 3>(0, 0, 0);
	const int h = l_shape[0], w = dDir == dispDirection::RightToLeft ? r_shape[1] : l_shape[1];
	Multidim::Array<TCV, 3> cv({h, w, disp_width}, {w * disp_width, 1, w});
	Multidim::Array<int, 2> mask = compressor.mask();
	stevi_array k = B200::describe(mask), l = B200::describe(img_l), r = B200::describe(img_r), o = B200::describe(cv);
	if (!B200::ok(stevi_cost_volume_compressed(B200::threadContext(), int(matchFunc), &k, &l, &r, disp_width, int(dDir), STEVI_COMPAT, &o)))
		return Multidim::Array<TCV, 3>(0, 0, 0);
	return cv;
}

// getFeatureVolumeForMatchFunc<matchFunc>(feature_vol) on a ready-made float volume, :645-722
template<matchingFunctions matchFunc>
Multidim::Array<float, 3> getFeatureVolumeForMatchFunc(Multidim::Array<float, 3> const& feature_vol) {
	static_assert(MatchingFunctionTraits<matchFunc>::onB200 && !MatchingFunctionTraits<matchFunc>::isCensusBased, "use censusFeatures");
	auto sh = feature_vol.shape();
	Multidim::Array<float, 3> out({sh[0], sh[1], sh[2]}, {sh[1] * sh[2], sh[2], 1});   // :427
	stevi_array i = B200::describe(feature_vol), o = B200::describe(out);
	if (!B200::ok(stevi_features_for_matchfunc(B200::threadContext(), int(matchFunc), &i, &o))) return Multidim::Array<float, 3>();
	return out;
}

// aggregateCost on processed float feature volumes: the searchOffset<1> overload :251-308 and the disp_width one :194-249
template<matchingFunctions matchFunc, class T_L, class T_R, dispDirection dDir = dispDirection::RightToLeft, typename TCV = float>
Multidim::Array<TCV, 3> aggregateCost(Multidim::Array<T_L, 3> const& feature_vol_l, Multidim::Array<T_R, 3> const& feature_vol_r,
									  searchOffset<1> searchRange) {
	static_assert(MatchingFunctionTraits<matchFunc>::onB200 && !MatchingFunctionTraits<matchFunc>::isCensusBased,
				  "census features: hammingCostVolume (stevi_hamming_cost_volume)");
	static_assert(std::is_same_v<T_L, float> && std::is_same_v<T_R, float> && std::is_same_v<TCV, float>, "float features and costs");
	auto l_shape = feature_vol_l.shape();
	auto r_shape = feature_vol_r.shape();
	const disp_t disp_width = searchRange.dimRange(0);
	if (l_shape[0] != r_shape[0] || disp_width <= 0) return Multidim::Array<TCV, 3>(0, 0, 0);
	const int h = l_shape[0], w = dDir == dispDirection::RightToLeft ? r_shape[1] : l_shape[1];
	Multidim::Array<TCV, 3> cv({h, w, disp_width}, {w * disp_width, 1, w});
	stevi_array l = B200::describe(feature_vol_l), r = B200::describe(feature_vol_r), o = B200::describe(cv);
	const int lower = searchRange.idx2disp(0, 0);
	if (!B200::ok(stevi_aggregate_cost(B200::threadContext(), int(matchFunc), &l, &r, lower, lower + disp_width - 1, int(dDir), &o)))
		return Multidim::Array<TCV, 3>(0, 0, 0);
	return cv;
}

template<matchingFunctions matchFunc, class T_L, class T_R, dispDirection dDir = dispDirection::RightToLeft, typename TCV = float>
Multidim::Array<TCV, 3> aggregateCost(Multidim::Array<T_L, 3> const& feature_vol_l, Multidim::Array<T_R, 3> const& feature_vol_r,
									  disp_t disp_width) {
	return aggregateCost<matchFunc, T_L, T_R, dDir, TCV>(feature_vol_l, feature_vol_r, searchOffset<1>(0, disp_width - 1));
}

// featureVolume2CostVolume, :724-738 (disp_t search range)
template<matchingFunctions matchFunc, class T_L, class T_R, typename SearchRangeType = disp_t, dispDirection dDir = dispDirection::RightToLeft,
		 typename TCV = float>
Multidim::Array<TCV, 3> featureVolume2CostVolume(Multidim::Array<T_L, 3> const& feature_vol_l, Multidim::Array<T_R, 3> const& feature_vol_r,
												 SearchRangeType disp_width) {
	static_assert(std::is_same_v<SearchRangeType, disp_t>, "the 1-D search takes a disparity width");
	static_assert(MatchingFunctionTraits<matchFunc>::onB200, "matching function not covered by the B200 library");
	static_assert(std::is_same_v<T_L, float> && std::is_same_v<T_R, float>, "float feature volumes");
	auto l_shape = feature_vol_l.shape();
	auto r_shape = feature_vol_r.shape();
	if (l_shape[0] != r_shape[0]) return Multidim::Array<TCV, 3>(0, 0, 0);
	const int h = l_shape[0], w = dDir == dispDirection::RightToLeft ? r_shape[1] : l_shape[1];
	Multidim::Array<TCV, 3> cv({h, w, disp_width}, {w * disp_width, 1, w});
	stevi_array l = B200::describe(feature_vol_l), r = B200::describe(feature_vol_r), o = B200::describe(cv);
	if (!B200::ok(stevi_feature_volume_to_cost_volume(B200::threadContext(), int(matchFunc), &l, &r, disp_width, int(dDir), STEVI_COMPAT, &o)))
		return Multidim::Array<TCV, 3>(0, 0, 0);
	return cv;
}

// getFeatureVolumeForMatchFunc<matchFunc>(unfold<T, T>(h_radius, v_radius, img)) for float images, strides {W*F, F, 1} (:427)
template<matchingFunctions matchFunc>
Multidim::Array<float, 3> featureVolumeForMatchFunc(Multidim::Array<float, 2> const& img, uint8_t h_radius, uint8_t v_radius) {
	static_assert(MatchingFunctionTraits<matchFunc>::onB200 && !MatchingFunctionTraits<matchFunc>::isCensusBased, "use censusTransform2D");
	auto sh = img.shape();
	const int F = (2 * h_radius + 1) * (2 * v_radius + 1);
	Multidim::Array<float, 3> out({sh[0], sh[1], F}, {sh[1] * F, F, 1});
	stevi_array i = B200::describe(img), o = B200::describe(out);
	if (!B200::ok(stevi_feature_volume(B200::threadContext(), int(matchFunc), &i, h_radius, v_radius, &o))) return Multidim::Array<float, 3>();
	return out;
}

template<matchingFunctions matchFunc, class T_L, class T_R, int nImDim = 2, dispDirection dDir = dispDirection::RightToLeft, typename TCV = float>
Multidim::Array<TCV, 4> unfoldBased2dDisparityCostVolume(Multidim::Array<T_L, nImDim> const& img_l, Multidim::Array<T_R, nImDim> const& img_r,
														 uint8_t h_radius, uint8_t v_radius, searchOffset<2> const& searchWindows) {
	static_assert(MatchingFunctionTraits<matchFunc>::onB200 && !MatchingFunctionTraits<matchFunc>::isCensusBased,
				  "2-D search runs the float matching functions on the B200 library");
	static_assert(nImDim == 2, "multi-channel images: unfold them and use the feature-level entry points");
	static_assert(std::is_same_v<T_L, float> && std::is_same_v<T_R, float> && std::is_same_v<TCV, float>, "float images and costs");
	auto l_shape = img_l.shape();
	auto r_shape = img_r.shape();
	if (l_shape[0] != r_shape[0] || l_shape[1] != r_shape[1]) return Multidim::Array<TCV, 4>(0, 0, 0, 0);   // :804-810
	const int h = l_shape[0], w = l_shape[1], dH = searchWindows.dimRange(0), dW = searchWindows.dimRange(1);
	if (dH <= 0 || dW <= 0) return Multidim::Array<TCV, 4>(0, 0, 0, 0);   // :336-338
	Multidim::Array<TCV, 4> cv({h, w, dH, dW}, {w * dW * dH, 1, w * dW, w});
	const auto so = B200Detail::search4(searchWindows);
	stevi_array l = B200::describe(img_l), r = B200::describe(img_r), o = B200::describe(cv);
	if (!B200::ok(stevi_cost_volume_2d(B200::threadContext(), int(matchFunc), &l, &r, h_radius, v_radius, so.data(), int(dDir), &o)))
		return Multidim::Array<TCV, 4>(0, 0, 0, 0);
	return cv;
}

} // namespace Correlation
} // namespace StereoVision

#endif
