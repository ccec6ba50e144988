// B200 facade of LibStevi's correlation/image_based_refinement.h: feature-domain subpixel refinement. Same names, template
// parameters and defaults; the per-pixel fits run on the GPU (libstevi_b200/csrc/barycentric.cu). Reference lines:
// refineBarycentricSymmetricDisp :31-105, refineBarycentricDisp :184-355, refineBarycentric2dDisp :675-880,
// refineBarycentricSymmetric2dDisp :1068-1241, refineCostSymmetricDisp :1591-1690, refinedBarycentricSymmetricDispFeatureVol
// :1692-1765, refinedBarycentricDispFeatureVol :1854-1927, refinedCostSymmetricDispFeatureVol :2351-2424,
// refinedBarycentricSymmetricDisp :2426-2500, refinedBarycentricDisp :2502-2571, refinedCostSymmetricDisp :2573-2624.
// Not provided (the reference instantiates them for MEDAD / ZMEDAD only, which are not on the B200 path): the Subpart variants
// (:107-182, :515-673, :882-1066, :1243-1416); nor refineArbitraryInterpolationDisp / 2dDisp (:357-513, :1418-1589: a host
// function pointer as template argument, "not recommanded for production" by its own comment).
#ifndef STEVI_B200_CORRELATION_IMAGE_BASED_REFINEMENT_H
#define STEVI_B200_CORRELATION_IMAGE_BASED_REFINEMENT_H

#include "./cross_correlations.h"
#include "./unfold.h"

namespace StereoVision {

// utils/contiguity.h:28-51 (the part the 2-D refinements name)
class Contiguity {
public:
	enum bidimensionalContiguity { Queen = 0, Rook = 1, Bishop = 2 };
	constexpr static int nDirections(bidimensionalContiguity contiguity) { return contiguity == Queen ? 8 : 4; }
	constexpr static int nCornerDirections(bidimensionalContiguity contiguity) { return contiguity == Queen ? 3 : (contiguity == Rook ? 2 : 1); }
};

namespace Correlation {

namespace B200Detail {
template<matchingFunctions matchFunc>
constexpr bool hasBarycentric = matchFunc == matchingFunctions::NCC || matchFunc == matchingFunctions::ZNCC || matchFunc == matchingFunctions::SSD ||
								matchFunc == matchingFunctions::ZSSD || matchFunc == matchingFunctions::SAD || matchFunc == matchingFunctions::ZSAD;
template<matchingFunctions matchFunc>
constexpr bool isL1 = matchFunc == matchingFunctions::SAD || matchFunc == matchingFunctions::ZSAD;
template<dispDirection dDir, class A>
std::array<int, 2> sourceShape(A const& l, A const& r) { auto s = (dDir == dispDirection::RightToLeft ? r : l).shape(); return {s[0], s[1]}; }

// the volumes refined*FeatureVol match on (first pair) and refine on (second pair): :1700-1762
template<matchingFunctions matchFunc>
struct RefinementVolumes { Multidim::Array<float, 3> match_l, match_r, fine_l, fine_r; };
template<matchingFunctions matchFunc>
RefinementVolumes<matchFunc> refinementVolumes(Multidim::Array<float, 3> const& l, Multidim::Array<float, 3> const& r, bool preNormalize) {
	typedef MatchingFunctionTraits<matchFunc> T;
	RefinementVolumes<matchFunc> v;
	v.match_l = getFeatureVolumeForMatchFunc<matchFunc>(l);
	v.match_r = getFeatureVolumeForMatchFunc<matchFunc>(r);
	if (!T::Normalized || preNormalize) { v.fine_l = v.match_l; v.fine_r = v.match_r; }
	else if (T::ZeroMean) { v.fine_l = getFeatureVolumeForMatchFunc<matchingFunctions::ZCC>(l); v.fine_r = getFeatureVolumeForMatchFunc<matchingFunctions::ZCC>(r); }
	else { v.fine_l = l; v.fine_r = r; }
	return v;
}
} // namespace B200Detail

template<matchingFunctions matchFunc, int refineRadius = 1, dispDirection dDir = dispDirection::RightToLeft>
Multidim::Array<float, 2> refineBarycentricSymmetricDisp(Multidim::Array<float, 3> const& feature_vol_l, Multidim::Array<float, 3> const& feature_vol_r,
														 Multidim::Array<disp_t, 2> const& selectedIndex, disp_t disp_width) {
	static_assert(refineRadius > 0, "refineBarycentricSymmetricDisp cannot proceed with a refinement radius smaller than 1.");
	static_assert(B200Detail::hasBarycentric<matchFunc> && !B200Detail::isL1<matchFunc>, "NCC, ZNCC, SSD, ZSSD on the B200 path");
	Multidim::Array<float, 2> out(B200Detail::sourceShape<dDir>(feature_vol_l, feature_vol_r));
	stevi_array l = B200::describe(feature_vol_l), r = B200::describe(feature_vol_r), i = B200::describe(selectedIndex), o = B200::describe(out);
	if (!B200::ok(stevi_refine_barycentric_symmetric_disp(B200::threadContext(), int(matchFunc), refineRadius, &l, &r, &i, disp_width, int(dDir), &o)))
		return Multidim::Array<float, 2>();
	return out;
}

template<matchingFunctions matchFunc, dispDirection dDir = dispDirection::RightToLeft>
Multidim::Array<float, 2> refineBarycentricDisp(Multidim::Array<float, 3> const& feature_vol_l, Multidim::Array<float, 3> const& feature_vol_r,
												Multidim::Array<disp_t, 2> const& selectedIndex) {
	static_assert(B200Detail::hasBarycentric<matchFunc>, "the matching function has no barycentric approximation (matching_costs.h:419-589)");
	Multidim::Array<float, 2> out(B200Detail::sourceShape<dDir>(feature_vol_l, feature_vol_r));
	stevi_array l = B200::describe(feature_vol_l), r = B200::describe(feature_vol_r), i = B200::describe(selectedIndex), o = B200::describe(out);
	if (!B200::ok(stevi_refine_barycentric_disp(B200::threadContext(), int(matchFunc), &l, &r, &i, int(dDir), &o))) return Multidim::Array<float, 2>();
	return out;
}

namespace B200Detail {
template<matchingFunctions matchFunc, Contiguity::bidimensionalContiguity contiguity, dispDirection dDir>
Multidim::Array<float, 3> barycentric2d(Multidim::Array<float, 3> const& feature_vol_l, Multidim::Array<float, 3> const& feature_vol_r,
										Multidim::Array<disp_t, 3> const& selectedIndices, searchOffset<2> const& searchWindows, int symmetric) {
	static_assert(hasBarycentric<matchFunc> && !isL1<matchFunc>, "NCC, ZNCC, SSD, ZSSD on the B200 path");
	Multidim::Array<float, 3> out(selectedIndices.shape());
	auto so = search4(searchWindows);
	stevi_array l = B200::describe(feature_vol_l), r = B200::describe(feature_vol_r), i = B200::describe(selectedIndices), o = B200::describe(out);
	if (!B200::ok(stevi_refine_barycentric_2d_disp(B200::threadContext(), int(matchFunc), int(contiguity), symmetric, &l, &r, &i, so.data(), int(dDir), &o)))
		return Multidim::Array<float, 3>();
	return out;
}
} // namespace B200Detail

template<matchingFunctions matchFunc, Contiguity::bidimensionalContiguity contiguity = Contiguity::Queen, dispDirection dDir = dispDirection::RightToLeft>
Multidim::Array<float, 3> refineBarycentric2dDisp(Multidim::Array<float, 3> const& feature_vol_l, Multidim::Array<float, 3> const& feature_vol_r,
												  Multidim::Array<disp_t, 3> const& selectedIndices, searchOffset<2> const& searchWindows) {
	return B200Detail::barycentric2d<matchFunc, contiguity, dDir>(feature_vol_l, feature_vol_r, selectedIndices, searchWindows, 0);
}

template<matchingFunctions matchFunc, Contiguity::bidimensionalContiguity contiguity = Contiguity::Queen, dispDirection dDir = dispDirection::RightToLeft>
Multidim::Array<float, 3> refineBarycentricSymmetric2dDisp(Multidim::Array<float, 3> const& feature_vol_l, Multidim::Array<float, 3> const& feature_vol_r,
														   Multidim::Array<disp_t, 3> const& selectedIndices, searchOffset<2> const& searchWindows) {
	return B200Detail::barycentric2d<matchFunc, contiguity, dDir>(feature_vol_l, feature_vol_r, selectedIndices, searchWindows, 1);
}

template<matchingFunctions matchFunc, dispDirection dDir = dispDirection::RightToLeft>
Multidim::Array<float, 2> refineCostSymmetricDisp(Multidim::Array<float, 3> const& feature_vol_l, Multidim::Array<float, 3> const& feature_vol_r,
												  Multidim::Array<disp_t, 2> const& selectedIndex, Multidim::Array<float, 3> const& cost_volume) {
	Multidim::Array<float, 2> out(B200Detail::sourceShape<dDir>(feature_vol_l, feature_vol_r));
	stevi_array l = B200::describe(feature_vol_l), r = B200::describe(feature_vol_r), i = B200::describe(selectedIndex), c = B200::describe(cost_volume),
				o = B200::describe(out);
	if (!B200::ok(stevi_refine_cost_symmetric_disp(B200::threadContext(), int(matchFunc), &l, &r, &i, &c, int(dDir), &o))) return Multidim::Array<float, 2>();
	return out;
}

template<matchingFunctions matchFunc, dispDirection dDir = dispDirection::RightToLeft, int refineRadius = 1>
Multidim::Array<float, 2> refinedBarycentricSymmetricDispFeatureVol(Multidim::Array<float, 3> const& feature_vol_l, Multidim::Array<float, 3> const& feature_vol_r,
																	disp_t searchRange, bool preNormalize = false) {
	auto v = B200Detail::refinementVolumes<matchFunc>(feature_vol_l, feature_vol_r, preNormalize);
	auto CV = aggregateCost<matchFunc, float, float, dDir>(v.match_l, v.match_r, searchRange);
	if (CV.empty()) return Multidim::Array<float, 2>();
	auto disp = extractSelectedIndex<MatchingFunctionTraits<matchFunc>::extractionStrategy>(CV);
	return refineBarycentricSymmetricDisp<matchFunc, refineRadius, dDir>(v.fine_l, v.fine_r, disp, searchRange);
}

template<matchingFunctions matchFunc, dispDirection dDir = dispDirection::RightToLeft>
Multidim::Array<float, 2> refinedBarycentricDispFeatureVol(Multidim::Array<float, 3> const& feature_vol_l, Multidim::Array<float, 3> const& feature_vol_r,
														   disp_t searchRange, bool preNormalize = false) {
	auto v = B200Detail::refinementVolumes<matchFunc>(feature_vol_l, feature_vol_r, preNormalize);
	auto CV = aggregateCost<matchFunc, float, float, dDir>(v.match_l, v.match_r, searchRange);
	if (CV.empty()) return Multidim::Array<float, 2>();
	auto disp = extractSelectedIndex<MatchingFunctionTraits<matchFunc>::extractionStrategy>(CV);
	return refineBarycentricDisp<matchFunc, dDir>(v.fine_l, v.fine_r, disp);
}

template<matchingFunctions matchFunc, dispDirection dDir = dispDirection::RightToLeft>
Multidim::Array<float, 2> refinedCostSymmetricDispFeatureVol(Multidim::Array<float, 3> const& feature_vol_l, Multidim::Array<float, 3> const& feature_vol_r,
															 disp_t searchRange, bool preNormalize = false) {
	auto v = B200Detail::refinementVolumes<matchFunc>(feature_vol_l, feature_vol_r, preNormalize);
	auto CV = aggregateCost<matchFunc, float, float, dDir>(v.match_l, v.match_r, searchRange);
	if (CV.empty()) return Multidim::Array<float, 2>();
	auto disp = extractSelectedIndex<MatchingFunctionTraits<matchFunc>::extractionStrategy>(CV);
	return refineCostSymmetricDisp<matchFunc, dDir>(v.fine_l, v.fine_r, disp, CV);
}

// the image-level forms: unfold both images, then the feature-volume form; empty array when the heights (or channel counts) differ
template<matchingFunctions matchFunc, class T_L, class T_R, int nImDim = 2, dispDirection dDir = dispDirection::RightToLeft, int refineRadius = 1>
Multidim::Array<float, 2> refinedBarycentricSymmetricDisp(Multidim::Array<T_L, nImDim> const& img_l, Multidim::Array<T_R, nImDim> const& img_r,
														  uint8_t h_radius, uint8_t v_radius, disp_t searchRange, bool preNormalize = false) {
	if (img_l.shape()[0] != img_r.shape()[0]) return Multidim::Array<float, 2>();
	if (nImDim == 3 && img_l.shape()[nImDim - 1] != img_r.shape()[nImDim - 1]) return Multidim::Array<float, 2>();
	return refinedBarycentricSymmetricDispFeatureVol<matchFunc, dDir, refineRadius>(unfold<T_L, float>(h_radius, v_radius, img_l),
																				   unfold<T_R, float>(h_radius, v_radius, img_r), searchRange, preNormalize);
}

template<matchingFunctions matchFunc, class T_L, class T_R, int nImDim = 2, dispDirection dDir = dispDirection::RightToLeft>
Multidim::Array<float, 2> refinedBarycentricDisp(Multidim::Array<T_L, nImDim> const& img_l, Multidim::Array<T_R, nImDim> const& img_r,
												 uint8_t h_radius, uint8_t v_radius, disp_t searchRange, bool preNormalize = false) {
	if (img_l.shape()[0] != img_r.shape()[0]) return Multidim::Array<float, 2>();
	if (nImDim == 3 && img_l.shape()[nImDim - 1] != img_r.shape()[nImDim - 1]) return Multidim::Array<float, 2>();
	return refinedBarycentricDispFeatureVol<matchFunc, dDir>(unfold<T_L, float>(h_radius, v_radius, img_l), unfold<T_R, float>(h_radius, v_radius, img_r),
															 searchRange, preNormalize);
}

template<matchingFunctions matchFunc, class T_L, class T_R, int nImDim = 2, dispDirection dDir = dispDirection::RightToLeft>
Multidim::Array<float, 2> refinedBarycentricDisp(Multidim::Array<T_L, nImDim> const& img_l, Multidim::Array<T_R, nImDim> const& img_r,
												 UnFoldCompressor const& compressor, disp_t searchRange, bool preNormalize = false) {
	if (img_l.shape()[0] != img_r.shape()[0]) return Multidim::Array<float, 2>();
	if (nImDim == 3 && img_l.shape()[nImDim - 1] != img_r.shape()[nImDim - 1]) return Multidim::Array<float, 2>();
	return refinedBarycentricDispFeatureVol<matchFunc, dDir>(unfold<T_L, float>(compressor, img_l), unfold<T_R, float>(compressor, img_r), searchRange, preNormalize);
}

template<matchingFunctions matchFunc, class T_L, class T_R, int nImDim = 2, dispDirection dDir = dispDirection::RightToLeft>
Multidim::Array<float, 2> refinedCostSymmetricDisp(Multidim::Array<T_L, nImDim> const& img_l, Multidim::Array<T_R, nImDim> const& img_r,
												   uint8_t h_radius, uint8_t v_radius, disp_t searchRange, bool preNormalize = false) {
	if (img_l.shape()[0] != img_r.shape()[0]) return Multidim::Array<float, 2>();
	if (nImDim == 3 && img_l.shape()[nImDim - 1] != img_r.shape()[nImDim - 1]) return Multidim::Array<float, 2>();
	return refinedCostSymmetricDispFeatureVol<matchFunc, dDir>(unfold<T_L, float>(h_radius, v_radius, img_l), unfold<T_R, float>(h_radius, v_radius, img_r),
															   searchRange, preNormalize);
}

} // namespace Correlation
} // namespace StereoVision

#endif
