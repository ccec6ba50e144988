// B200 facade of LibStevi's correlation/correlation_base.h (dense-stereo part): same names, template parameters and
// defaults; bodies forward to the C ABI. Reference lines: enums :31-45, disp_t :47, extractSelectedIndex :427-464,
// selectedIndexToDisp :511-532, selectedCost :557-577, truncatedCostVolume :579-674, extractInBoundDomain :727-822;
// 2-D search: searchOffset :288-409, extractSelected2dIndex :466-509, selected2dIndexToDisp :534-555,
// truncatedBidirectionaCostVolume :677-725; box filters: meanFilter2D :881-1004, sigmaFilter :1006-1077.
#ifndef STEVI_B200_CORRELATION_BASE_H
#define STEVI_B200_CORRELATION_BASE_H

#include <cstdlib>
#include "../stevi_b200_bridge.h"

namespace StereoVision {
namespace Correlation {

enum class dispExtractionStartegy { Cost = 0, Score = 1 };   // (sic)
enum class dispDirection { LeftToRight = 0, RightToLeft = 1 };
enum class truncatedCostVolumeDirection { Same = 0, Reversed = 1, Both = 2 };

typedef int32_t disp_t;

// Inclusive [lower, upper] search range per axis (reference :288-409): same constructor order (lower0, upper0, lower1, upper1, ...),
// same accessors. Plain host-side value type, nothing to offload.
template<int nDim>
class searchOffset {
public:
	searchOffset() : _valid(false) { _up.fill(0); _lo.fill(0); }
	template<typename... Ds>
	searchOffset(disp_t lowerOffset0, disp_t upperOffset0, Ds... nextOffsets) : _valid(true) {
		static_assert(sizeof...(nextOffsets) == 2 * (nDim - 1), "two offsets per dimension");
		const disp_t all[2 * nDim] = {lowerOffset0, upperOffset0, disp_t(nextOffsets)...};
		for (int k = 0; k < nDim; k++) { _lo[k] = all[2 * k]; _up[k] = all[2 * k + 1]; }
	}
	bool isValid() const { return _valid; }
	template<int dim> int& upperOffset() { return _up[dim]; }
	template<int dim> int& lowerOffset() { return _lo[dim]; }
	template<int dim> int const& upperOffset() const { return _up[dim]; }
	template<int dim> int const& lowerOffset() const { return _lo[dim]; }
	int& upperOffset(int dim) { return _up[dim]; }
	int& lowerOffset(int dim) { return _lo[dim]; }
	int const& upperOffset(int dim) const { return _up[dim]; }
	int const& lowerOffset(int dim) const { return _lo[dim]; }
	template<int dim> int dimRange() const { return dimRange(dim); }
	int dimRange(int dim) const { return _up[dim] - _lo[dim] + 1; }
	template<int dim> bool valueInRange(int val) const { return valueInRange(val, dim); }
	bool valueInRange(int val, int dim) const { return val >= _lo[dim] && val <= _up[dim]; }
	template<int dim> int setValueInRange(int val) const { return setValueInRange(val, dim); }
	int setValueInRange(int val, int dim) const { return std::abs(val % dimRange(dim)) + _lo[dim]; }
	template<int dim> int idx2disp(int idx) const { return _lo[dim] + idx; }
	int idx2disp(int dim, int idx) const { return _lo[dim] + idx; }
	template<int dim> int disp2idx(int disp) const { return disp - _lo[dim]; }
	int disp2idx(int dim, int disp) const { return disp - _lo[dim]; }
private:
	std::array<disp_t, nDim> _up, _lo;
	bool _valid;
};

namespace B200Detail {
inline std::array<int32_t, 4> search4(searchOffset<2> const& s) {
	return {s.lowerOffset(0), s.upperOffset(0), s.lowerOffset(1), s.upperOffset(1)};
}
} // namespace B200Detail

template<dispExtractionStartegy strategy, class T_CV>
Multidim::Array<disp_t, 2> extractSelectedIndex(Multidim::Array<T_CV, 3> const& costVolume) {
	auto sh = costVolume.shape();
	Multidim::Array<disp_t, 2> out(sh[0], sh[1]);
	stevi_array v = B200::describe(costVolume), o = B200::describe(out);
	if (!B200::ok(stevi_wta(B200::threadContext(), &v, int(strategy), &o))) return Multidim::Array<disp_t, 2>();
	return out;
}

template<typename DT = disp_t, dispDirection dDir = dispDirection::RightToLeft>
Multidim::Array<DT, 2> selectedIndexToDisp(Multidim::Array<disp_t, 2> const& selectedIndex, disp_t disp_offset = 0) {
	static_assert(std::is_same_v<DT, disp_t>, "the B200 library returns disp_t disparities");
	auto sh = selectedIndex.shape();
	Multidim::Array<DT, 2> out(sh[0], sh[1]);
	stevi_array i = B200::describe(selectedIndex), o = B200::describe(out);
	if (!B200::ok(stevi_index_to_disp(B200::threadContext(), &i, int(dDir), disp_offset, &o))) return Multidim::Array<DT, 2>();
	return out;
}

template<class T_CV>
Multidim::Array<T_CV, 2> selectedCost(Multidim::Array<T_CV, 3> const& costVolume, Multidim::Array<disp_t, 2> const& selectedIndex) {
	auto sh = costVolume.shape();
	Multidim::Array<T_CV, 2> out(sh[0], sh[1]);
	stevi_array v = B200::describe(costVolume), i = B200::describe(selectedIndex), o = B200::describe(out);
	if (!B200::ok(stevi_selected_cost(B200::threadContext(), &v, &i, &o))) return Multidim::Array<T_CV, 2>();
	return out;
}

template<class T_CV, dispDirection dir = dispDirection::RightToLeft, truncatedCostVolumeDirection sdir = truncatedCostVolumeDirection::Same>
Multidim::Array<T_CV, 3> truncatedCostVolume(Multidim::Array<T_CV, 3> const& costVolume, Multidim::Array<disp_t, 2> const& selectedIndex,
											 uint8_t h_radius, uint8_t v_radius, uint8_t cost_vol_radius) {
	static_assert(std::is_same_v<T_CV, float>, "truncatedCostVolume runs on float volumes on the B200 library");
	auto sh = costVolume.shape();
	const int depth = sdir == truncatedCostVolumeDirection::Both ? cost_vol_radius * 4 + 1 : cost_vol_radius * 2 + 1;
	Multidim::Array<T_CV, 3> out(sh[0], sh[1], depth);
	stevi_array v = B200::describe(costVolume), i = B200::describe(selectedIndex), o = B200::describe(out);
	if (!B200::ok(stevi_truncated_cost_volume(B200::threadContext(), &v, &i, h_radius, v_radius, cost_vol_radius, int(dir), int(sdir), &o)))
		return Multidim::Array<T_CV, 3>();
	return out;
}

template<class T_IB, dispDirection dir = dispDirection::RightToLeft, truncatedCostVolumeDirection sdir = truncatedCostVolumeDirection::Same>
Multidim::Array<T_IB, 3> extractInBoundDomain(Multidim::Array<disp_t, 2> const& selectedIndex, uint32_t width, uint8_t h_radius,
											  uint8_t v_radius, uint8_t cost_vol_radius) {
	auto sh = selectedIndex.shape();
	const int depth = sdir == truncatedCostVolumeDirection::Both ? cost_vol_radius * 4 + 1 : cost_vol_radius * 2 + 1;
	Multidim::Array<float, 3> mask(sh[0], sh[1], depth);
	stevi_array i = B200::describe(selectedIndex), o = B200::describe(mask);
	if (!B200::ok(stevi_in_bound_domain(B200::threadContext(), &i, int32_t(width), h_radius, v_radius, cost_vol_radius, int(dir), int(sdir), &o)))
		return Multidim::Array<T_IB, 3>();
	if constexpr (std::is_same_v<T_IB, float>) return mask;
	else return mask.template cast<T_IB>();
}

// :466-509 — (H,W,D1,D2) -> (H,W,2), last extremum in row-major (d1,d2) order
template<dispExtractionStartegy strategy, class T_CV>
Multidim::Array<disp_t, 3> extractSelected2dIndex(Multidim::Array<T_CV, 4> const& costVolume) {
	static_assert(std::is_same_v<T_CV, float>, "2-D search volumes are float on the B200 library");
	auto sh = costVolume.shape();
	Multidim::Array<disp_t, 3> out(sh[0], sh[1], 2);
	stevi_array v = B200::describe(costVolume), o = B200::describe(out);
	if (!B200::ok(stevi_wta_2d(B200::threadContext(), &v, int(strategy), &o))) return Multidim::Array<disp_t, 3>();
	return out;
}

// :534-555
template<typename DT>
Multidim::Array<DT, 3> selected2dIndexToDisp(Multidim::Array<DT, 3> const& selectedIndex, searchOffset<2> const& offset) {
	static_assert(std::is_same_v<DT, disp_t>, "the B200 library returns disp_t disparities");
	Multidim::Array<DT, 3> out(selectedIndex.shape());
	const auto so = B200Detail::search4(offset);
	stevi_array i = B200::describe(selectedIndex), o = B200::describe(out);
	if (!B200::ok(stevi_index2d_to_disp(B200::threadContext(), &i, so.data(), &o))) return Multidim::Array<DT, 3>();
	return out;
}

// :677-725 (sic). The radii are uint8_t with default uint8_t(-1) = 255, as in the reference.
template<class T_CV, dispDirection dir = dispDirection::RightToLeft>
Multidim::Array<T_CV, 4> truncatedBidirectionaCostVolume(Multidim::Array<T_CV, 4> const& costVolume, Multidim::Array<disp_t, 3> const& selectedIndex,
														 uint8_t cost_vol_radius0 = -1, uint8_t cost_vol_radius1 = -1) {
	static_assert(std::is_same_v<T_CV, float>, "2-D search volumes are float on the B200 library");
	auto sh = costVolume.shape();
	const int r0 = cost_vol_radius0 < 1 ? 1 : cost_vol_radius0, r1 = cost_vol_radius1 < 1 ? r0 : cost_vol_radius1;
	Multidim::Array<T_CV, 4> out(sh[0], sh[1], 2 * r0 + 1, 2 * r1 + 1);
	stevi_array v = B200::describe(costVolume), i = B200::describe(selectedIndex), o = B200::describe(out);
	if (!B200::ok(stevi_truncated_cost_volume_2d(B200::threadContext(), &v, &i, cost_vol_radius0, cost_vol_radius1, &o)))
		return Multidim::Array<T_CV, 4>();
	return out;
}

// meanFilter2D :881-1004 ((H,W) and (H,W,C) overloads) and sigmaFilter :1006-1077: same running-sum arithmetic as the reference
// (bit-identical interior); the border the reference leaves uninitialised is 0. An empty array when the window does not fit.
template<class T_I, int nDim>
Multidim::Array<float, 2> meanFilter2D(uint8_t h_radius, uint8_t v_radius, Multidim::Array<T_I, nDim> const& in_data) {
	static_assert(nDim == 2 || nDim == 3, "meanFilter2D takes (H,W) or (H,W,C) images");
	static_assert(std::is_same_v<T_I, float> || std::is_same_v<T_I, uint8_t>, "u8 or float images on the B200 library");
	auto sh = in_data.shape();
	Multidim::Array<float, 2> mean(sh[0], sh[1]);
	stevi_array i = B200::describe(in_data), o = B200::describe(mean);
	if (!B200::ok(stevi_mean_filter_2d(B200::threadContext(), &i, h_radius, v_radius, &o))) return Multidim::Array<float, 2>();
	return mean;
}

template<class T_I, int nDim>
Multidim::Array<float, 2> sigmaFilter(uint8_t h_radius, uint8_t v_radius, Multidim::Array<float, 2> const& mean,
									  Multidim::Array<T_I, nDim> const& in_data) {
	static_assert(nDim == 2 || nDim == 3, "sigmaFilter takes (H,W) or (H,W,C) images");
	static_assert(std::is_same_v<T_I, float> || std::is_same_v<T_I, uint8_t>, "u8 or float images on the B200 library");
	auto sh = in_data.shape();
	Multidim::Array<float, 2> sigma(sh[0], sh[1]);
	stevi_array m = B200::describe(mean), i = B200::describe(in_data), o = B200::describe(sigma);
	if (!B200::ok(stevi_sigma_filter(B200::threadContext(), &m, &i, h_radius, v_radius, &o))) return Multidim::Array<float, 2>();
	return sigma;
}

} // namespace Correlation
} // namespace StereoVision

#endif
