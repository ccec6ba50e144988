// B200 facade of LibStevi's correlation/correlation_base.h (dense-stereo part): same names, template parameters and
// defaults; bodies forward to the C ABI. Reference lines: enums :31-45, disp_t :47, extractSelectedIndex :427-464,
// selectedIndexToDisp :511-532, selectedCost :557-577, truncatedCostVolume :579-674, extractInBoundDomain :727-822.
#ifndef STEVI_B200_CORRELATION_BASE_H
#define STEVI_B200_CORRELATION_BASE_H

#include "../stevi_b200_bridge.h"

namespace StereoVision {
namespace Correlation {

enum class dispExtractionStartegy { Cost = 0, Score = 1 };   // (sic)
enum class dispDirection { LeftToRight = 0, RightToLeft = 1 };
enum class truncatedCostVolumeDirection { Same = 0, Reversed = 1, Both = 2 };

typedef int32_t disp_t;

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

} // namespace Correlation
} // namespace StereoVision

#endif
