// B200 facade of LibStevi's correlation/sgm.h: sgmDirections :29-46, sgmCostVolume :360-404. Compat semantics: the
// reference's behaviour as written, bit-exact for integer-valued volumes.
#ifndef STEVI_B200_CORRELATION_SGM_H
#define STEVI_B200_CORRELATION_SGM_H

#include "./correlation_base.h"
#include "../utils/margins.h"

namespace StereoVision {
namespace Correlation {

enum class sgmDirections {
	Up2Down, Down2Up, Left2Right, Right2Left, UpLeft2DownRight, DownRight2UpLeft, UpRight2DownLeft, DownLeft2UpRight,
	UpLeft2Right, DownRight2Left, UpRight2Left, DownLeft2Right, UpLeft2Down, DownRight2Up, UpRight2Down, DownLeft2Up
};

template<int nDirections = 8, dispExtractionStartegy strategy = dispExtractionStartegy::Cost, class T_CV = float>
Multidim::Array<float, 3> sgmCostVolume(Multidim::Array<T_CV, 3> const& cost_volume, float P1, float P2, Margins const& margins,
										float Pout = 100) {
	static_assert(nDirections == 4 || nDirections == 8 || nDirections == 16, "sgmCostVolume: 4, 8 or 16 directions");
	static_assert(std::is_same_v<T_CV, uint16_t> || std::is_same_v<T_CV, float>, "cost volume type not covered by the B200 library");
	auto sh = cost_volume.shape();
	Multidim::Array<float, 3> out(sh);   // default strides, sgm.h:369
	stevi_margins m{margins.left(), margins.top(), margins.right(), margins.bottom()};
	stevi_array v = B200::describe(cost_volume), o = B200::describe(out);
	if (!B200::ok(stevi_sgm(B200::threadContext(), &v, nDirections, int(strategy), P1, P2, Pout, &m, STEVI_COMPAT, &o)))
		return Multidim::Array<float, 3>();
	return out;
}

} // namespace Correlation
} // namespace StereoVision

#endif
