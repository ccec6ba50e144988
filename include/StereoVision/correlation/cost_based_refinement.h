// B200 facade of LibStevi's correlation/cost_based_refinement.h: InterpolationKernel :30-35,
// refineDispCostInterpolation :128-163 (empty array when the depth is even or < 3, :141-143).
#ifndef STEVI_B200_CORRELATION_COST_BASED_REFINEMENT_H
#define STEVI_B200_CORRELATION_COST_BASED_REFINEMENT_H

#include "./correlation_base.h"

namespace StereoVision {
namespace Correlation {

enum class InterpolationKernel { Equiangular, Parabola, Gaussian };

template<InterpolationKernel kernel = InterpolationKernel::Equiangular>
Multidim::Array<float, 2> refineDispCostInterpolation(Multidim::Array<float, 3> const& truncatedCostVolume,
													  Multidim::Array<disp_t, 2> const& rawDisparity) {
	auto sh = truncatedCostVolume.shape();
	Multidim::Array<float, 2> out(sh[0], sh[1]);
	stevi_array t = B200::describe(truncatedCostVolume), i = B200::describe(rawDisparity), o = B200::describe(out);
	if (!B200::ok(stevi_refine_disp_cost_interp(B200::threadContext(), &t, &i, int(kernel), &o))) return Multidim::Array<float, 2>();
	return out;
}

} // namespace Correlation
} // namespace StereoVision

#endif
