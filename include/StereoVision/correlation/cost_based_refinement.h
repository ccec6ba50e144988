// B200 facade of LibStevi's correlation/cost_based_refinement.h: InterpolationKernel :30-35,
// IsotropyHypothesis :37-41, refineDispCostInterpolation :128-163 (empty array when the depth is even or < 3, :141-143),
// refineDisp2dCostInterpolation :165-369 (empty array when a side is even or < 3, :180-182),
// refineDisp2dCostPatchInterpolation :371-433.
#ifndef STEVI_B200_CORRELATION_COST_BASED_REFINEMENT_H
#define STEVI_B200_CORRELATION_COST_BASED_REFINEMENT_H

#include "./correlation_base.h"

namespace StereoVision {
namespace Correlation {

enum class InterpolationKernel { Equiangular, Parabola, Gaussian };
enum class IsotropyHypothesis { Isotropic, Anisotropic };

template<InterpolationKernel kernel = InterpolationKernel::Equiangular>
Multidim::Array<float, 2> refineDispCostInterpolation(Multidim::Array<float, 3> const& truncatedCostVolume,
													  Multidim::Array<disp_t, 2> const& rawDisparity) {
	auto sh = truncatedCostVolume.shape();
	Multidim::Array<float, 2> out(sh[0], sh[1]);
	stevi_array t = B200::describe(truncatedCostVolume), i = B200::describe(rawDisparity), o = B200::describe(out);
	if (!B200::ok(stevi_refine_disp_cost_interp(B200::threadContext(), &t, &i, int(kernel), &o))) return Multidim::Array<float, 2>();
	return out;
}

template<InterpolationKernel kernel, IsotropyHypothesis isotropHypothesis = IsotropyHypothesis::Isotropic>
Multidim::Array<float, 3> refineDisp2dCostInterpolation(Multidim::Array<float, 4> const& truncatedCostVolume,
														Multidim::Array<disp_t, 3> const& rawDisparity) {
	Multidim::Array<float, 3> out(rawDisparity.shape());
	stevi_array t = B200::describe(truncatedCostVolume), i = B200::describe(rawDisparity), o = B200::describe(out);
	if (!B200::ok(stevi_refine_disp_2d_cost_interp(B200::threadContext(), &t, &i, int(kernel), int(isotropHypothesis), &o)))
		return Multidim::Array<float, 3>();
	return out;
}

// :371-433 (refineCostPatch :71-126): Parabola or Gaussian only, as the reference static_asserts
template<InterpolationKernel kernel>
Multidim::Array<float, 3> refineDisp2dCostPatchInterpolation(Multidim::Array<float, 4> const& truncatedCostVolume,
															 Multidim::Array<disp_t, 3> const& rawDisparity) {
	static_assert(kernel == InterpolationKernel::Parabola || kernel == InterpolationKernel::Gaussian, "Unsupported kernel used for patch cost refinement");
	Multidim::Array<float, 3> out(rawDisparity.shape());
	stevi_array t = B200::describe(truncatedCostVolume), i = B200::describe(rawDisparity), o = B200::describe(out);
	if (!B200::ok(stevi_refine_disp_2d_cost_patch_interp(B200::threadContext(), &t, &i, int(kernel), &o))) return Multidim::Array<float, 3>();
	return out;
}

} // namespace Correlation
} // namespace StereoVision

#endif
