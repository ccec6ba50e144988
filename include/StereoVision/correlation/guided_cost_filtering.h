// B200 facade of LibStevi's correlation/guided_cost_filtering.h: computeAggregationWindows :26-100, variableCostVolumeAggregation
// :102-176. The volume-sized work — the weighted window means over the (H,W,D) cost volume — runs on the GPU
// (libstevi_b200/csrc/aggregate.cu). computeAggregationWindows evaluates two functors supplied by the caller once per window
// cell: that is caller code, which only the host can run; it is evaluated here, one image row per OpenMP thread when the
// translation unit is built with -fopenmp, and its result goes to the GPU with the aggregation call.
#ifndef STEVI_B200_CORRELATION_GUIDED_COST_FILTERING_H
#define STEVI_B200_CORRELATION_GUIDED_COST_FILTERING_H

#include "./correlation_base.h"

namespace StereoVision {
namespace Correlation {

// weights(i, j, di + v_radius, dj + h_radius) = spatialWeightFunc(di, dj) * matchWeightFunc(guide(i,j,:), guide(i+di,j+dj,:)), 0 for
// window cells outside the image (:42-90). Default strides, like the reference's Array(dims...).
template<typename ComputeT, typename GuideT, int guideDims, typename SpatialWeightFuncT, typename MatchWeightFuncT>
Multidim::Array<ComputeT, 4> computeAggregationWindows(Multidim::Array<GuideT, guideDims> const& guide, uint8_t h_radius, uint8_t v_radius,
													   SpatialWeightFuncT const& spatialWeightFunc, MatchWeightFuncT const& matchWeightFunc) {
	static_assert(guideDims == 2 || guideDims == 3, "Can only use a guide with two image dimensions, or two image dimensions plus one channel dimension");
	const int H = guide.shape()[0], W = guide.shape()[1], nChannels = guideDims == 3 ? guide.shape()[guideDims - 1] : 1;
	const int hr = h_radius, vr = v_radius;
	Multidim::Array<ComputeT, 4> ret(H, W, 2 * vr + 1, 2 * hr + 1);
	auto pixel = [&](int i, int j, Multidim::Array<GuideT, 1>& v) {
		for (int c = 0; c < nChannels; c++) {
			std::array<int, guideDims> idx{};
			idx[0] = i; idx[1] = j;
			if (guideDims == 3) idx[guideDims - 1] = c;
			v.atUnchecked(c) = guide.valueUnchecked(idx);
		}
	};
#ifdef _OPENMP
#pragma omp parallel for
#endif
	for (int i = 0; i < H; i++) {
		Multidim::Array<GuideT, 1> refVec(nChannels), targetVec(nChannels);
		for (int j = 0; j < W; j++) {
			pixel(i, j, refVec);
			for (int di = -vr; di <= vr; di++) for (int dj = -hr; dj <= hr; dj++) {
				ComputeT w = 0;
				if (i + di >= 0 && i + di < H && j + dj >= 0 && j + dj < W) {
					pixel(i + di, j + dj, targetVec);
					const ComputeT wSpatial = spatialWeightFunc(di, dj);
					w = wSpatial * matchWeightFunc(refVec, targetVec);
				}
				ret.atUnchecked(i, j, di + vr, dj + hr) = w;
			}
		}
	}
	return ret;
}

// :102-176: out(i,j,d) = sum of w * cv over the in-image part of the pixel's window / sum of those w; empty array for windows of
// another image size or with an even / empty side (:108-130). Result in the cost volume's own layout.
template<typename T_CV, typename ComputeT>
Multidim::Array<T_CV, 3> variableCostVolumeAggregation(Multidim::Array<T_CV, 3> const& costVolume, Multidim::Array<ComputeT, 4> const& aggregationWindows) {
	static_assert(std::is_same_v<T_CV, float> && std::is_same_v<ComputeT, float>, "float volumes and weights on the B200 path");
	Multidim::Array<T_CV, 3> out(costVolume.shape(), costVolume.strides());
	stevi_array c = B200::describe(costVolume), w = B200::describe(aggregationWindows), o = B200::describe(out);
	if (!B200::ok(stevi_variable_cost_volume_aggregation(B200::threadContext(), &c, &w, &o))) return Multidim::Array<T_CV, 3>();
	return out;
}

} // namespace Correlation
} // namespace StereoVision

#endif
