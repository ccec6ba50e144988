// B200 facade of LibStevi's correlation/patchmatch.h and correlation/on_demand_cost_volume.h.
//   patchMatch<matchFunc, searchSpaceDim>(feature_vol_s, feature_vol_t, searchOffset, nIter, nRandomSearch, initializer)   patchmatch.h:498-556
//   cachelessPatchMatch                                                                                               :561-626 (same results: the
//                                                                                       cacheless volume only changes where the reference keeps costs)
//   OnDemandStereoCostVolume / OnDemandImageFlowVolume ::truncatedCostVolume(disp, radius)                           on_demand_cost_volume.h:187-330
// The reference asks a lazily filled host-side volume for one cell at a time; here cells are evaluated on the GPU where they are
// needed (libstevi_b200/csrc/patchmatch.cu) and no volume exists. costValue(pos, disp) for a single cell is not offered: a
// device call per cell is the wrong granularity — ask for the cells around a whole disparity map with truncatedCostVolume.
// Random draws: the reference seeds std::default_random_engine from std::random_device, so two of its runs differ; here they
// come from a counter-based generator keyed by B200::patchMatchSeed() (set it for other draws). The NumbersCache argument of the
// reference is accepted and ignored.
#ifndef STEVI_B200_CORRELATION_PATCHMATCH_H
#define STEVI_B200_CORRELATION_PATCHMATCH_H

#include <functional>
#include <optional>
#include "./cross_correlations.h"

namespace StereoVision {
namespace Random { template<typename T> class NumbersCache {}; }   // utils/randomcache.h: accepted for source compatibility

namespace B200 { inline uint64_t& patchMatchSeed() { static thread_local uint64_t seed = 0; return seed; } }

namespace Correlation {

template<matchingFunctions matchFunc, class FV_S_T, class FV_T_T, int searchSpaceDim, bool useUncachedCV = false>
struct PatchMatchTraits {
	static_assert(searchSpaceDim == 1 || searchSpaceDim == 2, "invalid searchSpaceDim");
	typedef std::function<Multidim::Array<disp_t, 3>(FV_S_T const&, FV_T_T const&)> PatchMatchInitializer;
};

template<matchingFunctions matchFunc, int searchSpaceDim>
Multidim::Array<disp_t, 3> patchMatch(Multidim::Array<float, 3> const& feature_vol_s_p, Multidim::Array<float, 3> const& feature_vol_t_p,
									  searchOffset<searchSpaceDim> searchOffset, int nIter = 5, int nRandomSearch = 4,
									  std::optional<typename PatchMatchTraits<matchFunc, Multidim::Array<float, 3>, Multidim::Array<float, 3>, searchSpaceDim>::PatchMatchInitializer> initializer = std::nullopt,
									  std::optional<Random::NumbersCache<int>> randcache = std::nullopt) {
	static_assert(searchSpaceDim == 1 || searchSpaceDim == 2, "patchMatch function can only be used to search in 1 or two dimension !");
	static_assert(MatchingFunctionTraits<matchFunc>::onB200 && !MatchingFunctionTraits<matchFunc>::isCensusBased, "float matching functions");
	(void)randcache;
	auto s_shape = feature_vol_s_p.shape();
	Multidim::Array<disp_t, 3> disp({s_shape[0], s_shape[1], searchSpaceDim}, {s_shape[1] * searchSpaceDim, searchSpaceDim, 1});
	int32_t so[2 * searchSpaceDim];
	for (int k = 0; k < searchSpaceDim; k++) { so[2 * k] = searchOffset.lowerOffset(k); so[2 * k + 1] = searchOffset.upperOffset(k); }
	Multidim::Array<disp_t, 3> init;
	if (initializer.has_value()) init = initializer.value()(feature_vol_s_p, feature_vol_t_p);
	stevi_array s = B200::describe(feature_vol_s_p), t = B200::describe(feature_vol_t_p), d = B200::describe(disp), i = B200::describe(init);
	if (!B200::ok(stevi_patch_match(B200::threadContext(), int(matchFunc), &s, &t, searchSpaceDim, so, nIter, nRandomSearch, B200::patchMatchSeed(),
									initializer.has_value() ? &i : nullptr, &d)))
		return Multidim::Array<disp_t, 3>();   // :531-539
	return disp;
}

template<matchingFunctions matchFunc, int searchSpaceDim>
Multidim::Array<disp_t, 3> cachelessPatchMatch(Multidim::Array<float, 3> const& f_s_p, Multidim::Array<float, 3> const& f_t_p,
											   searchOffset<searchSpaceDim> searchOffset, int nIter = 5, int nRandomSearch = 4,
											   std::optional<typename PatchMatchTraits<matchFunc, Multidim::Array<float, 3>, Multidim::Array<float, 3>, searchSpaceDim>::PatchMatchInitializer> initializer = std::nullopt,
											   std::optional<Random::NumbersCache<int>> randcache = std::nullopt) {
	return patchMatch<matchFunc, searchSpaceDim>(f_s_p, f_t_p, searchOffset, nIter, nRandomSearch, initializer, randcache);
}

namespace B200Detail {
template<matchingFunctions matchFunc, int searchSpaceDim>
class OnDemandVolume {
public:
	OnDemandVolume(Multidim::Array<float, 3> const& source, Multidim::Array<float, 3> const& target, searchOffset<searchSpaceDim> const& searchSpace)
		: _source(source), _target(target), _search(searchSpace) {}
	searchOffset<searchSpaceDim> const& searchSpace() const { return _search; }
	// (H, W, range...) — the shape of the volume the cells belong to (on_demand_cost_volume.h:101-103)
	std::array<int, 2 + searchSpaceDim> shape() const {
		std::array<int, 2 + searchSpaceDim> s{};
		s[0] = _source.shape()[0]; s[1] = _source.shape()[1];
		for (int k = 0; k < searchSpaceDim; k++) s[2 + k] = _search.dimRange(k);
		return s;
	}
	// :187-310: disp (H,W,searchSpaceDim), or (H,W) for a 1-D search
	template<int dispDims>
	Multidim::Array<float, 2 + searchSpaceDim> truncatedCostVolume(Multidim::Array<disp_t, dispDims> const& disp, int radius = 1) const {
		static_assert(dispDims == 3 || (dispDims == 2 && searchSpaceDim == 1), "disparity map of the search's dimensions");
		std::array<int, 2 + searchSpaceDim> shape{};
		shape[0] = _source.shape()[0]; shape[1] = _source.shape()[1];
		for (int k = 0; k < searchSpaceDim; k++) shape[2 + k] = 2 * radius + 1;
		Multidim::Array<float, 2 + searchSpaceDim> out(shape);
		int32_t so[2 * searchSpaceDim];
		for (int k = 0; k < searchSpaceDim; k++) { so[2 * k] = _search.lowerOffset(k); so[2 * k + 1] = _search.upperOffset(k); }
		stevi_array s = B200::describe(_source), t = B200::describe(_target), d = B200::describe(disp), o = B200::describe(out);
		if (!B200::ok(stevi_on_demand_truncated_cost_volume(B200::threadContext(), int(matchFunc), &s, &t, searchSpaceDim, so, &d, radius, &o)))
			return Multidim::Array<float, 2 + searchSpaceDim>();
		return out;
	}
private:
	Multidim::Array<float, 3> const& _source;
	Multidim::Array<float, 3> const& _target;
	searchOffset<searchSpaceDim> _search;
};
} // namespace B200Detail

// on_demand_cost_volume.h:329-337 (float features, float costs; the search space is given as the searchOffset it is built from)
template<matchingFunctions matchFunc> using OnDemandStereoCostVolume = B200Detail::OnDemandVolume<matchFunc, 1>;
template<matchingFunctions matchFunc> using OnDemandImageFlowVolume = B200Detail::OnDemandVolume<matchFunc, 2>;

} // namespace Correlation
} // namespace StereoVision

#endif
