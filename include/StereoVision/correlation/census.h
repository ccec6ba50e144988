// B200 facade of LibStevi's correlation/census.h: censusFeatures :69-115, censusTransform2D :117-131.
// Compat semantics: only completely filled 32-bit words are written, the trailing word is 0.
#ifndef STEVI_B200_CORRELATION_CENSUS_H
#define STEVI_B200_CORRELATION_CENSUS_H

#include "./matching_costs.h"

namespace StereoVision {
namespace Correlation {

template<class T_I, Multidim::ArrayDataAccessConstness constness = Multidim::NonConstView>
Multidim::Array<census_data_t, 3> censusFeatures(Multidim::Array<T_I, 3, constness> const& feature_volume) {
	auto sh = feature_volume.shape();
	if (sh[2] <= 1) return Multidim::Array<census_data_t, 3>();   // census.h:75-77
	Multidim::Array<census_data_t, 3> out(sh[0], sh[1], (sh[2] - 1) / 32 + 1);
	stevi_array f = B200::describe(feature_volume), o = B200::describe(out);
	if (!B200::ok(stevi_census_features(B200::threadContext(), &f, STEVI_COMPAT, &o))) return Multidim::Array<census_data_t, 3>();
	return out;
}

template<class T_I, Multidim::ArrayDataAccessConstness constness = Multidim::NonConstView>
Multidim::Array<census_data_t, 3> censusTransform2D(Multidim::Array<T_I, 2, constness> const& input, int8_t h_radius, int8_t v_radius) {
	auto sh = input.shape();
	const int F = (2 * h_radius + 1) * (2 * v_radius + 1);
	if (F <= 1) return Multidim::Array<census_data_t, 3>();
	Multidim::Array<census_data_t, 3> out(sh[0], sh[1], (F - 1) / 32 + 1);
	stevi_array i = B200::describe(input), o = B200::describe(out);
	if (!B200::ok(stevi_census_transform(B200::threadContext(), &i, h_radius, v_radius, STEVI_COMPAT, &o))) return Multidim::Array<census_data_t, 3>();
	return out;
}

} // namespace Correlation
} // namespace StereoVision

#endif
