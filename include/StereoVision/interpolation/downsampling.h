// B200 facade of LibStevi's interpolation/downsampling.h: DownSampleWindows :30-65, averagePoolingDownsample :67-118.
#ifndef STEVI_B200_INTERPOLATION_DOWNSAMPLING_H
#define STEVI_B200_INTERPOLATION_DOWNSAMPLING_H

#include "../stevi_b200_bridge.h"

namespace StereoVision {
namespace Interpolation {

class DownSampleWindows {
public:
	DownSampleWindows(int size) : _horizontal(size), _vertical(size) {}
	DownSampleWindows(int h_size, int v_size) : _horizontal(h_size), _vertical(v_size) {}
	int horizontal() const { return _horizontal; }
	int vertical() const { return _vertical; }
private:
	int _horizontal, _vertical;
};

template<class T_I, class T_O = float, Multidim::ArrayDataAccessConstness Cstn = Multidim::NonConstView>
Multidim::Array<T_O, 2> averagePoolingDownsample(Multidim::Array<T_I, 2, Cstn> const& input, DownSampleWindows const& windows) {
	static_assert(std::is_same_v<T_O, float>, "the B200 library pools into float");
	auto sh = input.shape();
	Multidim::Array<T_O, 2> out((sh[0] + windows.vertical() - 1) / windows.vertical(), (sh[1] + windows.horizontal() - 1) / windows.horizontal());
	stevi_array i = B200::describe(input), o = B200::describe(out);
	if (!B200::ok(stevi_avgpool_downsample(B200::threadContext(), &i, windows.horizontal(), windows.vertical(), &o))) return Multidim::Array<T_O, 2>();
	return out;
}

} // namespace Interpolation
} // namespace StereoVision

#endif
