// B200 facade of LibStevi's correlation/unfold.h: UnfoldPatchOrientation :139-144, channelFromCord :171-191,
// unfold(h_radius, v_radius, img, padding, orientation) :248-344, UnFoldCompressor :36-137, unfold(compressor, img, padding)
// :346-471. The cost-volume entry points never materialise these volumes; they are here for callers of the feature API.
#ifndef STEVI_B200_CORRELATION_UNFOLD_H
#define STEVI_B200_CORRELATION_UNFOLD_H

#include <vector>
#include <algorithm>
#include "../stevi_b200_bridge.h"
#include "../utils/margins.h"

namespace StereoVision {
namespace Correlation {

enum UnfoldPatchOrientation { Rotate0, Rotate90, Rotate180, Rotate270 };

inline int channelFromCord(int vertical, int horizontal, int channel, int hSize, int vSize, int channels,
						   UnfoldPatchOrientation orientation = Rotate0) {
	switch (orientation) {
	case Rotate0: return channels * hSize * vertical + channels * horizontal + channel;
	case Rotate90: return channels * vSize * (hSize - horizontal - 1) + channels * vertical + channel;
	case Rotate180: return channels * hSize * (vSize - vertical - 1) + channels * (hSize - horizontal - 1) + channel;
	case Rotate270: return channels * vSize * horizontal + channels * (vSize - vertical - 1) + channel;
	}
	return -1;
}

namespace B200Detail {
inline const int32_t* paddingArg(PaddingMargins const& p, int32_t (&buf)[4]) {
	if (p.isAuto()) return nullptr;
	buf[0] = p.left(); buf[1] = p.top(); buf[2] = p.right(); buf[3] = p.bottom();
	return buf;
}
}

template<class T_I, class T_O = float, int nDim, Multidim::ArrayDataAccessConstness constness = Multidim::NonConstView>
Multidim::Array<T_O, 3> unfold(uint8_t h_radius, uint8_t v_radius, Multidim::Array<T_I, nDim, constness> const& in_data,
							   PaddingMargins const& padding = PaddingMargins(), UnfoldPatchOrientation orientation = Rotate0) {
	static_assert(nDim == 2 || nDim == 3, "unfold takes (H,W) or (H,W,C) images");
	const int h = 2 * h_radius + 1, v = 2 * v_radius + 1;
	auto sh = in_data.shape();
	const int f = nDim == 3 ? sh[nDim - 1] : 1;
	const int pl = padding.isAuto() ? h_radius : padding.left(), pr = padding.isAuto() ? h_radius : padding.right();
	const int pt = padding.isAuto() ? v_radius : padding.top(), pb = padding.isAuto() ? v_radius : padding.bottom();
	const int oh = sh[0] - v + pt + pb + 1, ow = sh[1] - h + pl + pr + 1, F = h * v * f;
	// unfold.h:272: the 2-D overload lays its result out with strides {W*F, F, 1}; the 3-D one uses the default strides
	Multidim::Array<T_O, 3> out = nDim == 2 ? Multidim::Array<T_O, 3>({oh, ow, F}, {ow * F, F, 1}) : Multidim::Array<T_O, 3>(oh, ow, F);
	int32_t buf[4];
	stevi_array i = B200::describe(in_data), o = B200::describe(out);
	if (!B200::ok(stevi_unfold(B200::threadContext(), &i, h_radius, v_radius, B200Detail::paddingArg(padding, buf), int(orientation), &o)))
		return Multidim::Array<T_O, 3>();
	return out;
}

class UnFoldCompressor {
public:
	template<Multidim::ArrayDataAccessConstness viewConstness>
	explicit UnFoldCompressor(Multidim::Array<int, 2, viewConstness> const& mask) {
		_mh = mask.shape()[0]; _mw = mask.shape()[1];
		_mask.resize(size_t(_mh) * _mw);
		const int vo = _mh / 2, ho = _mw / 2;
		int minH = 0, maxH = 0, minW = 0, maxW = 0;
		std::vector<int> labels;
		for (int i = 0; i < _mh; i++) for (int j = 0; j < _mw; j++) {
			const int feat = mask.valueUnchecked(i, j);
			_mask[size_t(i) * _mw + j] = feat;
			if (feat <= 0) continue;
			minH = std::min(minH, i - vo); maxH = std::max(maxH, i - vo);
			minW = std::min(minW, j - ho); maxW = std::max(maxW, j - ho);
			if (std::find(labels.begin(), labels.end(), feat) == labels.end()) labels.push_back(feat);
		}
		_height = maxH - minH + 1; _width = maxW - minW + 1;
		_margins = PaddingMargins(-minW, -minH, maxW, maxH);
		_nFeatures = int(labels.size());
	}
	int nFeatures() const { return _nFeatures; }
	int width() const { return _width; }
	int height() const { return _height; }
	PaddingMargins margins() const { return _margins; }
	Multidim::Array<int, 2> mask() const { return Multidim::Array<int, 2>(const_cast<int*>(_mask.data()), {_mh, _mw}, {_mw, 1}, false); }
protected:
	int _nFeatures = 0, _width = 0, _height = 0, _mh = 0, _mw = 0;
	PaddingMargins _margins;
	std::vector<int> _mask;
};

template<class T_I, class T_O = float, int nDim, Multidim::ArrayDataAccessConstness constness = Multidim::NonConstView>
Multidim::Array<T_O, 3> unfold(UnFoldCompressor const& compressor, Multidim::Array<T_I, nDim, constness> const& in_data,
							   PaddingMargins const& padding = PaddingMargins()) {
	static_assert(std::is_same_v<T_O, float>, "the compressed unfold returns float features");
	static_assert(nDim == 2 || nDim == 3, "unfold takes (H,W) or (H,W,C) images");
	PaddingMargins m = padding.isAuto() ? compressor.margins() : padding;
	auto sh = in_data.shape();
	const int f = nDim == 3 ? sh[nDim - 1] : 1;
	const int oh = sh[0] - compressor.height() + m.top() + m.bottom() + 1, ow = sh[1] - compressor.width() + m.left() + m.right() + 1;
	Multidim::Array<T_O, 3> out(oh, ow, compressor.nFeatures() * f);
	Multidim::Array<int, 2> mask = compressor.mask();
	int32_t buf[4];
	stevi_array k = B200::describe(mask), i = B200::describe(in_data), o = B200::describe(out);
	if (!B200::ok(stevi_unfold_compressed(B200::threadContext(), &k, &i, B200Detail::paddingArg(padding, buf), &o))) return Multidim::Array<T_O, 3>();
	return out;
}

// The two super-pixel masks the reference ships for UnFoldCompressor (unfold.h:476-691): 17 super-pixels over a 7x7 / 9x9
// window. Data, kept as tables (checked against the reference's functions by tests/test_facade.py).
namespace CompressorGenerators {

namespace Detail {
inline Multidim::Array<int, 2> maskFromTable(const int* t, int n) {
	Multidim::Array<int, 2> out(n, n);
	for (int i = 0; i < n; i++) for (int j = 0; j < n; j++) out.atUnchecked(i, j) = t[i * n + j];
	return out;
}
} // namespace Detail

inline Multidim::Array<int, 2> GrPix17R3Filter() {
	static const int t[49] = {
		14, 14, 10, 10, 10, 16, 16,
		14, 14,  6,  4,  7, 16, 16,
		11,  6,  6,  4,  7,  7, 13,
		11,  2,  2,  1,  3,  3, 13,
		11,  8,  8,  5,  9,  9, 13,
		15, 15,  8,  5,  9, 17, 17,
		15, 15, 12, 12, 12, 17, 17};
	return Detail::maskFromTable(t, 7);
}

inline Multidim::Array<int, 2> GrPix17R4Filter() {
	static const int t[81] = {
		14, 14, 14, 10, 10, 10, 16, 16, 16,
		14, 14, 14, 10, 10, 10, 16, 16, 16,
		14, 14,  6,  6,  4,  7,  7, 16, 16,
		11, 11,  6,  6,  4,  7,  7, 13, 13,
		11, 11,  2,  2,  1,  3,  3, 13, 13,
		11, 11,  8,  8,  5,  9,  9, 13, 13,
		15, 15,  8,  8,  5,  9,  9, 17, 17,
		15, 15, 15, 12, 12, 12, 17, 17, 17,
		15, 15, 15, 12, 12, 12, 17, 17, 17};
	return Detail::maskFromTable(t, 9);
}

} // namespace CompressorGenerators

} // namespace Correlation
} // namespace StereoVision

#endif
