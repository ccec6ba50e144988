// StereoVision::Margins / PaddingMargins — the API vocabulary of LibStevi's utils/margins.h:24-163 that the dense-stereo
// entry points take.
#ifndef STEVI_B200_UTILS_MARGINS_H
#define STEVI_B200_UTILS_MARGINS_H

namespace StereoVision {

class Margins {
public:
	Margins(int all = 0) : _left(all), _top(all), _right(all), _bottom(all) {}
	Margins(int leftright, int topbottom) : _left(leftright), _top(topbottom), _right(leftright), _bottom(topbottom) {}
	Margins(int left, int top, int right, int bottom) : _left(left), _top(top), _right(right), _bottom(bottom) {}
	int left() const { return _left; }
	int top() const { return _top; }
	int right() const { return _right; }
	int bottom() const { return _bottom; }
protected:
	int _left, _top, _right, _bottom;
};

// default-constructed = automatic padding (the unfold operators then pad by the window radii)
class PaddingMargins : public Margins {
public:
	PaddingMargins() : Margins(), _auto(true) {}
	PaddingMargins(int padding) : Margins(padding), _auto(false) {}
	PaddingMargins(int leftright, int topbottom) : Margins(leftright, topbottom), _auto(false) {}
	PaddingMargins(int left, int top, int right, int bottom) : Margins(left, top, right, bottom), _auto(false) {}
	PaddingMargins(Margins const& m) : Margins(m), _auto(false) {}
	bool isAuto() const { return _auto; }
protected:
	bool _auto;
};

} // namespace StereoVision

#endif
