// TEST INFRASTRUCTURE — not product code. Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline /
// --impl reference legs may load the library built from this file.
//
// Thin extern "C" wrapper over the UNMODIFIED reference headers (included from $STEVI_REF, default
// /root/reference, never copied into this repo). Built by oracle/Makefile into oracle/_ref/libstevi_ref.so.
// All buffers are dense C-order (last index fastest); wrappers build Multidim::Array views with explicit
// strides, call the reference template exactly as test/benchmarks/benchmarkCrossCorrelationAlgorithms.cpp
// (:93-95, :216-221, :289-293) does, and copy the result out element by element.
#include <cstdint>
#include <cstring>
#include <cmath>
#include <omp.h>

#include <correlation/cross_correlations.h>
#include <correlation/sgm.h>
#include <correlation/hierarchical.h>
#include <correlation/cost_based_refinement.h>
#include <correlation/census.h>
#include <correlation/unfold.h>
#include <interpolation/downsampling.h>
#include <correlation/guided_cost_filtering.h>

using namespace StereoVision;
using namespace StereoVision::Correlation;
namespace MD = Multidim;

namespace {

template<typename T>
MD::Array<T, 2> view2(const T* p, int h, int w) {
	return MD::Array<T, 2>(const_cast<T*>(p), {h, w}, {w, 1}, false);
}
template<typename T>
MD::Array<T, 3> view3(const T* p, int h, int w, int c) {
	return MD::Array<T, 3>(const_cast<T*>(p), {h, w, c}, {w * c, c, 1}, false);
}
template<typename T, typename TO>
void out3(MD::Array<T, 3> const& a, TO* out) {
	auto s = a.shape();
	for (int i = 0; i < s[0]; i++) for (int j = 0; j < s[1]; j++) for (int k = 0; k < s[2]; k++)
		out[(size_t(i) * s[1] + j) * s[2] + k] = static_cast<TO>(a.valueUnchecked(i, j, k));
}
template<typename T, typename TO>
void out2(MD::Array<T, 2> const& a, TO* out) {
	auto s = a.shape();
	for (int i = 0; i < s[0]; i++) for (int j = 0; j < s[1]; j++)
		out[size_t(i) * s[1] + j] = static_cast<TO>(a.valueUnchecked(i, j));
}

// copies the interior [vr, h-vr) x [hr, w-hr) of a filter result (the reference never writes the rest of its array)
template<typename T>
int interior(MD::Array<T, 2> const& a, int hr, int vr, float* out) {
	auto s = a.shape();
	if (s[1] < 2 * hr + 1 || s[0] < 2 * vr + 1) return 1;
	for (int i = vr; i < s[0] - vr; i++) for (int j = hr; j < s[1] - hr; j++) out[size_t(i) * s[1] + j] = a.valueUnchecked(i, j);
	return 0;
}

template<matchingFunctions F, typename TI, typename TCV>
int cv_dir(const TI* l, const TI* r, int h, int w, int wr, int hr, int vr, int D, int dir, TCV* out) {
	auto L = view2(l, h, w);
	auto R = view2(r, h, wr);
	if (dir == int(dispDirection::RightToLeft)) {
		auto cv = unfoldBasedCostVolume<F, TI, TI, 2, dispDirection::RightToLeft, TCV>(L, R, hr, vr, D);
		if (cv.empty()) return 1;
		out3(cv, out);
	} else {
		auto cv = unfoldBasedCostVolume<F, TI, TI, 2, dispDirection::LeftToRight, TCV>(L, R, hr, vr, D);
		if (cv.empty()) return 1;
		out3(cv, out);
	}
	return 0;
}

template<matchingFunctions F>
int cv2d_dir(const float* l, const float* r, int h, int w, int hr, int vr, const int* so, int dir, float* out) {
	auto L = view2(l, h, w);
	auto R = view2(r, h, w);
	searchOffset<2> range(so[0], so[1], so[2], so[3]);
	MD::Array<float, 4> cv = dir == int(dispDirection::RightToLeft)
		? unfoldBased2dDisparityCostVolume<F, float, float, 2, dispDirection::RightToLeft, float>(L, R, uint8_t(hr), uint8_t(vr), range)
		: unfoldBased2dDisparityCostVolume<F, float, float, 2, dispDirection::LeftToRight, float>(L, R, uint8_t(hr), uint8_t(vr), range);
	if (cv.empty()) return 1;
	auto s = cv.shape();
	for (int i = 0; i < s[0]; i++) for (int j = 0; j < s[1]; j++) for (int a = 0; a < s[2]; a++) for (int b = 0; b < s[3]; b++)
		out[((size_t(i) * s[1] + j) * s[2] + a) * s[3] + b] = cv.valueUnchecked(i, j, a, b);
	return 0;
}

template<dispExtractionStartegy S, typename T>
int sgm_n(const T* cv, int h, int w, int d, int ndir, float P1, float P2, float Pout, const int* m, float* out) {
	auto CV = view3(cv, h, w, d);
	Margins mg(m[0], m[1], m[2], m[3]);
	if (ndir == 4) out3(sgmCostVolume<4, S>(CV, P1, P2, mg, Pout), out);
	else if (ndir == 8) out3(sgmCostVolume<8, S>(CV, P1, P2, mg, Pout), out);
	else if (ndir == 16) out3(sgmCostVolume<16, S>(CV, P1, P2, mg, Pout), out);
	else return 2;
	return 0;
}

template<typename T>
int wta_t(const T* v, int h, int w, int d, int strategy, int32_t* out) {
	auto V = view3(v, h, w, d);
	if (strategy == 0) out2(extractSelectedIndex<dispExtractionStartegy::Cost>(V), out);
	else out2(extractSelectedIndex<dispExtractionStartegy::Score>(V), out);
	return 0;
}

template<matchingFunctions F, int depth>
int hier_d(const float* l, const float* r, int h, int w, int hr, int vr, int D, int radius, int dir,
		   float* tcv, int32_t* disp) {
	auto L = view2(l, h, w);
	auto R = view2(r, h, w);
	if (dir == int(dispDirection::RightToLeft)) {
		auto o = hiearchicalTruncatedCostVolume<F, depth, float, float, 2, dispDirection::RightToLeft, float>(L, R, hr, vr, D, radius);
		if (o.truncated_cost_volume.empty()) return 1;
		out3(o.truncated_cost_volume, tcv); out2(o.disp_estimate, disp);
	} else {
		auto o = hiearchicalTruncatedCostVolume<F, depth, float, float, 2, dispDirection::LeftToRight, float>(L, R, hr, vr, D, radius);
		if (o.truncated_cost_volume.empty()) return 1;
		out3(o.truncated_cost_volume, tcv); out2(o.disp_estimate, disp);
	}
	return 0;
}

} // namespace

template<matchingFunctions F>
int agg_dir(const float* l, const float* r, int h, int wl, int wr, int f, int lower, int upper, int dir, float* out) {
	auto L = view3(l, h, wl, f);
	auto R = view3(r, h, wr, f);
	searchOffset<1> range(lower, upper);
	MD::Array<float, 3> cv = dir == int(dispDirection::RightToLeft)
		? aggregateCost<F, float, float, dispDirection::RightToLeft, float>(L, R, range)
		: aggregateCost<F, float, float, dispDirection::LeftToRight, float>(L, R, range);
	if (cv.empty()) return 1;
	out3(cv, out);
	return 0;
}

template<matchingFunctions F>
int fv2cv_dir(const float* l, const float* r, int h, int wl, int wr, int f, int D, int dir, float* out) {
	auto L = view3(l, h, wl, f);
	auto R = view3(r, h, wr, f);
	MD::Array<float, 3> cv = dir == int(dispDirection::RightToLeft)
		? featureVolume2CostVolume<F, float, float, disp_t, dispDirection::RightToLeft, float>(L, R, disp_t(D))
		: featureVolume2CostVolume<F, float, float, disp_t, dispDirection::LeftToRight, float>(L, R, disp_t(D));
	if (cv.empty()) return 1;
	out3(cv, out);
	return 0;
}

template<matchingFunctions F, int ND>
int cv_comp_dir(const int* mask, int mh, int mw, const float* l, const float* r, int h, int wl, int wr, int c, int D, int dir, float* out) {
	UnFoldCompressor comp(view2(mask, mh, mw));
	MD::Array<float, 3> cv;
	if constexpr (ND == 2) {
		auto L = view2(l, h, wl); auto R = view2(r, h, wr);
		cv = dir == int(dispDirection::RightToLeft)
			? unfoldBasedCostVolume<F, float, float, 2, dispDirection::RightToLeft, float>(L, R, comp, D)
			: unfoldBasedCostVolume<F, float, float, 2, dispDirection::LeftToRight, float>(L, R, comp, D);
	} else {
		auto L = view3(l, h, wl, c); auto R = view3(r, h, wr, c);
		cv = dir == int(dispDirection::RightToLeft)
			? unfoldBasedCostVolume<F, float, float, 3, dispDirection::RightToLeft, float>(L, R, comp, D)
			: unfoldBasedCostVolume<F, float, float, 3, dispDirection::LeftToRight, float>(L, R, comp, D);
	}
	if (cv.empty()) return 1;
	out3(cv, out);
	return 0;
}

template<matchingFunctions F>
int cv3_dir(const float* l, const float* r, int h, int wl, int wr, int c, int hr, int vr, int D, int dir, float* out) {
	auto L = view3(l, h, wl, c); auto R = view3(r, h, wr, c);
	MD::Array<float, 3> cv = dir == int(dispDirection::RightToLeft)
		? unfoldBasedCostVolume<F, float, float, 3, dispDirection::RightToLeft, float>(L, R, uint8_t(hr), uint8_t(vr), D)
		: unfoldBasedCostVolume<F, float, float, 3, dispDirection::LeftToRight, float>(L, R, uint8_t(hr), uint8_t(vr), D);
	if (cv.empty()) return 1;
	out3(cv, out);
	return 0;
}

extern "C" {

int ref_num_threads() { return omp_get_max_threads(); }
void ref_set_num_threads(int n) { omp_set_num_threads(n); }

// unfold.h:248-291 (2-D image, auto padding). out is (h, w, F) C-order, F=(2hr+1)(2vr+1).
int ref_unfold_f32(const float* img, int h, int w, int hr, int vr, float* out) {
	out3(unfold<float, float>(uint8_t(hr), uint8_t(vr), view2(img, h, w)), out); return 0;
}
int ref_unfold_u8(const uint8_t* img, int h, int w, int hr, int vr, uint8_t* out) {
	out3(unfold<uint8_t, uint8_t>(uint8_t(hr), uint8_t(vr), view2(img, h, w)), out); return 0;
}

// unfold.h:248-344, any padding / orientation, 2-D (c == 0) or 3-D image. pad = {left, top, right, bottom}, pad[0] < 0 = auto.
// out (h', w', F*max(c,1)) C-order; returns the output height in *oh, width in *ow.
int ref_unfold_general_f32(const float* img, int h, int w, int c, int hr, int vr, const int* pad, int orientation, float* out, int* oh, int* ow) {
	PaddingMargins pm = pad[0] < 0 ? PaddingMargins() : PaddingMargins(pad[0], pad[1], pad[2], pad[3]);
	MD::Array<float, 3> r = c == 0 ? unfold<float, float>(uint8_t(hr), uint8_t(vr), view2(img, h, w), pm, UnfoldPatchOrientation(orientation))
								   : unfold<float, float>(uint8_t(hr), uint8_t(vr), view3(img, h, w, c), pm, UnfoldPatchOrientation(orientation));
	*oh = r.shape()[0]; *ow = r.shape()[1];
	if (out) out3(r, out);
	return 0;
}

// unfold.h:36-137 + :346-471: compressed unfold with UnFoldCompressor(mask). out (h', w', nFeatures*max(c,1)) float.
int ref_unfold_compressed_f32(const int* mask, int mh, int mw, const float* img, int h, int w, int c, const int* pad, float* out, int* oh, int* ow, int* nf) {
	UnFoldCompressor comp(view2(mask, mh, mw));
	PaddingMargins pm = pad[0] < 0 ? PaddingMargins() : PaddingMargins(pad[0], pad[1], pad[2], pad[3]);
	MD::Array<float, 3> r = c == 0 ? unfold<float, float>(comp, view2(img, h, w), pm) : unfold<float, float>(comp, view3(img, h, w, c), pm);
	*oh = r.shape()[0]; *ow = r.shape()[1]; *nf = comp.nFeatures();
	if (out) out3(r, out);
	return 0;
}

// census.h:69-115 on a ready-made (h, w, F) float feature volume
int ref_census_features_f32(const float* fv, int h, int w, int F, uint32_t* out) {
	out3(censusFeatures(view3(fv, h, w, F)), out); return 0;
}

// census.h:117-131. out is (h, w, nwords) uint32 C-order, nwords = (F-2)/32+1... i.e. (F-1-1)/32+1 per census.h:79.
int ref_census_u8(const uint8_t* img, int h, int w, int hr, int vr, uint32_t* out) {
	out3(censusTransform2D(view2(img, h, w), int8_t(hr), int8_t(vr)), out); return 0;
}
int ref_census_f32(const float* img, int h, int w, int hr, int vr, uint32_t* out) {
	out3(censusTransform2D(view2(img, h, w), int8_t(hr), int8_t(vr)), out); return 0;
}

// cross_correlations.h:645-722 on the unfolded volume (float images). out (h, w, F) float.
int ref_feature_volume_f32(int matchfunc, const float* img, int h, int w, int hr, int vr, float* out) {
	auto fv = unfold<float, float>(uint8_t(hr), uint8_t(vr), view2(img, h, w));
	switch (matchingFunctions(matchfunc)) {
	case matchingFunctions::ZNCC: out3(getFeatureVolumeForMatchFunc<matchingFunctions::ZNCC>(fv), out); break;
	case matchingFunctions::NCC: out3(getFeatureVolumeForMatchFunc<matchingFunctions::NCC>(fv), out); break;
	case matchingFunctions::ZCC: out3(getFeatureVolumeForMatchFunc<matchingFunctions::ZCC>(fv), out); break;
	case matchingFunctions::ZSSD: out3(getFeatureVolumeForMatchFunc<matchingFunctions::ZSSD>(fv), out); break;
	case matchingFunctions::ZSAD: out3(getFeatureVolumeForMatchFunc<matchingFunctions::ZSAD>(fv), out); break;
	case matchingFunctions::CC: out3(getFeatureVolumeForMatchFunc<matchingFunctions::CC>(fv), out); break;
	case matchingFunctions::SSD: out3(getFeatureVolumeForMatchFunc<matchingFunctions::SSD>(fv), out); break;
	case matchingFunctions::SAD: out3(getFeatureVolumeForMatchFunc<matchingFunctions::SAD>(fv), out); break;
	default: return 2;
	}
	return 0;
}

// cross_correlations.h:740-765, float images, float cost volume. out (h, w, D) C-order.
// The left image may be wider/narrower than the right one (only heights must match, :751).
int ref_cost_volume_f32(int matchfunc, const float* l, const float* r, int h, int w, int wr,
						int hr, int vr, int D, int dir, float* out) {
	switch (matchingFunctions(matchfunc)) {
	case matchingFunctions::CC: return cv_dir<matchingFunctions::CC>(l, r, h, w, wr, hr, vr, D, dir, out);
	case matchingFunctions::NCC: return cv_dir<matchingFunctions::NCC>(l, r, h, w, wr, hr, vr, D, dir, out);
	case matchingFunctions::SSD: return cv_dir<matchingFunctions::SSD>(l, r, h, w, wr, hr, vr, D, dir, out);
	case matchingFunctions::SAD: return cv_dir<matchingFunctions::SAD>(l, r, h, w, wr, hr, vr, D, dir, out);
	case matchingFunctions::ZCC: return cv_dir<matchingFunctions::ZCC>(l, r, h, w, wr, hr, vr, D, dir, out);
	case matchingFunctions::ZNCC: return cv_dir<matchingFunctions::ZNCC>(l, r, h, w, wr, hr, vr, D, dir, out);
	case matchingFunctions::ZSSD: return cv_dir<matchingFunctions::ZSSD>(l, r, h, w, wr, hr, vr, D, dir, out);
	case matchingFunctions::ZSAD: return cv_dir<matchingFunctions::ZSAD>(l, r, h, w, wr, hr, vr, D, dir, out);
	default: return 2;
	}
}

// the same with uint8 images through the reference's own uint8 instantiation (CC / SSD / SAD: the functions whose feature type
// for uint8 input is not the fixed-point one, cross_correlations.h:716-720)
int ref_cost_volume_u8(int matchfunc, const uint8_t* l, const uint8_t* r, int h, int w, int wr,
					   int hr, int vr, int D, int dir, float* out) {
	switch (matchingFunctions(matchfunc)) {
	case matchingFunctions::CC: return cv_dir<matchingFunctions::CC>(l, r, h, w, wr, hr, vr, D, dir, out);
	case matchingFunctions::SSD: return cv_dir<matchingFunctions::SSD>(l, r, h, w, wr, hr, vr, D, dir, out);
	case matchingFunctions::SAD: return cv_dir<matchingFunctions::SAD>(l, r, h, w, wr, hr, vr, D, dir, out);
	default: return 2;
	}
}

// unfold.h:476-691: the two super-pixel masks the reference ships for UnFoldCompressor. radius 3 -> 7x7, radius 4 -> 9x9.
int ref_grpix17_filter(int radius, int* out) {
	MD::Array<int, 2> m = radius == 3 ? CompressorGenerators::GrPix17R3Filter() : CompressorGenerators::GrPix17R4Filter();
	const int n = 2 * radius + 1;
	for (int i = 0; i < n; i++) for (int j = 0; j < n; j++) out[i * n + j] = m.value<MD::AccessCheck::Nocheck>(i, j);
	return 0;
}

// ---- feature-level forms: the reference templates on caller-supplied float feature volumes -----------------------------------
#define STEVI_MF_SWITCH(CALL) \
	switch (matchingFunctions(matchfunc)) { \
	case matchingFunctions::CC: CALL(matchingFunctions::CC); break; \
	case matchingFunctions::NCC: CALL(matchingFunctions::NCC); break; \
	case matchingFunctions::SSD: CALL(matchingFunctions::SSD); break; \
	case matchingFunctions::SAD: CALL(matchingFunctions::SAD); break; \
	case matchingFunctions::ZCC: CALL(matchingFunctions::ZCC); break; \
	case matchingFunctions::ZNCC: CALL(matchingFunctions::ZNCC); break; \
	case matchingFunctions::ZSSD: CALL(matchingFunctions::ZSSD); break; \
	case matchingFunctions::ZSAD: CALL(matchingFunctions::ZSAD); break; \
	default: return 2; }

// cross_correlations.h:645-722
int ref_features_for_matchfunc_f32(int matchfunc, const float* fv, int h, int w, int F, float* out) {
	auto V = view3(fv, h, w, F);
#define CALL(MF) out3(getFeatureVolumeForMatchFunc<MF>(V), out)
	STEVI_MF_SWITCH(CALL)
#undef CALL
	return 0;
}

// cross_correlations.h:251-308 (searchOffset<1> overload; lower = 0, upper = disp_width - 1 is the disp_width overload :194-249)
int ref_aggregate_cost_f32(int matchfunc, const float* fl, const float* fr, int h, int wl, int wr, int F, int lower, int upper, int dir, float* out) {
#define CALL(MF) return agg_dir<MF>(fl, fr, h, wl, wr, F, lower, upper, dir, out)
	STEVI_MF_SWITCH(CALL)
#undef CALL
	return 0;
}

// cross_correlations.h:724-738
int ref_feature_volume_to_cost_volume_f32(int matchfunc, const float* fl, const float* fr, int h, int wl, int wr, int F, int D, int dir, float* out) {
#define CALL(MF) return fv2cv_dir<MF>(fl, fr, h, wl, wr, F, D, dir, out)
	STEVI_MF_SWITCH(CALL)
#undef CALL
	return 0;
}

// cross_correlations.h:767-791: the UnFoldCompressor overload of unfoldBasedCostVolume; c = 0: (h,w) images, else (h,w,c)
int ref_cost_volume_compressed_f32(int matchfunc, const int* mask, int mh, int mw, const float* l, const float* r, int h, int wl, int wr, int c,
								   int D, int dir, float* out) {
#define CALL(MF) return c == 0 ? cv_comp_dir<MF, 2>(mask, mh, mw, l, r, h, wl, wr, c, D, dir, out) : cv_comp_dir<MF, 3>(mask, mh, mw, l, r, h, wl, wr, c, D, dir, out)
	STEVI_MF_SWITCH(CALL)
#undef CALL
	return 0;
}

// cross_correlations.h:740-765 with nImDim = 3 (multi-channel images)
int ref_cost_volume_multichannel_f32(int matchfunc, const float* l, const float* r, int h, int wl, int wr, int c, int hr, int vr, int D, int dir, float* out) {
#define CALL(MF) return cv3_dir<MF>(l, r, h, wl, wr, c, hr, vr, D, dir, out)
	STEVI_MF_SWITCH(CALL)
#undef CALL
	return 0;
}

// cross_correlations.h:794-822: 2-D search cost volume, float images of equal size. so = {lower0, upper0, lower1, upper1}
// (searchOffset<2> ctor order, correlation_base.h:299-315). out (h, w, upper0-lower0+1, upper1-lower1+1) C-order.
int ref_cost_volume_2d_f32(int matchfunc, const float* l, const float* r, int h, int w, int hr, int vr, const int* so, int dir, float* out) {
	switch (matchingFunctions(matchfunc)) {
	case matchingFunctions::CC: return cv2d_dir<matchingFunctions::CC>(l, r, h, w, hr, vr, so, dir, out);
	case matchingFunctions::NCC: return cv2d_dir<matchingFunctions::NCC>(l, r, h, w, hr, vr, so, dir, out);
	case matchingFunctions::SSD: return cv2d_dir<matchingFunctions::SSD>(l, r, h, w, hr, vr, so, dir, out);
	case matchingFunctions::SAD: return cv2d_dir<matchingFunctions::SAD>(l, r, h, w, hr, vr, so, dir, out);
	case matchingFunctions::ZCC: return cv2d_dir<matchingFunctions::ZCC>(l, r, h, w, hr, vr, so, dir, out);
	case matchingFunctions::ZNCC: return cv2d_dir<matchingFunctions::ZNCC>(l, r, h, w, hr, vr, so, dir, out);
	case matchingFunctions::ZSSD: return cv2d_dir<matchingFunctions::ZSSD>(l, r, h, w, hr, vr, so, dir, out);
	case matchingFunctions::ZSAD: return cv2d_dir<matchingFunctions::ZSAD>(l, r, h, w, hr, vr, so, dir, out);
	default: return 2;
	}
}

// correlation_base.h:466-508 on a (h, w, d1, d2) C-order volume; out (h, w, 2)
int ref_wta_2d_f32(const float* v, int h, int w, int d1, int d2, int strategy, int32_t* out) {
	MD::Array<float, 4> V(const_cast<float*>(v), {h, w, d1, d2}, {w * d1 * d2, d1 * d2, d2, 1}, false);
	if (strategy == 0) out3(extractSelected2dIndex<dispExtractionStartegy::Cost>(V), out);
	else out3(extractSelected2dIndex<dispExtractionStartegy::Score>(V), out);
	return 0;
}

// correlation_base.h:534-555
int ref_index2d_to_disp(const int32_t* idx, int h, int w, const int* so, int32_t* out) {
	searchOffset<2> range(so[0], so[1], so[2], so[3]);
	out3(selected2dIndexToDisp<disp_t>(view3(idx, h, w, 2), range), out);
	return 0;
}

// census cost volume, hamming_cv_t (uint16) costs, as the C2/C5 configs use it.
int ref_census_cost_volume_u8(const uint8_t* l, const uint8_t* r, int h, int w, int wr,
							  int hr, int vr, int D, int dir, uint16_t* out) {
	return cv_dir<matchingFunctions::CENSUS, uint8_t, uint16_t>(l, r, h, w, wr, hr, vr, D, dir, out);
}
int ref_census_cost_volume_f32(const float* l, const float* r, int h, int w, int wr,
							   int hr, int vr, int D, int dir, uint16_t* out) {
	return cv_dir<matchingFunctions::CENSUS, float, uint16_t>(l, r, h, w, wr, hr, vr, D, dir, out);
}

// sgm.h:360-404. margins = {left, top, right, bottom}, the ctor order of utils/margins.h:52.
int ref_sgm_u16(const uint16_t* cv, int h, int w, int d, int ndir, int strategy,
				float P1, float P2, float Pout, const int* margins, float* out) {
	if (strategy == 0) return sgm_n<dispExtractionStartegy::Cost>(cv, h, w, d, ndir, P1, P2, Pout, margins, out);
	return sgm_n<dispExtractionStartegy::Score>(cv, h, w, d, ndir, P1, P2, Pout, margins, out);
}
int ref_sgm_f32(const float* cv, int h, int w, int d, int ndir, int strategy,
				float P1, float P2, float Pout, const int* margins, float* out) {
	if (strategy == 0) return sgm_n<dispExtractionStartegy::Cost>(cv, h, w, d, ndir, P1, P2, Pout, margins, out);
	return sgm_n<dispExtractionStartegy::Score>(cv, h, w, d, ndir, P1, P2, Pout, margins, out);
}

// correlation_base.h:427-464
int ref_wta_f32(const float* v, int h, int w, int d, int strategy, int32_t* out) { return wta_t(v, h, w, d, strategy, out); }
int ref_wta_u16(const uint16_t* v, int h, int w, int d, int strategy, int32_t* out) { return wta_t(v, h, w, d, strategy, out); }

// correlation_base.h:511-532
int ref_index_to_disp(const int32_t* idx, int h, int w, int dir, int offset, int32_t* out) {
	auto I = view2(idx, h, w);
	if (dir == int(dispDirection::RightToLeft)) out2(selectedIndexToDisp<disp_t, dispDirection::RightToLeft>(I, offset), out);
	else out2(selectedIndexToDisp<disp_t, dispDirection::LeftToRight>(I, offset), out);
	return 0;
}

// correlation_base.h:557-577
int ref_selected_cost_f32(const float* cv, const int32_t* idx, int h, int w, int d, float* out) {
	out2(selectedCost(view3(cv, h, w, d), view2(idx, h, w)), out); return 0;
}

// correlation_base.h:579-674; out (h, w, 2*radius+1)
int ref_truncated_cost_volume_f32(const float* cv, const int32_t* idx, int h, int w, int d,
								  int hr, int vr, int radius, int dir, int tdir, float* out) {
	auto CV = view3(cv, h, w, d);
	auto I = view2(idx, h, w);
#define TCV_CASE(DD, TD) out3(truncatedCostVolume<float, dispDirection::DD, truncatedCostVolumeDirection::TD>(CV, I, hr, vr, radius), out)
	bool r2l = dir == int(dispDirection::RightToLeft);
	if (tdir == 0) { if (r2l) TCV_CASE(RightToLeft, Same); else TCV_CASE(LeftToRight, Same); }
	else if (tdir == 1) { if (r2l) TCV_CASE(RightToLeft, Reversed); else TCV_CASE(LeftToRight, Reversed); }
	else { if (r2l) TCV_CASE(RightToLeft, Both); else TCV_CASE(LeftToRight, Both); }
#undef TCV_CASE
	return 0;
}

// correlation_base.h:727-822; out (h, w, 2*radius+1 | 4*radius+1) as float 1/0
int ref_in_bound_domain_f32(const int32_t* idx, int h, int w, int width, int hr, int vr, int radius, int dir, int tdir, float* out) {
	auto I = view2(idx, h, w);
#define IB_CASE(DD, TD) out3(extractInBoundDomain<float, dispDirection::DD, truncatedCostVolumeDirection::TD>(I, uint32_t(width), hr, vr, radius), out)
	bool r2l = dir == int(dispDirection::RightToLeft);
	if (tdir == 0) { if (r2l) IB_CASE(RightToLeft, Same); else IB_CASE(LeftToRight, Same); }
	else if (tdir == 1) { if (r2l) IB_CASE(RightToLeft, Reversed); else IB_CASE(LeftToRight, Reversed); }
	else { if (r2l) IB_CASE(RightToLeft, Both); else IB_CASE(LeftToRight, Both); }
#undef IB_CASE
	return 0;
}

// cost_based_refinement.h:128-163; returns 1 (and leaves out untouched) when the reference returns an empty array
int ref_refine_disp_cost_interp(const float* tcv, const int32_t* idx, int h, int w, int t, int kernel, float* out) {
	auto T = view3(tcv, h, w, t);
	auto I = view2(idx, h, w);
	MD::Array<float, 2> r;
	if (kernel == int(InterpolationKernel::Equiangular)) r = refineDispCostInterpolation<InterpolationKernel::Equiangular>(T, I);
	else if (kernel == int(InterpolationKernel::Parabola)) r = refineDispCostInterpolation<InterpolationKernel::Parabola>(T, I);
	else r = refineDispCostInterpolation<InterpolationKernel::Gaussian>(T, I);
	if (r.empty()) return 1;
	out2(r, out);
	return 0;
}

// correlation_base.h:677-725 truncatedBidirectionaCostVolume; cv (h, w, d1, d2) C-order, idx (h, w, 2); out (h, w, 2*r0+1, 2*r1+1)
// with r0 = max(radius0, 1) and r1 = radius1 < 1 ? r0 : radius1 (the uint8_t arguments of the reference)
int ref_truncated_cost_volume_2d_f32(const float* cv, const int32_t* idx, int h, int w, int d1, int d2, int radius0, int radius1, float* out) {
	MD::Array<float, 4> CV(const_cast<float*>(cv), {h, w, d1, d2}, {w * d1 * d2, d1 * d2, d2, 1}, false);
	MD::Array<float, 4> t = truncatedBidirectionaCostVolume<float>(CV, view3(idx, h, w, 2), uint8_t(radius0), uint8_t(radius1));
	auto s = t.shape();
	for (int i = 0; i < s[0]; i++) for (int j = 0; j < s[1]; j++) for (int a = 0; a < s[2]; a++) for (int b = 0; b < s[3]; b++)
		out[((size_t(i) * s[1] + j) * s[2] + a) * s[3] + b] = t.valueUnchecked(i, j, a, b);
	return 0;
}

// cost_based_refinement.h:165-369 refineDisp2dCostInterpolation<kernel, isotropy>; tcv (h, w, n0, n1) C-order, raw (h, w, 2);
// returns 1 (out untouched) when the reference returns an empty array
int ref_refine_disp_2d_cost_interp(const float* tcv, const int32_t* raw, int h, int w, int n0, int n1, int kernel, int anisotropic, float* out) {
	MD::Array<float, 4> T(const_cast<float*>(tcv), {h, w, n0, n1}, {w * n0 * n1, n0 * n1, n1, 1}, false);
	auto I = view3(raw, h, w, 2);
	MD::Array<float, 3> r;
#define R2D_CASE(K) r = anisotropic ? refineDisp2dCostInterpolation<InterpolationKernel::K, IsotropyHypothesis::Anisotropic>(T, I) \
							   : refineDisp2dCostInterpolation<InterpolationKernel::K, IsotropyHypothesis::Isotropic>(T, I)
	if (kernel == int(InterpolationKernel::Equiangular)) R2D_CASE(Equiangular);
	else if (kernel == int(InterpolationKernel::Parabola)) R2D_CASE(Parabola);
	else R2D_CASE(Gaussian);
#undef R2D_CASE
	if (r.empty()) return 1;
	out3(r, out);
	return 0;
}

// guided_cost_filtering.h:102-176 variableCostVolumeAggregation<float, float>; cv (h, w, d), windows (hw, ww, n0, n1), out (h, w, d),
// all C-order; returns 1 (out untouched) when the reference returns an empty array
int ref_variable_cost_volume_aggregation_f32(const float* cv, int h, int w, int d, const float* win, int hw, int ww, int n0, int n1, float* out) {
	MD::Array<float, 4> Wn(const_cast<float*>(win), {hw, ww, n0, n1}, {ww * n0 * n1, n0 * n1, n1, 1}, false);
	MD::Array<float, 3> r = variableCostVolumeAggregation<float, float>(view3(cv, h, w, d), Wn);
	if (r.empty()) return 1;
	out3(r, out);
	return 0;
}

// interpolation/downsampling.h:67-118; out is (ceil(h/2), ceil(w/2))
int ref_avgpool2_f32(const float* img, int h, int w, float* out) {
	out2(Interpolation::averagePoolingDownsample<float, float>(view2(img, h, w), Interpolation::DownSampleWindows(2)), out);
	return 0;
}

// hierarchical.h:297-314; tcv (h, w, 2*radius+1) float, disp (h, w) int32
int ref_hierarchical_f32(int matchfunc, int depth, const float* l, const float* r, int h, int w,
						 int hr, int vr, int D, int radius, int dir, float* tcv, int32_t* disp) {
#define HIER(F) \
	if (depth == 1) return hier_d<matchingFunctions::F, 1>(l, r, h, w, hr, vr, D, radius, dir, tcv, disp); \
	if (depth == 2) return hier_d<matchingFunctions::F, 2>(l, r, h, w, hr, vr, D, radius, dir, tcv, disp); \
	if (depth == 3) return hier_d<matchingFunctions::F, 3>(l, r, h, w, hr, vr, D, radius, dir, tcv, disp); \
	return 2;
	switch (matchingFunctions(matchfunc)) {
	case matchingFunctions::ZNCC: HIER(ZNCC)
	case matchingFunctions::NCC: HIER(NCC)
	case matchingFunctions::SSD: HIER(SSD)
	case matchingFunctions::SAD: HIER(SAD)
	case matchingFunctions::ZSSD: HIER(ZSSD)
	case matchingFunctions::ZSAD: HIER(ZSAD)
	default: return 2;
	}
#undef HIER
}

// correlation_base.h:881-1004 meanFilter2D, :1006-1077 sigmaFilter; in (h,w) when c == 0, else (h,w,c). Only the interior
// [vr, h-vr) x [hr, w-hr) of the reference's result is ever written; the wrapper copies that part and leaves the rest of out alone.
int ref_mean_filter_2d_f32(const float* in, int h, int w, int c, int hr, int vr, float* out) {
	return c > 0 ? interior(meanFilter2D<float>(hr, vr, view3(in, h, w, c)), hr, vr, out) : interior(meanFilter2D<float>(hr, vr, view2(in, h, w)), hr, vr, out);
}
int ref_mean_filter_2d_u8(const uint8_t* in, int h, int w, int c, int hr, int vr, float* out) {
	return c > 0 ? interior(meanFilter2D<uint8_t>(hr, vr, view3(in, h, w, c)), hr, vr, out) : interior(meanFilter2D<uint8_t>(hr, vr, view2(in, h, w)), hr, vr, out);
}
int ref_sigma_filter_f32(const float* mean, const float* in, int h, int w, int c, int hr, int vr, float* out) {
	auto m = view2(mean, h, w);
	return c > 0 ? interior(sigmaFilter<float>(hr, vr, m, view3(in, h, w, c)), hr, vr, out) : interior(sigmaFilter<float>(hr, vr, m, view2(in, h, w)), hr, vr, out);
}
int ref_sigma_filter_u8(const float* mean, const uint8_t* in, int h, int w, int c, int hr, int vr, float* out) {
	auto m = view2(mean, h, w);
	return c > 0 ? interior(sigmaFilter<uint8_t>(hr, vr, m, view3(in, h, w, c)), hr, vr, out) : interior(sigmaFilter<uint8_t>(hr, vr, m, view2(in, h, w)), hr, vr, out);
}

} // extern "C"
