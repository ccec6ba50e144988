// TEST INFRASTRUCTURE — not product code. Only tests/ may load the library built from this file.
//
// extern "C" wrappers over the UNMODIFIED reference header correlation/image_based_refinement.h (included from $STEVI_REF,
// never copied into this repo), built by oracle/Makefile into oracle/_ref/libstevi_ref_ibr.so (its own translation unit so
// that it compiles beside ref_wrap.cpp). The reference's code runs as it is — its border rules, its comparisons, its float
// loops; the Eigen calls inside it (small vectors, FullPivHouseholderQR::solve) are served by the stand-in of
// oracle/shims/Eigen (+ oracle/small_qr.h): Eigen itself is not in this image. All buffers are dense C-order.
#include <cstdint>
#include <cstring>
#include <cmath>
#include <cassert>
#include <omp.h>

#include <correlation/image_based_refinement.h>
#include <io/image_io.h>
#include <correlation/patchmatch.h>
#include <correlation/on_demand_cost_volume.h>
#include <correlation/hierarchical.h>
#include <correlation/cost_based_refinement.h>

using namespace StereoVision;
using namespace StereoVision::Correlation;
namespace MD = Multidim;

namespace {

template<typename T> MD::Array<T, 2> view2(const T* p, int h, int w) { return MD::Array<T, 2>(const_cast<T*>(p), {h, w}, {w, 1}, false); }
template<typename T> MD::Array<T, 3> view3(const T* p, int h, int w, int c) { return MD::Array<T, 3>(const_cast<T*>(p), {h, w, c}, {w * c, c, 1}, false); }
void out2(MD::Array<float, 2> const& a, float* out) {
	auto s = a.shape();
	for (int i = 0; i < s[0]; i++) for (int j = 0; j < s[1]; j++) out[size_t(i) * s[1] + j] = a.valueUnchecked(i, j);
}
void out3(MD::Array<float, 3> const& a, float* out) {
	auto s = a.shape();
	for (int i = 0; i < s[0]; i++) for (int j = 0; j < s[1]; j++) for (int k = 0; k < s[2]; k++)
		out[(size_t(i) * s[1] + j) * s[2] + k] = a.valueUnchecked(i, j, k);
}

// the 1-column overload of leastAbsoluteDifferences (l1optimization.h:257-283) sorts break points that can be NaN (0/0);
// the wrapper runs it single-threaded like everything else here so a result is at least repeatable
template<matchingFunctions F, dispDirection D>
int bary1d(const float* fl, const float* fr, int h, int wl, int wr, int f, const int32_t* idx, int wi, float* out) {
	auto r = refineBarycentricDisp<F, D>(view3(fl, h, wl, f), view3(fr, h, wr, f), view2(idx, h, wi));
	out2(r, out);
	return 0;
}
template<matchingFunctions F, int R, dispDirection D>
int barysym1d(const float* fl, const float* fr, int h, int wl, int wr, int f, const int32_t* idx, int wi, int width, float* out) {
	auto r = refineBarycentricSymmetricDisp<F, R, D>(view3(fl, h, wl, f), view3(fr, h, wr, f), view2(idx, h, wi), width);
	out2(r, out);
	return 0;
}
template<matchingFunctions F, Contiguity::bidimensionalContiguity C, dispDirection D>
int bary2d(const float* fl, const float* fr, int h, int w, int f, const int32_t* idx, const int* so, int sym, float* out) {
	searchOffset<2> win(so[0], so[1], so[2], so[3]);
	MD::Array<int32_t, 3> sel(const_cast<int32_t*>(idx), {h, w, 2}, {w * 2, 2, 1}, false);
	auto r = sym ? refineBarycentricSymmetric2dDisp<F, C, D>(view3(fl, h, w, f), view3(fr, h, w, f), sel, win)
				 : refineBarycentric2dDisp<F, C, D>(view3(fl, h, w, f), view3(fr, h, w, f), sel, win);
	out3(r, out);
	return 0;
}
template<matchingFunctions F, dispDirection D>
int costsym(const float* fl, const float* fr, int h, int wl, int wr, int f, const int32_t* idx, const float* cv, int wi, int d, float* out) {
	auto r = refineCostSymmetricDisp<F, D>(view3(fl, h, wl, f), view3(fr, h, wr, f), view2(idx, h, wi), view3(cv, h, wi, d));
	out2(r, out);
	return 0;
}

// patchMatch<matchFunc, dims>(source, target, searchOffset, nIter, nRandomSearch, initializer) (patchmatch.h:498-556). init
// (nullable) is returned by the initializer; without it the reference draws the initial map (and, with nRand > 0, the search
// candidates) from std::random_device: such runs are not repeatable.
template<matchingFunctions F, int DIMS>
int pm_run(const float* fs, const float* ft, int h, int w, int ht, int wt, int f, const int* so, int nIter, int nRand, const int32_t* init, int32_t* out) {
	auto s = view3(fs, h, w, f), t = view3(ft, ht, wt, f);
	typedef MD::Array<float, 3> FV;
	std::optional<typename PatchMatchTraits<F, FV, FV, DIMS>::PatchMatchInitializer> initializer = std::nullopt;
	if (init) initializer = [init, h, w](FV const&, FV const&) {
		MD::Array<disp_t, 3> d(h, w, DIMS);
		for (int i = 0; i < h; i++) for (int j = 0; j < w; j++) for (int k = 0; k < DIMS; k++) d.atUnchecked(i, j, k) = init[(size_t(i) * w + j) * DIMS + k];
		return d;
	};
	MD::Array<disp_t, 3> d;
	if constexpr (DIMS == 1) d = patchMatch<F, 1>(s, t, searchOffset<1>(so[0], so[1]), nIter, nRand, initializer);
	else d = patchMatch<F, 2>(s, t, searchOffset<2>(so[0], so[1], so[2], so[3]), nIter, nRand, initializer);
	if (d.empty()) return 1;
	for (int i = 0; i < h; i++) for (int j = 0; j < w; j++) for (int k = 0; k < DIMS; k++) out[(size_t(i) * w + j) * DIMS + k] = d.valueUnchecked(i, j, k);
	return 0;
}

// OnDemandStereoCostVolume / OnDemandImageFlowVolume ::truncatedCostVolume(disp, radius) (on_demand_cost_volume.h:187-310)
template<matchingFunctions F>
int od_run(int dims, const float* fs, const float* ft, int h, int w, int ht, int wt, int f, const int* so, const int32_t* disp, int radius, float* out) {
	typedef MD::Array<float, 3> FV;
	auto s = view3(fs, h, w, f), t = view3(ft, ht, wt, f);
	const int n = 2 * radius + 1;
	if (dims == 1) {
		typedef OnDemandStereoCostVolume<F, float, FV, FV> CV;
		typename CV::SearchSpaceType space(SearchSpaceBase::IgnoredDim(), SearchSpaceBase::SearchDim(so[0], so[1]), SearchSpaceBase::FeatureDim());
		CV cv(s, t, space);
		auto r = cv.truncatedCostVolume(view2(disp, h, w), radius);
		if (r.empty()) return 1;
		for (int i = 0; i < h; i++) for (int j = 0; j < w; j++) for (int k = 0; k < n; k++) out[(size_t(i) * w + j) * n + k] = r.valueUnchecked(i, j, k);
	} else {
		typedef OnDemandImageFlowVolume<F, float, FV, FV> CV;
		typename CV::SearchSpaceType space(SearchSpaceBase::SearchDim(so[0], so[1]), SearchSpaceBase::SearchDim(so[2], so[3]), SearchSpaceBase::FeatureDim());
		CV cv(s, t, space);
		MD::Array<int32_t, 3> d(const_cast<int32_t*>(disp), {h, w, 2}, {w * 2, 2, 1}, false);
		auto r = cv.truncatedCostVolume(d, radius);
		if (r.empty()) return 1;
		for (int i = 0; i < h; i++) for (int j = 0; j < w; j++) for (int a = 0; a < n; a++) for (int b = 0; b < n; b++)
			out[((size_t(i) * w + j) * n + a) * n + b] = r.valueUnchecked(i, j, a, b);
	}
	return 0;
}

} // namespace

extern "C" {

void ref_ibr_set_num_threads(int n) { omp_set_num_threads(n); }

#define IBR_MF(...) \
	switch (matchfunc) { \
	case 1: { constexpr auto F = matchingFunctions::NCC; __VA_ARGS__; } \
	case 5: { constexpr auto F = matchingFunctions::ZNCC; __VA_ARGS__; } \
	case 2: { constexpr auto F = matchingFunctions::SSD; __VA_ARGS__; } \
	case 6: { constexpr auto F = matchingFunctions::ZSSD; __VA_ARGS__; } \
	default: break; }
#define IBR_MF_L1(...) \
	switch (matchfunc) { \
	case 3: { constexpr auto F = matchingFunctions::SAD; __VA_ARGS__; } \
	case 7: { constexpr auto F = matchingFunctions::ZSAD; __VA_ARGS__; } \
	default: break; }

// image_based_refinement.h:184-355. matchfunc: the stevi_matchfunc numbering (c_api.h); dir 1 = RightToLeft
int ref_refine_barycentric_disp(int matchfunc, const float* fl, const float* fr, int h, int wl, int wr, int f, const int32_t* idx, int dir, float* out) {
	const int wi = dir ? wr : wl;
	if (dir) { IBR_MF(return (bary1d<F, dispDirection::RightToLeft>(fl, fr, h, wl, wr, f, idx, wi, out))) IBR_MF_L1(return (bary1d<F, dispDirection::RightToLeft>(fl, fr, h, wl, wr, f, idx, wi, out))) }
	else { IBR_MF(return (bary1d<F, dispDirection::LeftToRight>(fl, fr, h, wl, wr, f, idx, wi, out))) IBR_MF_L1(return (bary1d<F, dispDirection::LeftToRight>(fl, fr, h, wl, wr, f, idx, wi, out))) }
	return 3;
}

// image_based_refinement.h:31-105 (refineRadius 1 or 2)
int ref_refine_barycentric_symmetric_disp(int matchfunc, int radius, const float* fl, const float* fr, int h, int wl, int wr, int f,
										  const int32_t* idx, int width, int dir, float* out) {
	const int wi = dir ? wr : wl;
	if (radius == 1) {
		if (dir) { IBR_MF(return (barysym1d<F, 1, dispDirection::RightToLeft>(fl, fr, h, wl, wr, f, idx, wi, width, out))) }
		else { IBR_MF(return (barysym1d<F, 1, dispDirection::LeftToRight>(fl, fr, h, wl, wr, f, idx, wi, width, out))) }
	} else if (radius == 2) {
		if (dir) { IBR_MF(return (barysym1d<F, 2, dispDirection::RightToLeft>(fl, fr, h, wl, wr, f, idx, wi, width, out))) }
		else { IBR_MF(return (barysym1d<F, 2, dispDirection::LeftToRight>(fl, fr, h, wl, wr, f, idx, wi, width, out))) }
	}
	return 3;
}

// image_based_refinement.h:675-880 (sym 0) and :1068-1241 (sym 1). contiguity: 0 Queen, 1 Rook, 2 Bishop
int ref_refine_barycentric_2d_disp(int matchfunc, int contiguity, int sym, const float* fl, const float* fr, int h, int w, int f,
								   const int32_t* idx, const int* so, int dir, float* out) {
#define IBR_C(CONT) \
	if (dir) { IBR_MF(return (bary2d<F, CONT, dispDirection::RightToLeft>(fl, fr, h, w, f, idx, so, sym, out))) } \
	else { IBR_MF(return (bary2d<F, CONT, dispDirection::LeftToRight>(fl, fr, h, w, f, idx, so, sym, out))) }
	if (contiguity == 0) { IBR_C(Contiguity::Queen) }
	else if (contiguity == 1) { IBR_C(Contiguity::Rook) }
	else if (contiguity == 2) { IBR_C(Contiguity::Bishop) }
	return 3;
}

// image_based_refinement.h:1591-1690
int ref_refine_cost_symmetric_disp(int matchfunc, const float* fl, const float* fr, int h, int wl, int wr, int f, const int32_t* idx,
								   const float* cv, int d, int dir, float* out) {
	const int wi = dir ? wr : wl;
	if (dir) { IBR_MF(return (costsym<F, dispDirection::RightToLeft>(fl, fr, h, wl, wr, f, idx, cv, wi, d, out))) IBR_MF_L1(return (costsym<F, dispDirection::RightToLeft>(fl, fr, h, wl, wr, f, idx, cv, wi, d, out))) }
	else { IBR_MF(return (costsym<F, dispDirection::LeftToRight>(fl, fr, h, wl, wr, f, idx, cv, wi, d, out))) IBR_MF_L1(return (costsym<F, dispDirection::LeftToRight>(fl, fr, h, wl, wr, f, idx, cv, wi, d, out))) }
	return 3;
}

// io/image_io.h:66-165 through the reference's own writer / reader: data is a dense C-order (d0,d1,d2) block; `ref_layout` != 0
// writes it through an array with the reference's cost-volume strides {d1*d2, 1, d1} (cross_correlations.h:220)
int ref_write_stevimg_f32(const char* path, const float* data, int d0, int d1, int d2, int ref_layout) {
	MD::Array<float, 3> a = ref_layout ? MD::Array<float, 3>({d0, d1, d2}, {d1 * d2, 1, d1}) : MD::Array<float, 3>({d0, d1, d2}, {d1 * d2, d2, 1});
	for (int i = 0; i < d0; i++) for (int j = 0; j < d1; j++) for (int k = 0; k < d2; k++) a.atUnchecked(i, j, k) = data[(size_t(i) * d1 + j) * d2 + k];
	return IO::writeStevimg<float, float, 3>(path, a) ? 0 : 1;
}
int ref_write_stevimg_u8_from_i32(const char* path, const int32_t* data, int d0, int d1) {
	MD::Array<int32_t, 2> a({d0, d1}, {d1, 1});
	for (int i = 0; i < d0; i++) for (int j = 0; j < d1; j++) a.atUnchecked(i, j) = data[size_t(i) * d1 + j];
	return IO::writeStevimg<uint8_t, int32_t, 2>(path, a) ? 0 : 1;
}
int ref_read_stevimg_f32(const char* path, float* out, int* shape) {
	auto a = IO::readStevimg<float, 3>(path);
	if (a.empty()) return 1;
	for (int k = 0; k < 3; k++) shape[k] = a.shape()[k];
	out3(a, out);
	return 0;
}
int ref_stevimg_matches(const char* path, int which) {
	return which == 0 ? IO::stevImgFileMatchTypeAndDim<float, 3>(path) : (which == 1 ? IO::stevImgFileMatchTypeAndDim<uint8_t, 2>(path) : IO::stevImgFileMatchTypeAndDim<float, 2>(path));
}

int ref_patch_match(int matchfunc, int dims, const float* fs, const float* ft, int h, int w, int ht, int wt, int f, const int* so, int nIter, int nRand,
					const int32_t* init, int32_t* out) {
	switch (matchfunc) {
	case 5: return dims == 1 ? pm_run<matchingFunctions::ZNCC, 1>(fs, ft, h, w, ht, wt, f, so, nIter, nRand, init, out) : pm_run<matchingFunctions::ZNCC, 2>(fs, ft, h, w, ht, wt, f, so, nIter, nRand, init, out);
	case 2: return dims == 1 ? pm_run<matchingFunctions::SSD, 1>(fs, ft, h, w, ht, wt, f, so, nIter, nRand, init, out) : pm_run<matchingFunctions::SSD, 2>(fs, ft, h, w, ht, wt, f, so, nIter, nRand, init, out);
	case 7: return dims == 1 ? pm_run<matchingFunctions::ZSAD, 1>(fs, ft, h, w, ht, wt, f, so, nIter, nRand, init, out) : pm_run<matchingFunctions::ZSAD, 2>(fs, ft, h, w, ht, wt, f, so, nIter, nRand, init, out);
	default: return 3;
	}
}

int ref_on_demand_truncated_cost_volume(int matchfunc, int dims, const float* fs, const float* ft, int h, int w, int ht, int wt, int f, const int* so,
										const int32_t* disp, int radius, float* out) {
	switch (matchfunc) {
	case 5: return od_run<matchingFunctions::ZNCC>(dims, fs, ft, h, w, ht, wt, f, so, disp, radius, out);
	case 2: return od_run<matchingFunctions::SSD>(dims, fs, ft, h, w, ht, wt, f, so, disp, radius, out);
	case 7: return od_run<matchingFunctions::ZSAD>(dims, fs, ft, h, w, ht, wt, f, so, disp, radius, out);
	default: return 3;
	}
}

// computeGuidedCV<matchFunc, float, float, dDir>(feature_vol_l, feature_vol_r, disp_guide, radius) (hierarchical.h:73-233)
int ref_guided_cv_features_f32(int matchfunc, const float* fl, const float* fr, int h, int wl, int wr, int f, const int32_t* guide, int hg, int wg,
							   int radius, int dir, float* tcv, int32_t* disp) {
	auto L = view3(fl, h, wl, f), R = view3(fr, h, wr, f);
	auto G = view2(guide, hg, wg);
	const int ws = dir ? wr : wl, n = 2 * radius + 1;
#define GCV(F_) { auto r = dir ? computeGuidedCV<F_, float, float, dispDirection::RightToLeft>(L, R, G, radius) : computeGuidedCV<F_, float, float, dispDirection::LeftToRight>(L, R, G, radius); \
		if (r.truncated_cost_volume.empty()) return 1; \
		for (int i = 0; i < h; i++) for (int j = 0; j < ws; j++) { disp[size_t(i) * ws + j] = r.disp_estimate.valueUnchecked(i, j); \
			for (int k = 0; k < n; k++) tcv[(size_t(i) * ws + j) * n + k] = r.truncated_cost_volume.valueUnchecked(i, j, k); } return 0; }
	switch (matchfunc) {
	case 5: GCV(matchingFunctions::ZNCC)
	case 2: GCV(matchingFunctions::SSD)
	case 7: GCV(matchingFunctions::ZSAD)
	default: return 3;
	}
#undef GCV
}

// refineDisp2dCostPatchInterpolation<kernel> (cost_based_refinement.h:371-433 over refineCostPatch :71-126): the reference's code; its
// two Eigen inverses (6x6 normal equations, 2x2 Hessian) run on the stand-in's Gauss-Jordan. kernel: 1 Parabola, 2 Gaussian
int ref_refine_disp_2d_cost_patch_interp(const float* tcv, const int32_t* raw, int h, int w, int n0, int n1, int kernel, float* out) {
	MD::Array<float, 4> t(const_cast<float*>(tcv), {h, w, n0, n1}, {w * n0 * n1, n0 * n1, n1, 1}, false);
	MD::Array<int32_t, 3> d(const_cast<int32_t*>(raw), {h, w, 2}, {w * 2, 2, 1}, false);
	auto r = kernel == 2 ? refineDisp2dCostPatchInterpolation<InterpolationKernel::Gaussian>(t, d) : refineDisp2dCostPatchInterpolation<InterpolationKernel::Parabola>(t, d);
	if (r.empty()) return 1;
	out3(r, out);
	return 0;
}

// refinedBarycentricDisp on images (image_based_refinement.h:2502-2544): unfold, features, cost volume, winner, refinement
int ref_refined_barycentric_disp_f32(int matchfunc, const float* l, const float* r, int h, int wl, int wr, int hr, int vr, int width,
									 int prenormalize, float* out) {
	IBR_MF({ auto res = refinedBarycentricDisp<F, float, float, 2, dispDirection::RightToLeft>(view2(l, h, wl), view2(r, h, wr), uint8_t(hr), uint8_t(vr), width, prenormalize != 0);
			 out2(res, out); return 0; })
	return 3;
}

}
