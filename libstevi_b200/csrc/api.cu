// extern "C" entry points (include/stevi_b200/c_api.h): argument checks, staging of host operands, stage chaining.
#include "common.cuh"
#include <algorithm>
#include <vector>
#include "kernels.cuh"
#include <cmath>

using namespace stevi;

namespace {

int32_t check_images(stevi_ctx* c, const View& l, const View& r) {
	if (l.ndim != 2 || r.ndim != 2) return fail(c, STEVI_E_UNSUPPORTED, "images must be 2-D (H,W)");
	if (l.dtype != r.dtype || (l.dtype != STEVI_U8 && l.dtype != STEVI_F32))
		return fail(c, STEVI_E_DTYPE, "images must both be u8 or both f32");
	if (l.shape[0] != r.shape[0]) return fail(c, STEVI_E_SHAPE, "image heights differ (cross_correlations.h:751)");
	return STEVI_OK;
}

// dense (H,W,n) u32 workspace array with the word axis slowest ([n][H][W]): coalesced for every consumer
int32_t alloc_census(stevi_ctx* c, int H, int W, int nw, View& v) {
	uint32_t* p = nullptr;
	STEVI_TRY(arena_alloc_t(c, (size_t)H * W * nw, &p));
	v.p = p; v.dtype = STEVI_U32; v.ndim = 3;
	v.shape[0] = H; v.shape[1] = W; v.shape[2] = nw;
	v.st[0] = W; v.st[1] = 1; v.st[2] = (int64_t)H * W;
	return STEVI_OK;
}

// census cost volume on device views; optionally emits the g map of the collapsed SGM
int32_t census_cv_dev(stevi_ctx* c, const View& l, const View& r, int hr, int vr, int D, int direction, int semantics,
					  const View& out, int32_t* gmap, int pout_i) {
	STEVI_TRY(check_images(c, l, r));
	const int nw = census_n_words(hr, vr);
	View cl, cr;
	STEVI_TRY(alloc_census(c, l.shape[0], l.shape[1], nw, cl));
	STEVI_TRY(alloc_census(c, r.shape[0], r.shape[1], nw, cr));
	const View imgs[2] = {l, r}, cens[2] = {cl, cr};
	STEVI_TRY(census_transform_pair_dev(c, imgs, cens, 2, hr, vr, semantics));
	// words past n_written are zero on both sides (and round to zero): no need to compare them
	return hamming_cv_dev(c, cl, cr, census_n_written(hr, vr, semantics), D, direction, semantics == STEVI_COMPAT, out, gmap, pout_i);
}

} // namespace

extern "C" {

int32_t stevi_census_transform(stevi_ctx* c, const stevi_array* img, int32_t hr, int32_t vr, int32_t semantics, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vi, vo;
	STEVI_TRY(st.in(img, vi, "img"));
	STEVI_TRY(st.out(out, vo, "out"));
	STEVI_TRY(census_transform_dev(c, vi, hr, vr, semantics, vo));
	return st.finish();
}

int32_t stevi_unfold(stevi_ctx* c, const stevi_array* img, int32_t hr, int32_t vr, const int32_t* padding, int32_t orientation, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vi, vo;
	STEVI_TRY(st.in(img, vi, "img"));
	STEVI_TRY(st.out(out, vo, "out"));
	STEVI_TRY(unfold_dev(c, vi, hr, vr, padding, orientation, vo));
	return st.finish();
}

int32_t stevi_unfold_compressed(stevi_ctx* c, const stevi_array* mask, const stevi_array* img, const int32_t* padding, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	if (!mask || !mask->data || mask->mem != STEVI_HOST || mask->ndim != 2 || mask->dtype != STEVI_I32)
		return fail(c, STEVI_E_ARG, "unfold_compressed: mask must be a 2-D i32 array in host memory");
	Stager st(c);
	View vi, vo;
	STEVI_TRY(st.in(img, vi, "img"));
	STEVI_TRY(st.out(out, vo, "out"));
	STEVI_TRY(unfold_compressed_dev(c, static_cast<const int32_t*>(mask->data), mask->shape[0], mask->shape[1], mask->strides[0],
									mask->strides[1], vi, padding, vo));
	return st.finish();
}

int32_t stevi_census_features(stevi_ctx* c, const stevi_array* feat, int32_t semantics, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vi, vo;
	STEVI_TRY(st.in(feat, vi, "feat"));
	STEVI_TRY(st.out(out, vo, "out"));
	STEVI_TRY(census_features_dev(c, vi, semantics & 1, vo));
	return st.finish();
}

int32_t stevi_feature_volume(stevi_ctx* c, int32_t mf, const stevi_array* img, int32_t hr, int32_t vr, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vi, vo;
	STEVI_TRY(st.in(img, vi, "img"));
	STEVI_TRY(st.out(out, vo, "out"));
	STEVI_TRY(feature_volume_dev(c, mf, vi, hr, vr, vo));
	return st.finish();
}

int32_t stevi_hamming_cost_volume(stevi_ctx* c, const stevi_array* fl, const stevi_array* fr, int32_t D, int32_t direction,
								  int32_t semantics, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vl, vr, vo;
	STEVI_TRY(st.in(fl, vl, "feat_l"));
	STEVI_TRY(st.in(fr, vr, "feat_r"));
	STEVI_TRY(st.out(out, vo, "out"));
	if (vl.ndim != 3) return fail(c, STEVI_E_SHAPE, "hamming: features must be 3-D");
	STEVI_TRY(hamming_cv_dev(c, vl, vr, vl.shape[2], D, direction, semantics == STEVI_COMPAT, vo, nullptr, 0));
	return st.finish();
}

int32_t stevi_features_for_matchfunc(stevi_ctx* c, int32_t mf, const stevi_array* feat, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vi, vo;
	STEVI_TRY(st.in(feat, vi, "feat"));
	STEVI_TRY(st.out(out, vo, "out"));
	STEVI_TRY(features_for_matchfunc_dev(c, mf, vi, vo));
	return st.finish();
}

int32_t stevi_aggregate_cost(stevi_ctx* c, int32_t mf, const stevi_array* fl, const stevi_array* fr, int32_t lower, int32_t upper,
							 int32_t direction, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vl, vr, vo;
	STEVI_TRY(st.in(fl, vl, "feat_l"));
	STEVI_TRY(st.in(fr, vr, "feat_r"));
	STEVI_TRY(st.out(out, vo, "out"));
	STEVI_TRY(aggregate_cost_dev(c, mf, vl, vr, lower, upper - lower + 1, direction, vo));
	return st.finish();
}

namespace stevi {
// a workspace (H,W,F) f32 volume in C order
static int32_t ws_volume(stevi_ctx* c, int H, int W, int F, int dtype, View& v) {
	void* p = nullptr;
	STEVI_TRY(arena_alloc(c, (size_t)H * W * F * dtype_size(dtype), &p));
	v.p = p; v.dtype = dtype; v.ndim = 3;
	v.shape[0] = H; v.shape[1] = W; v.shape[2] = F;
	v.st[0] = (int64_t)W * F; v.st[1] = F; v.st[2] = 1;
	return STEVI_OK;
}

// featureVolume2CostVolume (cross_correlations.h:724-738) on device views: features of both volumes, then aggregateCost
static int32_t feature_volumes_to_cv(stevi_ctx* c, int mf, const View& fl, const View& fr, int D, int direction, int semantics, const View& out) {
	if (fl.ndim != 3 || fr.ndim != 3) return fail(c, STEVI_E_SHAPE, "feature volumes must be (H,W,F)");
	if (fl.shape[0] != fr.shape[0] || fl.shape[2] != fr.shape[2])
		return fail(c, STEVI_E_SHAPE, "feature volumes disagree in height or feature count (cross_correlations.h:751-759)");
	if (mf == STEVI_MF_CENSUS) {   // :712-714: census bit strings, then the Hamming volume
		const int nw = (fl.shape[2] - 1) / 32 + 1;
		View cl, cr;
		STEVI_TRY(ws_volume(c, fl.shape[0], fl.shape[1], nw, STEVI_U32, cl));
		STEVI_TRY(ws_volume(c, fr.shape[0], fr.shape[1], nw, STEVI_U32, cr));
		STEVI_TRY(census_features_dev(c, fl, semantics & 1, cl));
		STEVI_TRY(census_features_dev(c, fr, semantics & 1, cr));
		return hamming_cv_dev(c, cl, cr, nw, D, direction, (semantics & 1) == STEVI_COMPAT, out, nullptr, 0);
	}
	View gl, gr;
	STEVI_TRY(ws_volume(c, fl.shape[0], fl.shape[1], fl.shape[2], STEVI_F32, gl));
	STEVI_TRY(ws_volume(c, fr.shape[0], fr.shape[1], fr.shape[2], STEVI_F32, gr));
	STEVI_TRY(features_for_matchfunc_dev(c, mf, fl, gl));
	STEVI_TRY(features_for_matchfunc_dev(c, mf, fr, gr));
	return aggregate_cost_dev(c, mf, gl, gr, 0, D, direction, out);
}
} // namespace stevi

int32_t stevi_feature_volume_to_cost_volume(stevi_ctx* c, int32_t mf, const stevi_array* fl, const stevi_array* fr, int32_t D,
											int32_t direction, int32_t semantics, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vl, vr, vo;
	STEVI_TRY(st.in(fl, vl, "feat_l"));
	STEVI_TRY(st.in(fr, vr, "feat_r"));
	STEVI_TRY(st.out(out, vo, "out"));
	if (vl.dtype != STEVI_F32 || vr.dtype != STEVI_F32) return fail(c, STEVI_E_DTYPE, "feature_volume_to_cost_volume: feature volumes must be f32");
	STEVI_TRY(feature_volumes_to_cv(c, mf, vl, vr, D, direction, semantics, vo));
	return st.finish();
}

int32_t stevi_cost_volume_compressed(stevi_ctx* c, int32_t mf, const stevi_array* mask, const stevi_array* l, const stevi_array* r,
									 int32_t D, int32_t direction, int32_t semantics, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	if (!mask || !mask->data || mask->mem != STEVI_HOST || mask->ndim != 2 || mask->dtype != STEVI_I32)
		return fail(c, STEVI_E_ARG, "cost_volume_compressed: mask must be a 2-D i32 array in host memory");
	Stager st(c);
	View vl, vr, vo;
	STEVI_TRY(st.in(l, vl, "img_l"));
	STEVI_TRY(st.in(r, vr, "img_r"));
	STEVI_TRY(st.out(out, vo, "out"));
	if (vl.ndim != vr.ndim || (vl.ndim != 2 && vl.ndim != 3)) return fail(c, STEVI_E_SHAPE, "cost_volume_compressed: images must both be (H,W) or (H,W,C)");
	// number of features = number of distinct positive labels (UnFoldCompressor, unfold.h:36-137)
	const int32_t* m = static_cast<const int32_t*>(mask->data);
	std::vector<int32_t> labels;
	for (int a = 0; a < mask->shape[0]; a++)
		for (int b = 0; b < mask->shape[1]; b++) {
			const int32_t v = m[a * mask->strides[0] + b * mask->strides[1]];
			if (v > 0 && std::find(labels.begin(), labels.end(), v) == labels.end()) labels.push_back(v);
		}
	const int C = vl.ndim == 3 ? vl.shape[2] : 1;
	if (vr.ndim == 3 && vr.shape[2] != C) return fail(c, STEVI_E_SHAPE, "cost_volume_compressed: channel counts differ (cross_correlations.h:775-777)");
	const int F = (int)labels.size() * C;
	View fl, fr;
	STEVI_TRY(ws_volume(c, vl.shape[0], vl.shape[1], F, STEVI_F32, fl));
	STEVI_TRY(ws_volume(c, vr.shape[0], vr.shape[1], F, STEVI_F32, fr));
	STEVI_TRY(unfold_compressed_dev(c, m, mask->shape[0], mask->shape[1], mask->strides[0], mask->strides[1], vl, nullptr, fl));
	STEVI_TRY(unfold_compressed_dev(c, m, mask->shape[0], mask->shape[1], mask->strides[0], mask->strides[1], vr, nullptr, fr));
	STEVI_TRY(feature_volumes_to_cv(c, mf, fl, fr, D, direction, semantics, vo));
	return st.finish();
}

int32_t stevi_cost_volume_multichannel(stevi_ctx* c, int32_t mf, const stevi_array* l, const stevi_array* r, int32_t hr, int32_t vr_,
									   int32_t D, int32_t direction, int32_t semantics, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vl, vr, vo;
	STEVI_TRY(st.in(l, vl, "img_l"));
	STEVI_TRY(st.in(r, vr, "img_r"));
	STEVI_TRY(st.out(out, vo, "out"));
	if (vl.ndim != 3 || vr.ndim != 3) return fail(c, STEVI_E_SHAPE, "cost_volume_multichannel: images must be (H,W,C)");
	if (vl.shape[2] != vr.shape[2]) return fail(c, STEVI_E_SHAPE, "cost_volume_multichannel: channel counts differ (cross_correlations.h:755-759)");
	if (hr < 0 || vr_ < 0 || hr > 127 || vr_ > 127) return fail(c, STEVI_E_ARG, "cost_volume_multichannel: radii");
	const int F = (2 * hr + 1) * (2 * vr_ + 1) * vl.shape[2];
	View fl, fr;
	STEVI_TRY(ws_volume(c, vl.shape[0], vl.shape[1], F, STEVI_F32, fl));
	STEVI_TRY(ws_volume(c, vr.shape[0], vr.shape[1], F, STEVI_F32, fr));
	STEVI_TRY(unfold_dev(c, vl, hr, vr_, nullptr, 0, fl));   // unfold.h:293-344 (3-D image overload), automatic padding
	STEVI_TRY(unfold_dev(c, vr, hr, vr_, nullptr, 0, fr));
	STEVI_TRY(feature_volumes_to_cv(c, mf, fl, fr, D, direction, semantics, vo));
	return st.finish();
}

int32_t stevi_cost_volume(stevi_ctx* c, int32_t mf, const stevi_array* l, const stevi_array* r, int32_t hr, int32_t vr,
						  int32_t D, int32_t direction, int32_t semantics, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vl, vr_, vo;
	STEVI_TRY(st.in(l, vl, "img_l"));
	STEVI_TRY(st.in(r, vr_, "img_r"));
	STEVI_TRY(st.out(out, vo, "out"));
	if (mf == STEVI_MF_CENSUS) {
		STEVI_TRY(census_cv_dev(c, vl, vr_, hr, vr, D, direction, semantics & 1, vo, nullptr, 0));
	} else {
		// semantics bit 8 (STEVI_EXACT_SUMS) asks for the reference's own O(F) per-cell sums instead of the box sums
		STEVI_TRY(float_cv_dev(c, mf, vl, vr_, hr, vr, D, direction, vo, (semantics & STEVI_EXACT_SUMS) != 0));
	}
	return st.finish();
}

int32_t stevi_sgm(stevi_ctx* c, const stevi_array* cv, int32_t ndir, int32_t strategy, float P1, float P2, float Pout,
				  const stevi_margins* margins, int32_t semantics, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vc, vo;
	STEVI_TRY(st.in(cv, vc, "cv"));
	STEVI_TRY(st.out(out, vo, "out"));
	stevi_margins mg = margins ? *margins : stevi_margins{0, 0, 0, 0};
	STEVI_TRY(sgm_dev(c, vc, ndir, strategy, P1, P2, Pout, mg, semantics, vo, nullptr, -1, nullptr));
	return st.finish();
}

int32_t stevi_wta(stevi_ctx* c, const stevi_array* vol, int32_t strategy, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vv, vo;
	STEVI_TRY(st.in(vol, vv, "vol"));
	STEVI_TRY(st.out_index(out, vo, vv.ndim == 3 ? vv.shape[2] : 0, 1, 0, "out"));
	STEVI_TRY(wta_dev(c, vv, strategy, vo));
	return st.finish();
}

int32_t stevi_index_to_disp(stevi_ctx* c, const stevi_array* idx, int32_t direction, int32_t offset, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vi, vo;
	STEVI_TRY(st.in(idx, vi, "idx"));
	STEVI_TRY(st.out(out, vo, "out"));
	STEVI_TRY(index_to_disp_dev(c, vi, direction, offset, vo));
	return st.finish();
}

int32_t stevi_cost_volume_2d_ex(stevi_ctx* c, int32_t mf, const stevi_array* l, const stevi_array* r, int32_t hr, int32_t vr,
								const int32_t* search, int32_t direction, int32_t semantics, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vl, vr_, vo;
	STEVI_TRY(st.in(l, vl, "img_l"));
	STEVI_TRY(st.in(r, vr_, "img_r"));
	STEVI_TRY(st.out(out, vo, "out"));
	STEVI_TRY(float_cv2d_dev(c, mf, vl, vr_, hr, vr, search, direction, vo, (semantics & STEVI_EXACT_SUMS) != 0));
	return st.finish();
}

int32_t stevi_cost_volume_2d(stevi_ctx* c, int32_t mf, const stevi_array* l, const stevi_array* r, int32_t hr, int32_t vr,
							 const int32_t* search, int32_t direction, stevi_array* out) {
	return stevi_cost_volume_2d_ex(c, mf, l, r, hr, vr, search, direction, STEVI_COMPAT, out);
}

int32_t stevi_wta_2d(stevi_ctx* c, const stevi_array* vol, int32_t strategy, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vv, vo;
	STEVI_TRY(st.in(vol, vv, "vol"));
	STEVI_TRY(st.out(out, vo, "out"));
	STEVI_TRY(wta2d_dev(c, vv, strategy, vo));
	return st.finish();
}

int32_t stevi_index2d_to_disp(stevi_ctx* c, const stevi_array* idx, const int32_t* search, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vi, vo;
	STEVI_TRY(st.in(idx, vi, "idx"));
	STEVI_TRY(st.out(out, vo, "out"));
	STEVI_TRY(index2d_to_disp_dev(c, vi, search, vo));
	return st.finish();
}

int32_t stevi_truncated_cost_volume_2d(stevi_ctx* c, const stevi_array* cv, const stevi_array* idx, int32_t radius0, int32_t radius1,
									   stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vc, vi, vo;
	STEVI_TRY(st.in(cv, vc, "cv"));
	STEVI_TRY(st.in(idx, vi, "idx"));
	STEVI_TRY(st.out(out, vo, "out"));
	STEVI_TRY(truncated_cv2d_dev(c, vc, vi, radius0, radius1, vo));
	return st.finish();
}

int32_t stevi_refine_disp_2d_cost_interp(stevi_ctx* c, const stevi_array* tcv, const stevi_array* raw, int32_t kernel, int32_t isotropy,
										 stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vt, vi, vo;
	STEVI_TRY(st.in(tcv, vt, "tcv"));
	STEVI_TRY(st.in(raw, vi, "raw_disparity"));
	STEVI_TRY(st.out(out, vo, "out"));
	STEVI_TRY(refine2d_interp_dev(c, vt, vi, kernel, isotropy, vo));
	return st.finish();
}

int32_t stevi_refine_disp_2d_cost_patch_interp(stevi_ctx* c, const stevi_array* tcv, const stevi_array* raw, int32_t kernel, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vt, vi, vo;
	STEVI_TRY(st.in(tcv, vt, "tcv"));
	STEVI_TRY(st.in(raw, vi, "raw_disparity"));
	STEVI_TRY(st.out(out, vo, "out"));
	STEVI_TRY(refine2d_patch_dev(c, vt, vi, kernel, vo));
	return st.finish();
}

int32_t stevi_variable_cost_volume_aggregation(stevi_ctx* c, const stevi_array* cv, const stevi_array* windows, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	if (cv && out && cv->data == out->data && cv->data) return fail(c, STEVI_E_ARG, "variable_cost_volume_aggregation: out must not alias cv");
	Stager st(c);
	View vc, vw, vo;
	STEVI_TRY(st.in(cv, vc, "cv"));
	STEVI_TRY(st.in(windows, vw, "windows"));
	STEVI_TRY(st.out(out, vo, "out"));
	STEVI_TRY(var_aggregate_dev(c, vc, vw, vo));
	return st.finish();
}

int32_t stevi_selected_cost(stevi_ctx* c, const stevi_array* cv, const stevi_array* idx, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vc, vi, vo;
	STEVI_TRY(st.in(cv, vc, "cv"));
	STEVI_TRY(st.in(idx, vi, "idx"));
	STEVI_TRY(st.out(out, vo, "out"));
	STEVI_TRY(selected_cost_dev(c, vc, vi, vo));
	return st.finish();
}

int32_t stevi_truncated_cost_volume(stevi_ctx* c, const stevi_array* cv, const stevi_array* idx, int32_t hr, int32_t vr, int32_t r,
									int32_t direction, int32_t tdir, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vc, vi, vo;
	STEVI_TRY(st.in(cv, vc, "cv"));
	STEVI_TRY(st.in(idx, vi, "idx"));
	STEVI_TRY(st.out(out, vo, "out"));
	STEVI_TRY(truncated_cv_dev(c, &vc, vi, 0, hr, vr, r, direction, tdir, vo));
	return st.finish();
}

int32_t stevi_in_bound_domain(stevi_ctx* c, const stevi_array* idx, int32_t width, int32_t hr, int32_t vr, int32_t r,
							  int32_t direction, int32_t tdir, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vi, vo;
	STEVI_TRY(st.in(idx, vi, "idx"));
	STEVI_TRY(st.out(out, vo, "out"));
	if (vi.ndim != 2) return fail(c, STEVI_E_SHAPE, "in_bound_domain: index map must be (H,W)");
	STEVI_TRY(truncated_cv_dev(c, nullptr, vi, width, hr, vr, r, direction, tdir, vo));
	return st.finish();
}

int32_t stevi_refine_disp_cost_interp(stevi_ctx* c, const stevi_array* tcv, const stevi_array* idx, int32_t kernel, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vt, vi, vo;
	STEVI_TRY(st.in(tcv, vt, "tcv"));
	STEVI_TRY(st.in(idx, vi, "idx"));
	STEVI_TRY(st.out(out, vo, "out"));
	STEVI_TRY(refine_interp_dev(c, vt, vi, kernel, vo));
	return st.finish();
}

int32_t stevi_refine_from_cost_volume(stevi_ctx* c, const stevi_array* cv, const stevi_array* idx, int32_t hr, int32_t vr,
									  int32_t kernel, stevi_array* tcv_out, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vc, vi, vt, vo;
	STEVI_TRY(st.in(cv, vc, "cv"));
	STEVI_TRY(st.in(idx, vi, "idx"));
	if (tcv_out) STEVI_TRY(st.out(tcv_out, vt, "tcv_out"));
	STEVI_TRY(st.out(out, vo, "out"));
	STEVI_TRY(refine_from_cv_dev(c, vc, vi, hr, vr, kernel, tcv_out ? &vt : nullptr, vo));
	return st.finish();
}

int32_t stevi_refine_barycentric_disp(stevi_ctx* c, int32_t mf, const stevi_array* fl, const stevi_array* fr, const stevi_array* idx,
									  int32_t direction, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vl, vr, vi, vo;
	STEVI_TRY(st.in(fl, vl, "feat_l"));
	STEVI_TRY(st.in(fr, vr, "feat_r"));
	STEVI_TRY(st.in(idx, vi, "idx"));
	STEVI_TRY(st.out(out, vo, "out"));
	STEVI_TRY(refine_barycentric_dev(c, mf, vl, vr, vi, direction, vo));
	return st.finish();
}

int32_t stevi_refine_barycentric_symmetric_disp(stevi_ctx* c, int32_t mf, int32_t radius, const stevi_array* fl, const stevi_array* fr,
												const stevi_array* idx, int32_t disp_width, int32_t direction, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vl, vr, vi, vo;
	STEVI_TRY(st.in(fl, vl, "feat_l"));
	STEVI_TRY(st.in(fr, vr, "feat_r"));
	STEVI_TRY(st.in(idx, vi, "idx"));
	STEVI_TRY(st.out(out, vo, "out"));
	STEVI_TRY(refine_barycentric_symmetric_dev(c, mf, radius, vl, vr, vi, disp_width, direction, vo));
	return st.finish();
}

int32_t stevi_refine_barycentric_2d_disp(stevi_ctx* c, int32_t mf, int32_t contiguity, int32_t symmetric, const stevi_array* fl,
										 const stevi_array* fr, const stevi_array* idx, const int32_t search[4], int32_t direction,
										 stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vl, vr, vi, vo;
	STEVI_TRY(st.in(fl, vl, "feat_l"));
	STEVI_TRY(st.in(fr, vr, "feat_r"));
	STEVI_TRY(st.in(idx, vi, "idx"));
	STEVI_TRY(st.out(out, vo, "out"));
	STEVI_TRY(refine_barycentric_2d_dev(c, mf, contiguity, symmetric, vl, vr, vi, search, direction, vo));
	return st.finish();
}

int32_t stevi_refine_cost_symmetric_disp(stevi_ctx* c, int32_t mf, const stevi_array* fl, const stevi_array* fr, const stevi_array* idx,
										 const stevi_array* cv, int32_t direction, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vl, vr, vi, vc, vo;
	STEVI_TRY(st.in(fl, vl, "feat_l"));
	STEVI_TRY(st.in(fr, vr, "feat_r"));
	STEVI_TRY(st.in(idx, vi, "idx"));
	STEVI_TRY(st.in(cv, vc, "cv"));
	STEVI_TRY(st.out(out, vo, "out"));
	STEVI_TRY(refine_cost_symmetric_dev(c, mf, vl, vr, vi, vc, direction, vo));
	return st.finish();
}

int32_t stevi_guided_cost_volume_features(stevi_ctx* c, int32_t mf, const stevi_array* fl, const stevi_array* fr, const stevi_array* guide,
										  int32_t radius, int32_t direction, stevi_array* tcv, stevi_array* disp) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vl, vr, vg, vt, vd;
	STEVI_TRY(st.in(fl, vl, "feat_l"));
	STEVI_TRY(st.in(fr, vr, "feat_r"));
	STEVI_TRY(st.in(guide, vg, "guide"));
	STEVI_TRY(st.out(tcv, vt, "tcv"));
	STEVI_TRY(st.out(disp, vd, "disp"));
	STEVI_TRY(guided_cv_features_dev(c, mf, vl, vr, vg, radius, direction, vt, vd));
	return st.finish();
}

int32_t stevi_patch_match(stevi_ctx* c, int32_t mf, const stevi_array* fs, const stevi_array* ft, int32_t dims, const int32_t* search,
						  int32_t n_iter, int32_t n_random, uint64_t seed, const stevi_array* init, stevi_array* disp) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vs, vt, vi, vd;
	STEVI_TRY(st.in(fs, vs, "feat_source"));
	STEVI_TRY(st.in(ft, vt, "feat_target"));
	if (init) STEVI_TRY(st.in(init, vi, "init"));
	STEVI_TRY(st.out(disp, vd, "disp"));
	STEVI_TRY(patch_match_dev(c, mf, vs, vt, dims, search, n_iter, n_random, seed, init ? &vi : nullptr, vd));
	return st.finish();
}

int32_t stevi_on_demand_truncated_cost_volume(stevi_ctx* c, int32_t mf, const stevi_array* fs, const stevi_array* ft, int32_t dims,
											  const int32_t* search, const stevi_array* disp, int32_t radius, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vs, vt, vd, vo;
	STEVI_TRY(st.in(fs, vs, "feat_source"));
	STEVI_TRY(st.in(ft, vt, "feat_target"));
	STEVI_TRY(st.in(disp, vd, "disp"));
	STEVI_TRY(st.out(out, vo, "out"));
	STEVI_TRY(on_demand_tcv_dev(c, mf, vs, vt, dims, search, vd, radius, vo));
	return st.finish();
}

int32_t stevi_mean_filter_2d(stevi_ctx* c, const stevi_array* img, int32_t hr, int32_t vr, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vi, vo;
	STEVI_TRY(st.in(img, vi, "img"));
	STEVI_TRY(st.out(out, vo, "out"));
	STEVI_TRY(mean_filter_dev(c, vi, hr, vr, vo));
	return st.finish();
}

int32_t stevi_sigma_filter(stevi_ctx* c, const stevi_array* mean, const stevi_array* img, int32_t hr, int32_t vr, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vm, vi, vo;
	STEVI_TRY(st.in(mean, vm, "mean"));
	STEVI_TRY(st.in(img, vi, "img"));
	STEVI_TRY(st.out(out, vo, "out"));
	STEVI_TRY(sigma_filter_dev(c, vm, vi, hr, vr, vo));
	return st.finish();
}

int32_t stevi_avgpool_downsample(stevi_ctx* c, const stevi_array* img, int32_t wh, int32_t wv, stevi_array* out) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vi, vo;
	STEVI_TRY(st.in(img, vi, "img"));
	STEVI_TRY(st.out(out, vo, "out"));
	STEVI_TRY(avgpool_dev(c, vi, wh, wv, vo));
	return st.finish();
}

int32_t stevi_guided_cost_volume(stevi_ctx* c, int32_t mf, const stevi_array* l, const stevi_array* r, int32_t hr, int32_t vr,
								 const stevi_array* guide, int32_t radius, int32_t direction, stevi_array* tcv, stevi_array* disp) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vl, vr_, vg, vt, vd;
	STEVI_TRY(st.in(l, vl, "img_l"));
	STEVI_TRY(st.in(r, vr_, "img_r"));
	STEVI_TRY(st.in(guide, vg, "guide"));
	STEVI_TRY(st.out(tcv, vt, "tcv"));
	STEVI_TRY(st.out(disp, vd, "disp"));
	STEVI_TRY(guided_cv_dev(c, mf, vl, vr_, hr, vr, vg, radius, direction, nullptr, vt, vd));
	return st.finish();
}

int32_t stevi_guided_cost_volume_band(stevi_ctx* c, int32_t mf, const stevi_array* l, const stevi_array* r, int32_t hr, int32_t vr,
									  const stevi_array* guide, int32_t radius, int32_t direction, const stevi_band* band,
									  stevi_array* tcv, stevi_array* disp) {
	STEVI_TRY(begin_call(c));
	if (!band) return fail(c, STEVI_E_ARG, "guided_cost_volume_band: band is required");
	Stager st(c);
	View vl, vr_, vg, vt, vd;
	STEVI_TRY(st.in(l, vl, "img_l"));
	STEVI_TRY(st.in(r, vr_, "img_r"));
	STEVI_TRY(st.in(guide, vg, "guide"));
	STEVI_TRY(st.out(tcv, vt, "tcv"));
	STEVI_TRY(st.out(disp, vd, "disp"));
	STEVI_TRY(guided_cv_dev(c, mf, vl, vr_, hr, vr, vg, radius, direction, band, vt, vd));
	return st.finish();
}

int32_t stevi_hierarchical(stevi_ctx* c, int32_t mf, int32_t depth, const stevi_array* l, const stevi_array* r, const int32_t* h_radii,
						   const int32_t* v_radii, int32_t hr, int32_t vr, int32_t D, int32_t radius, int32_t direction,
						   stevi_array* tcv, stevi_array* disp) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vl, vr_, vt, vd;
	STEVI_TRY(st.in(l, vl, "img_l"));
	STEVI_TRY(st.in(r, vr_, "img_r"));
	STEVI_TRY(st.out(tcv, vt, "tcv"));
	STEVI_TRY(st.out(disp, vd, "disp"));
	if (depth < 1 || depth > 8) return fail(c, STEVI_E_ARG, "hierarchical: depth must be 1..8");
	int32_t hrs[9], vrs[9];
	for (int k = 0; k <= depth; k++) { hrs[k] = h_radii ? h_radii[k] : hr; vrs[k] = v_radii ? v_radii[k] : vr; }
	STEVI_TRY(hierarchical_dev(c, mf, depth, vl, vr_, hrs, vrs, D, radius, direction, vt, vd));
	return st.finish();
}

int32_t stevi_census_sgm_pipeline(stevi_ctx* c, const stevi_array* l, const stevi_array* r, int32_t hr, int32_t vr, int32_t D,
								  int32_t direction, int32_t ndir, float P1, float P2, float Pout, const stevi_margins* margins,
								  int32_t semantics, stevi_array* cv_out, stevi_array* sgm_out, stevi_array* disp) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vl, vr_, vcv, vs, vd;
	STEVI_TRY(st.in(l, vl, "img_l"));
	STEVI_TRY(st.in(r, vr_, "img_r"));
	STEVI_TRY(check_images(c, vl, vr_));
	if (!disp) return fail(c, STEVI_E_ARG, "pipeline: disp is required");
	if (D < 1) return fail(c, STEVI_E_ARG, "pipeline: disp_width must be positive");   // before any workspace is sized by it
	// selectedIndexToDisp is folded into the result: disp = +-index (correlation_base.h:511-532)
	STEVI_TRY(st.out_index(disp, vd, D, direction == STEVI_RIGHT_TO_LEFT ? 1 : -1, 0, "disp"));
	const int H = vl.shape[0];
	const int Ws = direction == STEVI_RIGHT_TO_LEFT ? vr_.shape[1] : vl.shape[1];
	// volumes the caller does not want live in the workspace, in the reference's [H][D][W] layout
	auto workspace_volume = [&](int dtype, View& v) -> int32_t {
		void* p = nullptr;
		STEVI_TRY(arena_alloc(c, (size_t)H * Ws * D * dtype_size(dtype), &p));
		v.p = p; v.dtype = dtype; v.ndim = 3;
		v.shape[0] = H; v.shape[1] = Ws; v.shape[2] = D;
		v.st[0] = (int64_t)Ws * D; v.st[1] = 1; v.st[2] = Ws;
		return STEVI_OK;
	};
	stevi_margins mg = margins ? *margins : stevi_margins{0, 0, 0, 0};
	const bool pout_int = Pout == rintf(Pout) && fabsf(Pout) <= 1.0e6f;
	const bool collapse = semantics == STEVI_COMPAT && pout_int;
	if (vd.ndim != 2 || vd.dtype != STEVI_I32 || vd.shape[0] != H || vd.shape[1] != Ws) return fail(c, STEVI_E_SHAPE, "pipeline: disp must be (H,W) i32 (or u8 / u16 for RIGHT_TO_LEFT maps that fit)");
	// the WTA rides on the last SGM pass and selectedIndexToDisp (correlation_base.h:511-532) is folded into its output
	WtaSink sink{vd, direction == STEVI_RIGHT_TO_LEFT ? 1 : -1, 0};
	const int max_cost = 32 * census_n_written(hr, vr, semantics);
	if (!cv_out && !sgm_out && collapse && D >= 1) {
		// disparity only: neither volume is ever in memory; the Hamming costs are recomputed from the census words by the
		// g-map pass and by the aggregation + winner-take-all pass
		const int nw = census_n_words(hr, vr);
		View cl, cr;
		STEVI_TRY(alloc_census(c, vl.shape[0], vl.shape[1], nw, cl));
		STEVI_TRY(alloc_census(c, vr_.shape[0], vr_.shape[1], nw, cr));
		const View imgs[2] = {vl, vr_}, cens[2] = {cl, cr};
		STEVI_TRY(census_transform_pair_dev(c, imgs, cens, 2, hr, vr, semantics));
		HamSource hs{cl, cr, census_n_written(hr, vr, semantics), D, direction, semantics == STEVI_COMPAT};
		const int32_t s = sgm_census_disparity_dev(c, hs, ndir, (int)Pout, mg, max_cost, sink);
		if (s != STEVI_NOT_FUSED) {
			STEVI_TRY(s);
			return st.finish();
		}
	}
	if (cv_out) STEVI_TRY(st.out(cv_out, vcv, "cv_out")); else STEVI_TRY(workspace_volume(STEVI_U16, vcv));
	if (sgm_out) STEVI_TRY(st.out(sgm_out, vs, "sgm_out"));
	else {
		// not asked for: the aggregated costs are evaluated in registers by the winner-take-all pass and never stored
		// (sgm_dev allocates a workspace volume only on the paths that need one)
		vs.p = nullptr; vs.dtype = STEVI_F32; vs.ndim = 3;
		vs.shape[0] = H; vs.shape[1] = Ws; vs.shape[2] = D;
		vs.st[0] = (int64_t)Ws * D; vs.st[1] = 1; vs.st[2] = Ws;
	}
	if (vcv.dtype != STEVI_U16) return fail(c, STEVI_E_DTYPE, "pipeline: cv_out must be u16 (hamming_cv_t)");
	int32_t* gmap = nullptr;
	if (collapse) STEVI_TRY(arena_alloc_t(c, (size_t)H * Ws, &gmap));
	STEVI_TRY(census_cv_dev(c, vl, vr_, hr, vr, D, direction, semantics, vcv, gmap, (int)Pout));
	STEVI_TRY(sgm_dev(c, vcv, ndir, STEVI_COST, P1, P2, Pout, mg, semantics, vs, gmap, max_cost, &sink));
	return st.finish();
}

int32_t stevi_local_matching_pipeline(stevi_ctx* c, int32_t mf, const stevi_array* l, const stevi_array* r, int32_t hr, int32_t vr, int32_t D,
									  int32_t direction, int32_t kernel, stevi_array* cv_out, stevi_array* idx_out, stevi_array* tcv_out,
									  stevi_array* refined) {
	STEVI_TRY(begin_call(c));
	Stager st(c);
	View vl, vr_, vcv, vidx, vt, vo;
	STEVI_TRY(st.in(l, vl, "img_l"));
	STEVI_TRY(st.in(r, vr_, "img_r"));
	STEVI_TRY(check_images(c, vl, vr_));
	if (!refined) return fail(c, STEVI_E_ARG, "pipeline: refined is required");
	if (D < 1) return fail(c, STEVI_E_ARG, "pipeline: disp_width must be positive");
	STEVI_TRY(st.out(refined, vo, "refined"));
	const int H = vl.shape[0];
	const int Ws = direction == STEVI_RIGHT_TO_LEFT ? vr_.shape[1] : vl.shape[1];
	if (cv_out) STEVI_TRY(st.out(cv_out, vcv, "cv_out"));
	else {
		float* p = nullptr;
		STEVI_TRY(arena_alloc_t(c, (size_t)H * Ws * std::max(D, 0), &p));
		vcv.p = p; vcv.dtype = STEVI_F32; vcv.ndim = 3;
		vcv.shape[0] = H; vcv.shape[1] = Ws; vcv.shape[2] = D;
		vcv.st[0] = (int64_t)Ws * D; vcv.st[1] = 1; vcv.st[2] = Ws;   // the reference's [H][D][W] layout
	}
	if (idx_out) STEVI_TRY(st.out_index(idx_out, vidx, D, 1, 0, "idx_out"));
	else {
		int32_t* p = nullptr;
		STEVI_TRY(arena_alloc_t(c, (size_t)H * Ws, &p));
		vidx.p = p; vidx.dtype = STEVI_I32; vidx.ndim = 2; vidx.shape[0] = H; vidx.shape[1] = Ws; vidx.st[0] = Ws; vidx.st[1] = 1;
	}
	if (tcv_out) STEVI_TRY(st.out(tcv_out, vt, "tcv_out"));
	// the winner-take-all rides on the cost-volume kernel: the 4*W*H*D-byte volume is written once and never read back
	STEVI_TRY(float_cv_dev(c, mf, vl, vr_, hr, vr, D, direction, vcv, false, &vidx));
	STEVI_TRY(refine_from_cv_dev(c, vcv, vidx, hr, vr, kernel, tcv_out ? &vt : nullptr, vo));
	return st.finish();
}

} // extern "C"
