// Device-operand entry points of each stage (host wrappers that launch the kernels on ctx->stream). The C ABI
// functions in api.cu stage host operands and then call these; the fused pipelines chain them directly.
#pragma once
#include "common.cuh"

namespace stevi {

// census.cu
int32_t census_n_words(int hr, int vr);
int32_t census_n_written(int hr, int vr, int semantics);
int32_t census_transform_dev(stevi_ctx* c, const View& img, int hr, int vr, int semantics, const View& out);
int32_t census_transform_pair_dev(stevi_ctx* c, const View* imgs, const View* outs, int n, int hr, int vr, int semantics);
// feat_* (H,W*,nw) u32; n_cmp = how many leading words to compare (the rest are known to be zero on both sides);
// gmap (optional, int32 [H*Ws] dense): min_d (2*C + pout_i*[j+d >= Ws]) for the collapsed SGM (sgm.cu)
int32_t hamming_cv_dev(stevi_ctx* c, const View& feat_l, const View& feat_r, int n_cmp, int D, int direction,
					   bool round_target, const View& out, int32_t* gmap, int pout_i);

// costvol.cu: zero-padded float copies of an image pair plus two per-pixel window-statistics maps per image, all with
// the same geometry (pointers are at image row 0, padded column 0; `pitch` floats per row in the unit-de-interleaved
// layout of costvol_box.cuh; pads are zero). index 0 = left, 1 = right.
struct PaddedPair {
	int H = 0, W[2] = {0, 0};
	int pad_left = 0, pad_top = 0, rows = 0, half = 0;
	int64_t pitch = 0;
	size_t plane = 0;
	float* img[2] = {nullptr, nullptr};
	float* A[2] = {nullptr, nullptr};
	float* B[2] = {nullptr, nullptr};
	// extra_rows: further zero rows above and below (2-D search: the target planes are read shifted by whole rows)
	int32_t build(stevi_ctx* c, const View& l, const View& r, int hr, int vr, int D, int band_rows, int extra_rows = 0);
	int32_t stats(stevi_ctx* c, int which, int hr, int vr, int mf, int kind);
};
// unfoldBasedCostVolume for CC/NCC/ZCC/ZNCC/SSD/ZSSD/SAD/ZSAD, out (H,Wsource,D) f32
// wta_out (optional, (H,Wsource) i32): extractSelectedIndex of the volume with the function's own strategy, fused into the
// cost-volume kernel (the volume is not read back)
int32_t float_cv_dev(stevi_ctx* c, int mf, const View& l, const View& r, int hr, int vr, int D, int direction, const View& out,
					 bool force_direct, const View* wta_out = nullptr);

// sgm.cu
// optional fused winner-take-all of the aggregated volume: out(i,j) = sign * argmin_d + offset (selectedIndexToDisp folded in)
struct WtaSink { View out; int sign; int offset; };
// Hamming costs recomputed from census words instead of read from a cost volume (disparity-only census+SGM pipeline)
struct HamSource { View fl, fr; int n_cmp, D, direction; bool round; };
bool hamming_fused_capable(const HamSource& h);
int32_t hamming_gmap_dev(stevi_ctx* c, const HamSource& h, int32_t* gmap, int pout_i);
// one pass: per pixel the minima over d of cost * 8192 + (8191 - d), separately for j + d < W (kin) and j + d >= W (kout)
// (*nz slices of H*W keys each, one per disparity chunk of the pass, allocated in the ctx workspace: the consumer takes their minimum)
int32_t hamming_minkeys_dev(stevi_ctx* c, const HamSource& h, uint32_t** kin, uint32_t** kout, int* nz);
// r_max: upper bound of R (number of passes over a pixel)
int32_t hamming_sgm_wta_dev(stevi_ctx* c, const HamSource& h, const int32_t* T, const int32_t* R, int pout_i, int r_max,
							unsigned long long* keys);
// census -> SGM -> WTA with no volume in memory; STEVI_NOT_FUSED when the collapsed evaluation does not apply (the caller
// then runs the materialised path)
constexpr int32_t STEVI_NOT_FUSED = -1001;
int32_t sgm_census_disparity_dev(stevi_ctx* c, const HamSource& h, int ndir, int pout, const stevi_margins& mg, int max_cost,
								 const WtaSink& wta);
int32_t sgm_dev(stevi_ctx* c, const View& cv, int ndir, int strategy, float P1, float P2, float Pout,
				const stevi_margins& mg, int semantics, const View& out, const int32_t* gmap_ready, int max_cost_hint,
				const WtaSink* wta);

// costvol.cu: 2-D search cost volume, search = {lower0, upper0, lower1, upper1}, out (H,W,dH,dW) f32
// exact: the reference's O(F) sums in its channel order (2e-6); else the O(1) box sums per row offset (1e-5), where a box kernel exists
int32_t float_cv2d_dev(stevi_ctx* c, int mf, const View& l, const View& r, int hr, int vr, const int32_t* search, int direction, const View& out,
					   bool exact = false);

// features_cv.cu: feature-level forms (caller-supplied float feature volumes)
int32_t features_for_matchfunc_dev(stevi_ctx* c, int mf, const View& in, const View& out);
// cv(i,j,d) = compare(source(i,j,.), target(i, j + sign*(d + lower), .)), d < D
int32_t aggregate_cost_dev(stevi_ctx* c, int mf, const View& fl, const View& fr, int lower, int D, int direction, const View& out);

// wta.cu
int32_t wta2d_dev(stevi_ctx* c, const View& vol, int strategy, const View& out);
int32_t index2d_to_disp_dev(stevi_ctx* c, const View& idx, const int32_t* search, const View& out);
int32_t wta_dev(stevi_ctx* c, const View& vol, int strategy, const View& out);
// keys: dense [H][W], (order bits << 32 | d) merged with atomicMax, 0 = never written -> index 0
int32_t wta_decode_max_keys_dev(stevi_ctx* c, const unsigned long long* keys, int H, int W, const View& out);
int32_t index_to_disp_dev(stevi_ctx* c, const View& idx, int direction, int offset, const View& out);


// refine.cu
// cv == nullptr: extractInBoundDomain (0/1 mask, `width` = search width); else truncatedCostVolume of *cv
int32_t truncated_cv_dev(stevi_ctx* c, const View* cv, const View& idx, int width, int hr, int vr, int r, int direction, int tdir,
						 const View& out);
int32_t selected_cost_dev(stevi_ctx* c, const View& cv, const View& idx, const View& out);
int32_t refine_interp_dev(stevi_ctx* c, const View& tcv, const View& idx, int kernel, const View& out);
// truncatedCostVolume<Same>(radius 1) + refineDispCostInterpolation in one pass over the winners; tcv_out optional
int32_t refine_from_cv_dev(stevi_ctx* c, const View& cv, const View& idx, int hr, int vr, int kernel, const View* tcv_out, const View& out);

// 2-D search: truncatedBidirectionaCostVolume, refineDisp2dCostInterpolation
int32_t truncated_cv2d_dev(stevi_ctx* c, const View& cv, const View& idx, int radius0, int radius1, const View& out);
int32_t refine2d_interp_dev(stevi_ctx* c, const View& tcv, const View& raw, int kernel, int isotropy, const View& out);
int32_t refine2d_patch_dev(stevi_ctx* c, const View& tcv, const View& raw, int kernel, const View& out);

// barycentric.cu: feature-domain refinement (image_based_refinement.h); idx2 / out2 are (H,W,2)
int32_t refine_barycentric_dev(stevi_ctx* c, int mf, const View& fl, const View& fr, const View& idx, int direction, const View& out);
int32_t refine_barycentric_symmetric_dev(stevi_ctx* c, int mf, int radius, const View& fl, const View& fr, const View& idx, int disp_width,
										 int direction, const View& out);
int32_t refine_barycentric_2d_dev(stevi_ctx* c, int mf, int contiguity, int symmetric, const View& fl, const View& fr, const View& idx2,
								  const int32_t* search, int direction, const View& out2);
int32_t refine_cost_symmetric_dev(stevi_ctx* c, int mf, const View& fl, const View& fr, const View& idx, const View& cv, int direction, const View& out);

// patchmatch.cu: on-demand cost evaluation and PatchMatch (on_demand_cost_volume.h, patchmatch.h); search = {lower, upper} for a
// 1-D search (columns), {lower0, upper0, lower1, upper1} for a 2-D one; disp (H,W,dims) i32
int32_t patch_match_dev(stevi_ctx* c, int mf, const View& fs, const View& ft, int dims, const int32_t* search, int n_iter, int n_random,
						unsigned long long seed, const View* init, const View& disp);
int32_t on_demand_tcv_dev(stevi_ctx* c, int mf, const View& fs, const View& ft, int dims, const int32_t* search, const View& disp, int radius,
						  const View& out);

// aggregate.cu: variableCostVolumeAggregation (guided_cost_filtering.h:102-176)
int32_t var_aggregate_dev(stevi_ctx* c, const View& cv, const View& win, const View& out);

// hier.cu
int32_t avgpool_dev(stevi_ctx* c, const View& img, int win_h, int win_v, const View& out);
// filters.cu: meanFilter2D / sigmaFilter (correlation_base.h:881-1077)
int32_t mean_filter_dev(stevi_ctx* c, const View& img, int hr, int vr, const View& out);
int32_t sigma_filter_dev(stevi_ctx* c, const View& mean, const View& img, int hr, int vr, const View& out);
// band == nullptr: whole images; else the images / guide are row bands placed in global coordinates (c_api.h: stevi_band)
int32_t guided_cv_dev(stevi_ctx* c, int mf, const View& l, const View& r, int hr, int vr, const View& guide, int radius, int direction,
					  const stevi_band* band, const View& tcv, const View& disp);
// computeGuidedCV on caller-supplied (processed) feature volumes, the reference's own signature (hierarchical.h:73-233)
int32_t guided_cv_features_dev(stevi_ctx* c, int mf, const View& fl, const View& fr, const View& guide, int radius, int direction,
							   const View& tcv, const View& disp);
int32_t hierarchical_dev(stevi_ctx* c, int mf, int depth, const View& l, const View& r, const int32_t* h_radii, const int32_t* v_radii,
						 int D, int radius, int direction, const View& tcv, const View& disp);

// unfold.cu (materialised feature volumes; pad = {left, top, right, bottom}, NULL or any negative entry = auto)
int32_t unfold_dev(stevi_ctx* c, const View& img, int hr, int vr, const int32_t* pad, int orientation, const View& out);
int32_t unfold_compressed_dev(stevi_ctx* c, const int32_t* mask, int mh, int mw, int64_t m0, int64_t m1, const View& img,
							  const int32_t* pad, const View& out);
int32_t census_features_dev(stevi_ctx* c, const View& fv, int semantics, const View& out);
int32_t feature_volume_dev(stevi_ctx* c, int mf, const View& img, int hr, int vr, const View& out);

} // namespace stevi
