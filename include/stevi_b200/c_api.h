/* stevi_b200 — C ABI of the B200-native (sm_100a) dense-stereo hot path of LibStevi.
 *
 * This is the drop-in boundary: plain pointers, sizes, enums as int32 — no C++ or torch types. The reference has
 * no FFI of its own (it is a header-only template library); each entry point below names the
 * reference template it stands in for (file:line relative to the LibStevi tree). The C++ facade under
 * include/StereoVision/ re-creates those templates over Multidim::Array and forwards to these functions.
 *
 * Conventions
 *  - every function returns a stevi_status (0 = OK); stevi_last_error(ctx) has the text of the last failure;
 *  - arrays are described by stevi_array: element strides (not bytes), any order; `mem` says whether `data` is a
 *    host or a device pointer. Host operands are staged through the ctx workspace (H2D before, D2H after, the
 *    call returns when the results are in host memory). Device operands are used in place and the call is
 *    asynchronous on the ctx stream;
 *  - the library never allocates caller-visible memory: outputs are pre-shaped by the caller;
 *  - there is NO CPU fallback: without a usable CUDA device stevi_ctx_create fails with STEVI_E_CUDA.
 *  - `semantics`: STEVI_COMPAT reproduces the reference bit for bit including its quirks,
 *    STEVI_CORRECTED applies the one-line fixes listed there.
 */
#ifndef STEVI_B200_C_API_H
#define STEVI_B200_C_API_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define STEVI_API __attribute__((visibility("default")))
#else
#define STEVI_API
#endif

#define STEVI_MAX_DIMS 4

typedef enum {
	STEVI_OK = 0,
	STEVI_E_SHAPE = 1,       /* the reference returns an empty array in these cases */
	STEVI_E_DTYPE = 2,
	STEVI_E_UNSUPPORTED = 3,
	STEVI_E_CUDA = 4,
	STEVI_E_NCCL = 5,
	STEVI_E_ARG = 6
} stevi_status;

typedef enum { STEVI_U8 = 0, STEVI_U16 = 1, STEVI_I32 = 2, STEVI_U32 = 3, STEVI_F32 = 4 } stevi_dtype;
typedef enum { STEVI_HOST = 0, STEVI_DEVICE = 1 } stevi_mem;
typedef enum { STEVI_COMPAT = 0, STEVI_CORRECTED = 1,
	/* flag, OR-ed in: float cost volumes evaluate the reference's own O(F) per-cell sums in its channel order (slow, as
	 * close to the reference bits as IEEE arithmetic allows) instead of the O(1) box sums */
	STEVI_EXACT_SUMS = 256 } stevi_semantics;

/* mirrors StereoVision::Correlation::matchingFunctions, correlation/matching_costs.h:38-53 */
typedef enum {
	STEVI_MF_CC = 0, STEVI_MF_NCC = 1, STEVI_MF_SSD = 2, STEVI_MF_SAD = 3, STEVI_MF_ZCC = 4, STEVI_MF_ZNCC = 5,
	STEVI_MF_ZSSD = 6, STEVI_MF_ZSAD = 7, STEVI_MF_MEDAD = 8, STEVI_MF_ZMEDAD = 9, STEVI_MF_HAMMING = 10,
	STEVI_MF_CENSUS = 11, STEVI_MF_KERMI = 12
} stevi_matchfunc;
/* dispExtractionStartegy (sic), dispDirection, truncatedCostVolumeDirection: correlation/correlation_base.h:31-45 */
typedef enum { STEVI_COST = 0, STEVI_SCORE = 1 } stevi_strategy;
typedef enum { STEVI_LEFT_TO_RIGHT = 0, STEVI_RIGHT_TO_LEFT = 1 } stevi_direction;
typedef enum { STEVI_TCV_SAME = 0, STEVI_TCV_REVERSED = 1, STEVI_TCV_BOTH = 2 } stevi_tcv_direction;
/* InterpolationKernel: correlation/cost_based_refinement.h:30-35 */
typedef enum { STEVI_EQUIANGULAR = 0, STEVI_PARABOLA = 1, STEVI_GAUSSIAN = 2 } stevi_interp_kernel;
/* IsotropyHypothesis: correlation/cost_based_refinement.h:37-41 */
typedef enum { STEVI_ISOTROPIC = 0, STEVI_ANISOTROPIC = 1 } stevi_isotropy;

typedef struct {
	void* data;
	int32_t dtype;                    /* stevi_dtype */
	int32_t ndim;
	int32_t shape[STEVI_MAX_DIMS];
	int64_t strides[STEVI_MAX_DIMS];  /* in elements */
	int32_t mem;                      /* stevi_mem */
} stevi_array;

/* StereoVision::Margins(left, top, right, bottom), utils/margins.h:24-60 */
typedef struct { int32_t left, top, right, bottom; } stevi_margins;

typedef struct stevi_ctx stevi_ctx;

/* ---- context: one per host thread / CUDA stream (the reference's functions are stateless and re-entrant; the
 * ctx only owns the device id, the stream and a grow-only device workspace). `stream` is a cudaStream_t or NULL
 * to let the ctx create its own non-blocking stream. */
STEVI_API int32_t stevi_ctx_create(int32_t device, void* stream, stevi_ctx** out);
STEVI_API int32_t stevi_ctx_destroy(stevi_ctx* ctx);
STEVI_API int32_t stevi_ctx_set_stream(stevi_ctx* ctx, void* stream);
STEVI_API void* stevi_ctx_stream(stevi_ctx* ctx);
/* Row-band runs (one frame split over several GPUs): tells the ctx which global row of the frame is row 0 of the images
 * passed to the following calls. It only anchors the kernels' internal row blocking (the float cost volumes keep running
 * window sums per block of 32 rows), so that a band gives bit-identical rows to the whole-frame call. Default 0. */
STEVI_API int32_t stevi_ctx_set_row_origin(stevi_ctx* ctx, int32_t global_row_of_row0);
/* Streaming many pairs from host memory: with deferred host completion on, a call whose operands live in host memory
 * enqueues its host->device copies, kernels and device->host copies on the ctx stream and RETURNS WITHOUT WAITING; the
 * host buffers (page-locked, or the copies do not overlap) must stay alive and untouched, and the results are valid only
 * after stevi_ctx_synchronize. One ctx (own stream, own workspace) per pair in flight lets the copies of one pair overlap
 * the kernels of another. Default off: calls with host operands return with their results in place. */
STEVI_API int32_t stevi_ctx_set_deferred_host(stevi_ctx* ctx, int32_t enable);
STEVI_API int32_t stevi_ctx_synchronize(stevi_ctx* ctx);
/* Disparity / index maps that go back to HOST memory as dense (H,W) int32 arrays (disp_t, correlation_base.h:47) cross PCIe as
 * the narrowest exact type (u8 for up to 256 disparities, u16 up to 65536) and are widened into the caller's int32 array by the
 * host when the copy has landed (inside the call, or inside stevi_ctx_synchronize with deferred host completion): the same
 * values with a quarter / half of the device->host bytes. On by default; 0 sends the int32 map itself. The widening is done by a
 * process-wide pool of STEVI_HOST_THREADS (default 4) worker threads, started by the first call that needs them (they inherit
 * that thread's CPU affinity); a map is handed to them by the first library call that finds its copy landed, at the latest by
 * the synchronising call, which also takes part of the work. Which setting is faster depends on the box: narrow when many
 * GPUs share the host's PCIe path (8 GPUs: 0.69 vs 0.87 ms per 1080p map), the int32 map itself when host cores / memory
 * bandwidth are the scarcer resource (1 GPU: 0.27 vs 0.33 ms). Callers that can use a u8 / u16 map pass one and skip both. */
STEVI_API int32_t stevi_ctx_set_narrow_wire(stevi_ctx* ctx, int32_t enable);
STEVI_API const char* stevi_last_error(stevi_ctx* ctx);
/* number of kernels this ctx has launched so far (for the benchmark's gpu_launches claim) */
STEVI_API int64_t stevi_ctx_launch_count(stevi_ctx* ctx);
/* per-kernel device timing for benchmarks: when enabled every kernel launch of this ctx is bracketed by a CUDA event
 * pair on the ctx stream; the report is "name<TAB>milliseconds<NL>" per launch since the previous report (the call
 * synchronises the stream). Leave it off for production calls. */
STEVI_API int32_t stevi_ctx_set_timing(stevi_ctx* ctx, int32_t enable);
STEVI_API const char* stevi_ctx_timing_report(stevi_ctx* ctx);
/* bytes currently held by the workspace arena */
STEVI_API int64_t stevi_ctx_workspace_bytes(stevi_ctx* ctx);
STEVI_API int32_t stevi_device_count(void);
STEVI_API const char* stevi_version(void);

/* pinned host memory for callers that want DMA-speed staging of host operands */
STEVI_API int32_t stevi_host_alloc(size_t bytes, void** out);
STEVI_API int32_t stevi_host_free(void* p);

/* ---- features ------------------------------------------------------------------------------------------- */

/* censusTransform2D (correlation/census.h:117-131) = unfold (unfold.h:248-291) + censusFeatures (census.h:69-115).
 * img (H,W) u8|f32  ->  out (H,W,nwords) u32, nwords = (F-1)/32+1, F=(2*h_radius+1)*(2*v_radius+1).
 * COMPAT: only completely filled 32-bit words are written, the trailing word is 0 (Q1). */
STEVI_API int32_t stevi_census_transform(stevi_ctx* ctx, const stevi_array* img, int32_t h_radius, int32_t v_radius,
							   int32_t semantics, stevi_array* out);

/* unfold<T_I,T_O>(h_radius, v_radius, img, padding, orientation) (correlation/unfold.h:248-344): img (H,W) or (H,W,C)
 * u8|f32 -> out (H',W',F*C) u8|f32, out(i,j,c) = img(i+k-pad_top, j+l-pad_left, ch), 0 outside, c = channelFromCord
 * (:171-191; orientation 0..3 = Rotate0/90/180/270). padding = {left, top, right, bottom} like PaddingMargins; NULL or a
 * negative entry = automatic (= the radii: H' = H, W' = W). The reference gives the 2-D overload's result strides
 * {W*F, F, 1} (:272); any strides are accepted. */
STEVI_API int32_t stevi_unfold(stevi_ctx* ctx, const stevi_array* img, int32_t h_radius, int32_t v_radius, const int32_t* padding,
					 int32_t orientation, stevi_array* out);

/* unfold(UnFoldCompressor(mask), img, padding) (correlation/unfold.h:36-137, :346-471): mask (mh,mw) i32 in HOST memory
 * (labels > 0 form super-pixels, features in ascending label order); img (H,W) or (H,W,C) u8|f32 ->
 * out (H',W',nFeatures*C) f32 of super-pixel means. */
STEVI_API int32_t stevi_unfold_compressed(stevi_ctx* ctx, const stevi_array* mask, const stevi_array* img, const int32_t* padding,
								stevi_array* out);

/* censusFeatures (correlation/census.h:69-115) on a ready-made feature volume: feat (H,W,F) u8|f32 -> out (H,W,nwords) u32 */
STEVI_API int32_t stevi_census_features(stevi_ctx* ctx, const stevi_array* feat, int32_t semantics, stevi_array* out);

/* getFeatureVolumeForMatchFunc<matchFunc>(unfold(h_radius, v_radius, img)) (correlation/cross_correlations.h:645-722) for
 * float images: img (H,W) f32 -> out (H,W,F) f32 = v | v-mean | v/norm | (v-mean)/norm per the function's traits. */
STEVI_API int32_t stevi_feature_volume(stevi_ctx* ctx, int32_t matchfunc, const stevi_array* img, int32_t h_radius,
							 int32_t v_radius, stevi_array* out);

/* ---- cost volumes ---------------------------------------------------------------------------------------- */

/* unfoldBasedCostVolume<matchFunc, T_L, T_R, 2, dDir, TCV> (correlation/cross_correlations.h:740-765).
 * img_l (H,Wl), img_r (H,Wr): u8|f32 (same dtype). out (H, Wsource, D) of dtype TCV, Wsource = Wr for
 * RIGHT_TO_LEFT, Wl for LEFT_TO_RIGHT. CENSUS: out u16|i32|f32 (hamming_cv_t = u16, matching_costs.h:234).
 * CC, NCC, ZCC, ZNCC, SSD, ZSSD, SAD, ZSAD: out f32; u8 images only for CC/SSD/SAD (the reference's int16 fixed-point
 * ZNCC/NCC features for u8 images, cross_correlations.h:416-462, are not provided). Flat windows give NaN for the
 * normalised functions, like the reference (Q6). MEDAD, ZMEDAD, KERMI: STEVI_E_UNSUPPORTED.
 * The reference lays the volume out as strides {W*D, 1, W} (cross_correlations.h:220); any strides are accepted.
 * Returns STEVI_E_SHAPE where the reference returns an empty array (heights differ, :751). */
STEVI_API int32_t stevi_cost_volume(stevi_ctx* ctx, int32_t matchfunc, const stevi_array* img_l, const stevi_array* img_r,
						  int32_t h_radius, int32_t v_radius, int32_t disp_width, int32_t direction,
						  int32_t semantics, stevi_array* out);

/* aggregateCost<HAMMING|CENSUS> on ready-made census/binary features (correlation/cross_correlations.h:194-249):
 * feat_l (H,Wl,nw) u32, feat_r (H,Wr,nw) u32 -> out (H,Wsource,D). COMPAT rounds every target word through
 * float (Q2, :235-236). */
STEVI_API int32_t stevi_hamming_cost_volume(stevi_ctx* ctx, const stevi_array* feat_l, const stevi_array* feat_r,
								  int32_t disp_width, int32_t direction, int32_t semantics, stevi_array* out);

/* ---- feature-level forms: callers that bring their own (H,W,F) float feature volumes ---------------------------------- */

/* getFeatureVolumeForMatchFunc<matchFunc>(feature_vol) (correlation/cross_correlations.h:645-722) on a ready-made volume:
 * feat (H,W,F) f32 -> out (H,W,F) f32 = v | v-mean | v/norm | (v-mean)/norm per the function's traits (CC..ZSAD). */
STEVI_API int32_t stevi_features_for_matchfunc(stevi_ctx* ctx, int32_t matchfunc, const stevi_array* feat, stevi_array* out);

/* aggregateCost<matchFunc, float, float, dDir, float> (correlation/cross_correlations.h:194-249) and its searchOffset<1> overload
 * (:251-308): feat_l (H,Wl,F), feat_r (H,Wr,F) f32 -> out (H,Wsource,upper-lower+1) f32, cell d compares the source vector with the
 * target vector lower + d columns away (zero features outside the target image). The disp_width overload is lower = 0,
 * upper = disp_width - 1. The features are taken as they are (apply stevi_features_for_matchfunc first where the function wants it). */
STEVI_API int32_t stevi_aggregate_cost(stevi_ctx* ctx, int32_t matchfunc, const stevi_array* feat_l, const stevi_array* feat_r,
							 int32_t lower_offset, int32_t upper_offset, int32_t direction, stevi_array* out);

/* featureVolume2CostVolume<matchFunc> (correlation/cross_correlations.h:724-738): getFeatureVolumeForMatchFunc of both volumes, then
 * aggregateCost. feat_l, feat_r (H,W*,F) f32 (raw unfolded features) -> out (H,Wsource,D): f32 for CC..ZSAD; u16|i32|f32 for CENSUS
 * (censusFeatures + Hamming distances, :712-714). */
STEVI_API int32_t stevi_feature_volume_to_cost_volume(stevi_ctx* ctx, int32_t matchfunc, const stevi_array* feat_l, const stevi_array* feat_r,
											int32_t disp_width, int32_t direction, int32_t semantics, stevi_array* out);

/* unfoldBasedCostVolume(img_l, img_r, UnFoldCompressor const&, disp_width) (correlation/cross_correlations.h:767-791): mask (mh,mw)
 * i32 in HOST memory (as for stevi_unfold_compressed); images (H,W) or (H,W,C) u8|f32. */
STEVI_API int32_t stevi_cost_volume_compressed(stevi_ctx* ctx, int32_t matchfunc, const stevi_array* mask, const stevi_array* img_l,
									 const stevi_array* img_r, int32_t disp_width, int32_t direction, int32_t semantics, stevi_array* out);

/* unfoldBasedCostVolume<matchFunc, T_L, T_R, 3, ...> (correlation/cross_correlations.h:740-765 over the 3-D image unfold,
 * unfold.h:293-344): img_l, img_r (H,W,C) u8|f32 with equal channel counts. */
STEVI_API int32_t stevi_cost_volume_multichannel(stevi_ctx* ctx, int32_t matchfunc, const stevi_array* img_l, const stevi_array* img_r,
									   int32_t h_radius, int32_t v_radius, int32_t disp_width, int32_t direction, int32_t semantics,
									   stevi_array* out);

/* ---- semi-global matching -------------------------------------------------------------------------------- */

/* sgmCostVolume<nDirections, strategy, T_CV> (correlation/sgm.h:360-404). cv (H,W,D) u16|f32 -> out (H,W,D) f32.
 * n_directions in {4,8,16}. COMPAT reproduces Q3 (Cost mode ignores P1/P2) and Q4 (dead / half-run / doubled
 * directions); CORRECTED is the textbook recurrence over all directions. */
STEVI_API int32_t stevi_sgm(stevi_ctx* ctx, const stevi_array* cv, int32_t n_directions, int32_t strategy,
				  float P1, float P2, float Pout, const stevi_margins* margins, int32_t semantics, stevi_array* out);

/* ---- winner-take-all and friends -------------------------------------------------------------------------- */

/* extractSelectedIndex<strategy, T_CV> (correlation/correlation_base.h:427-464): vol (H,W,D) u16|i32|f32 ->
 * out (H,W) i32. `<=` / `>=`: the LAST extremum wins (Q5); a NaN at d=0 sticks. */
STEVI_API int32_t stevi_wta(stevi_ctx* ctx, const stevi_array* vol, int32_t strategy, stevi_array* out);

/* selectedIndexToDisp<disp_t, dDir> (correlation/correlation_base.h:511-532): disp = +-idx + offset */
STEVI_API int32_t stevi_index_to_disp(stevi_ctx* ctx, const stevi_array* idx, int32_t direction, int32_t offset,
							stevi_array* out);

/* ---- 2-D search ---------------------------------------------------------------------
 * search = {lower0, upper0, lower1, upper1}: the arguments of searchOffset<2> (correlation/correlation_base.h:299-315),
 * axis 0 = rows, axis 1 = columns; dH = upper0 - lower0 + 1, dW = upper1 - lower1 + 1. */

/* unfoldBased2dDisparityCostVolume<matchFunc, float, float, 2, dDir, float> (correlation/cross_correlations.h:794-822, over
 * aggregateCost :308-371): img_l, img_r (H,W) f32 of equal size -> out (H,W,dH,dW) f32 (the reference's strides are
 * {W*dW*dH, 1, W*dW, W}; any strides are accepted). Cell (i,j,a,b) compares the source window at (i,j) with the target
 * window at (i+a+lower0, j+b+lower1), zero features outside the target image; the direction only picks the source
 * image. Float matching functions (CC..ZSAD); sizes that differ or an empty window -> STEVI_E_SHAPE (reference: empty array). */
STEVI_API int32_t stevi_cost_volume_2d(stevi_ctx* ctx, int32_t matchfunc, const stevi_array* img_l, const stevi_array* img_r,
							 int32_t h_radius, int32_t v_radius, const int32_t search[4], int32_t direction, stevi_array* out);
/* The same with a semantics word, as stevi_cost_volume has: without STEVI_EXACT_SUMS (what stevi_cost_volume_2d passes) windows up
 * to 9 columns wide take the O(1)-per-cell box sums, one pass of the 1-D box kernel per row offset (1e-5 of the largest magnitude
 * against the reference, 1e-4 absolute for the normalised correlations); with it every cell is summed over its F features in the
 * reference's channel order (2e-6). ZSAD (not a box filter) and wider windows always take the second form. */
STEVI_API int32_t stevi_cost_volume_2d_ex(stevi_ctx* ctx, int32_t matchfunc, const stevi_array* img_l, const stevi_array* img_r,
								int32_t h_radius, int32_t v_radius, const int32_t search[4], int32_t direction, int32_t semantics,
								stevi_array* out);

/* extractSelected2dIndex<strategy, T_CV> (correlation/correlation_base.h:466-508): vol (H,W,D1,D2) f32 -> out (H,W,2) i32;
 * row-major scan of (d1,d2) with `<=` / `>=`: the LAST extremum wins, a NaN at (0,0) sticks. */
STEVI_API int32_t stevi_wta_2d(stevi_ctx* ctx, const stevi_array* vol, int32_t strategy, stevi_array* out);

/* selected2dIndexToDisp<disp_t> (correlation/correlation_base.h:534-555): disp(i,j,k) = idx(i,j,k) + lower offset of axis k */
STEVI_API int32_t stevi_index2d_to_disp(stevi_ctx* ctx, const stevi_array* idx, const int32_t search[4], stevi_array* out);

/* truncatedBidirectionaCostVolume<float> (correlation/correlation_base.h:677-725): cv (H,W,D1,D2) f32, idx (H,W,2) i32 ->
 * out (H,W,2*r0+1,2*r1+1) f32 = cv(i, j, idx0+a-r0, idx1+b-r1), NaN outside the search window (:718). radius0 / radius1 are
 * the reference's uint8_t arguments (0..255; its default uint8_t(-1) is 255): r0 = radius0 < 1 ? 1 : radius0 and
 * r1 = radius1 < 1 ? r0 : radius1 (:689-695). */
STEVI_API int32_t stevi_truncated_cost_volume_2d(stevi_ctx* ctx, const stevi_array* cv, const stevi_array* idx, int32_t radius0,
									   int32_t radius1, stevi_array* out);

/* refineDisp2dCostInterpolation<kernel, isotropHypothesis> (correlation/cost_based_refinement.h:165-369): tcv (H,W,T0,T1) f32,
 * raw_disparity (H,W,2) i32 -> out (H,W,2) f32 = raw + (delta0, delta1). Isotropic: one refineCostTriplet per axis through
 * the centre cell. Anisotropic: the per-column / per-row optima of the rows and columns next to the centre give two ridge
 * lines whose intersection is the estimate (:269-353); whether the volume holds scores is read off the centre pixel of the
 * image (:184-206). |delta| > 1 or NaN on either axis resets both to 0 (:356-360). STEVI_E_SHAPE where the reference returns
 * an empty array (a side even or < 3, :180-182). */
STEVI_API int32_t stevi_refine_disp_2d_cost_interp(stevi_ctx* ctx, const stevi_array* tcv, const stevi_array* raw_disparity,
										 int32_t kernel, int32_t isotropy, stevi_array* out);

/* refineDisp2dCostPatchInterpolation<kernel> (correlation/cost_based_refinement.h:371-433 over refineCostPatch :71-126):
 * least-squares quadric through the 3x3 cells around the centre of tcv (H,W,T0,T1) f32 (their logarithms for the Gaussian
 * kernel) and its stationary point; out (H,W,2) f32 = raw + (delta0, delta1), both reset to 0 when either is beyond one cell
 * or NaN. kernel = STEVI_PARABOLA | STEVI_GAUSSIAN (the reference static_asserts the same, :82). The reference solves the
 * normal equations with Eigen's general inverse; here they are solved in closed form (same estimator, rounding differs:
 * compare with a relative tolerance of 1e-4 on well-conditioned patches). */
STEVI_API int32_t stevi_refine_disp_2d_cost_patch_interp(stevi_ctx* ctx, const stevi_array* tcv, const stevi_array* raw_disparity,
											   int32_t kernel, stevi_array* out);

/* selectedCost<T_CV> (correlation/correlation_base.h:557-577): out(i,j) = cv(i,j,idx(i,j)); out (H,W) of cv's dtype */
STEVI_API int32_t stevi_selected_cost(stevi_ctx* ctx, const stevi_array* cv, const stevi_array* idx, stevi_array* out);

/* truncatedCostVolume<float, dir, sdir> (correlation/correlation_base.h:579-674): cv (H,W,D) f32, idx (H,W) i32 ->
 * out (H,W,2r+1) f32, or (H,W,4r+1) for STEVI_TCV_BOTH; NaN where the disparity is out of range or the matching window
 * (h_radius, v_radius) leaves the image (:606-609). */
STEVI_API int32_t stevi_truncated_cost_volume(stevi_ctx* ctx, const stevi_array* cv, const stevi_array* idx, int32_t h_radius,
									int32_t v_radius, int32_t cost_vol_radius, int32_t direction, int32_t tcv_direction,
									stevi_array* out);

/* extractInBoundDomain<float, dir, sdir> (correlation/correlation_base.h:727-822): the same tests as a 1/0 mask;
 * `width` is the disparity search width. out (H,W,2r+1 | 4r+1) f32. */
STEVI_API int32_t stevi_in_bound_domain(stevi_ctx* ctx, const stevi_array* idx, int32_t width, int32_t h_radius, int32_t v_radius,
							  int32_t cost_vol_radius, int32_t direction, int32_t tcv_direction, stevi_array* out);

/* ---- guided cost-volume aggregation ------------------------------------------------------ */

/* variableCostVolumeAggregation<float, float> (correlation/guided_cost_filtering.h:102-176): cv (H,W,D) f32, windows
 * (H,W,2v+1,2h+1) f32 (the per-pixel weights computeAggregationWindows :26-100 builds from the caller's functors) ->
 * out (H,W,D) f32 = sum of w * cv over the in-image part of the pixel's window / sum of those w. Sums run in the reference's
 * order (window rows outer, columns inner). STEVI_E_SHAPE where the reference returns an empty array (:108-130: windows
 * of another image size, even or empty window sides). out must not alias cv. */
STEVI_API int32_t stevi_variable_cost_volume_aggregation(stevi_ctx* ctx, const stevi_array* cv, const stevi_array* windows,
											   stevi_array* out);

/* ---- cost-based subpixel refinement ------------------------------------------------------------------------ */

/* refineDispCostInterpolation<kernel> (correlation/cost_based_refinement.h:128-163): tcv (H,W,2r+1) f32, idx (H,W) i32
 * -> out (H,W) f32 = idx + delta(c[r-1], c[r], c[r+1]) (refineCostTriplet, :43-69). STEVI_E_SHAPE where the reference
 * returns an empty array (depth even or < 3, :141-143). */
STEVI_API int32_t stevi_refine_disp_cost_interp(stevi_ctx* ctx, const stevi_array* tcv, const stevi_array* idx, int32_t kernel,
									  stevi_array* out);

/* truncatedCostVolume<Same>(cost_vol_radius = 1) followed by refineDispCostInterpolation in one pass (three gathers per
 * pixel): cv (H,W,D) f32, idx (H,W) i32 -> out (H,W) f32; tcv_out (H,W,3) f32 optional (NULL = not materialised). */
STEVI_API int32_t stevi_refine_from_cost_volume(stevi_ctx* ctx, const stevi_array* cv, const stevi_array* idx, int32_t h_radius,
									  int32_t v_radius, int32_t kernel, stevi_array* tcv_out, stevi_array* out);

/* ---- feature-domain subpixel refinement (correlation/image_based_refinement.h) --------------------
 * feat_l (H,Wl,F), feat_r (H,Wr,F) f32: the feature volumes the refinement works on (the reference's callers pass the zero-mean
 * and / or normalised volumes of refined*FeatureVol, :1692-2424; any strides). The source is feat_r for RIGHT_TO_LEFT, feat_l
 * otherwise; idx / out have the source's (H,W). The per-pixel fits are the reference's estimators — NCC, ZNCC:
 * Optimization::sphericalAffineBestApproximation (optimization/sphericaloptimization.h:35-63); SSD, ZSSD:
 * affineBestL2Approximation (l2optimization.h:48-63); SAD, ZSAD: affineBestL1Approximation (l1optimization.h:293-310) — solved
 * from the fp64 Gram matrix of the vectors instead of Eigen's float QR: same solution, other rounding (compare with a
 * tolerance: 1e-3 px). CC, ZCC and the other functions have no barycentric approximation in the reference either
 * (matching_costs.h:442-457, :505-520): STEVI_E_UNSUPPORTED. */
typedef enum { STEVI_QUEEN = 0, STEVI_ROOK = 1, STEVI_BISHOP = 2 } stevi_contiguity; /* utils/contiguity.h:31-35 */

/* refineBarycentricDisp<matchFunc, dDir> (correlation/image_based_refinement.h:184-355): the source vector is fitted by the two
 * affine pairs (target(jd), target(jd+1)) and (target(jd-1), target(jd)); a coefficient inside (0,1) whose interpolated vector
 * compares better than the winner moves the disparity by that fraction. idx (H,W) i32 -> out (H,W) f32; pixels whose target
 * column jd = j +- idx is within one column of the border keep idx (:214). */
STEVI_API int32_t stevi_refine_barycentric_disp(stevi_ctx* ctx, int32_t matchfunc, const stevi_array* feat_l, const stevi_array* feat_r,
									  const stevi_array* idx, int32_t direction, stevi_array* out);

/* refineBarycentricSymmetricDisp<matchFunc, refineRadius, dDir> (:31-105): one fit over the 2*radius+1 target vectors around jd,
 * out = idx + sum(coeff_p * p) when that is below one pixel; idx at the borders, at index 0 and at disp_width - 1 (:64-69).
 * radius 1..3. SAD / ZSAD: STEVI_E_UNSUPPORTED (with more than two columns the reference runs a randomised descent,
 * l1optimization.h:98-252, whose result depends on Eigen's Random()). */
STEVI_API int32_t stevi_refine_barycentric_symmetric_disp(stevi_ctx* ctx, int32_t matchfunc, int32_t radius, const stevi_array* feat_l,
												const stevi_array* feat_r, const stevi_array* idx, int32_t disp_width, int32_t direction,
												stevi_array* out);

/* refineBarycentric2dDisp<matchFunc, contiguity, dDir> (:675-880; symmetric = 0): for each of the four quadrants the source is
 * fitted by the corner directions of the contiguity (utils/contiguity.h:139-164) and the winner; the best-scoring admissible
 * fit moves the 2-D disparity. refineBarycentricSymmetric2dDisp (:1068-1241; symmetric = 1): one fit over all the directions of
 * the contiguity (:98-135). feat_l, feat_r (H,W,F) of one size, idx (H,W,2) i32 (row, column displacement), search = {lower0,
 * upper0, lower1, upper1} as for stevi_cost_volume_2d -> out (H,W,2) f32. NCC, ZNCC, SSD, ZSSD. */
STEVI_API int32_t stevi_refine_barycentric_2d_disp(stevi_ctx* ctx, int32_t matchfunc, int32_t contiguity, int32_t symmetric,
										 const stevi_array* feat_l, const stevi_array* feat_r, const stevi_array* idx,
										 const int32_t search[4], int32_t direction, stevi_array* out);

/* refineCostSymmetricDisp<matchFunc, dDir> (:1591-1690): the parabola through the winner's three costs, averaged with the
 * parabola through the comparisons of the half-pixel source vector with the three target vectors. cv (H,W,D) f32 is the
 * cost volume idx was taken from. CC .. ZSAD. */
STEVI_API int32_t stevi_refine_cost_symmetric_disp(stevi_ctx* ctx, int32_t matchfunc, const stevi_array* feat_l, const stevi_array* feat_r,
										 const stevi_array* idx, const stevi_array* cv, int32_t direction, stevi_array* out);

/* ---- on-demand costs and PatchMatch ----------------------------------------------------------------
 * feat_source (H,W,F), feat_target (Ht,Wt,F) f32: raw feature volumes (what unfold gives); the function's processing (zero mean,
 * normalisation) is applied inside, like the reference does. search_dims = 1: search along the columns, search = {lower, upper},
 * Ht must equal H; search_dims = 2: search = {lower0, upper0, lower1, upper1} (rows, columns). A cell (pixel, disparity) has no
 * value when the disparity is outside the search range or the target pixel outside the target image. CC .. ZSAD. */

/* patchMatch<matchFunc, searchSpaceDim>(feature_vol_s, feature_vol_t, searchOffset, nIter, nRandomSearch, initializer)
 * (correlation/patchmatch.h:498-556 over patchMatchImpl :449-496, patchMatchPropagate :371-446, patchMatchSearch :233-369,
 * patchMatchTestCost :166-231) -> disp (H,W,search_dims) i32. Per iteration: the reference's row scans and column scans in the
 * iteration's direction (each pixel tests its predecessor's disparity and keeps it when it compares at least as well), then
 * n_random_search random candidates per pixel pulled towards its current disparity; an iteration without a change ends the run.
 * init (nullable): the initial map (the reference's `initializer`); NULL = uniform random disparities in the search range
 * (randomDispInit, :57-160). With n_random_search = 0 and an init the result equals the reference's bit for bit; the random draws
 * come from a counter-based generator keyed by `seed` (the reference seeds from std::random_device: its own runs differ). */
STEVI_API int32_t stevi_patch_match(stevi_ctx* ctx, int32_t matchfunc, const stevi_array* feat_source, const stevi_array* feat_target,
						  int32_t search_dims, const int32_t* search, int32_t n_iter, int32_t n_random_search, uint64_t seed,
						  const stevi_array* init, stevi_array* disp);

/* OnDemand*CostVolume::truncatedCostVolume(disp, radius) (correlation/on_demand_cost_volume.h:187-310): the costs of the cells
 * around each pixel's disparity, evaluated where asked for — no cost volume exists. disp (H,W) or (H,W,1) i32 for a 1-D search,
 * (H,W,2) for a 2-D one -> out (H,W,2r+1) / (H,W,2r+1,2r+1) f32; cells without a value hold defaultCvValForMatchFunc
 * (matching_costs.h:706-713: FLT_MAX for costs, FLT_MIN for scores). The features are processed once, as the on-demand volume's
 * constructor does (:63-68). Like the reference (:224-241) the window is shifted by the lower search offset when that is not 0. */
STEVI_API int32_t stevi_on_demand_truncated_cost_volume(stevi_ctx* ctx, int32_t matchfunc, const stevi_array* feat_source,
											  const stevi_array* feat_target, int32_t search_dims, const int32_t* search,
											  const stevi_array* disp, int32_t radius, stevi_array* out);

/* ---- coarse-to-fine matching ---------------------------------------------------------------------------------- */

/* meanFilter2D(h_radius, v_radius, in_data) (correlation/correlation_base.h:881-938 for (H,W), :941-1004 for (H,W,C) images):
 * img u8|f32 -> out (H,W) f32, the mean over the (2*h_radius+1) x (2*v_radius+1) (x C) box, computed like the reference by running
 * sums in float along the rows and then along the columns (bit-identical to it: every value carries the rounding of the chain
 * before it). The reference writes only the interior [v_radius, H-v_radius) x [h_radius, W-h_radius) of its result and leaves the
 * rest uninitialised; here the border is 0. STEVI_E_SHAPE when the window does not fit the image (out of bounds in the reference). */
STEVI_API int32_t stevi_mean_filter_2d(stevi_ctx* ctx, const stevi_array* img, int32_t h_radius, int32_t v_radius, stevi_array* out);
/* sigmaFilter(h_radius, v_radius, mean, in_data) (correlation/correlation_base.h:1006-1077): out(i,j) = sqrtf(sum over the box
 * (rows, columns, channels in that order) of (in - mean(i,j))^2) on the interior, 0 on the border; product and sum rounded
 * separately (what g++ emits for three of the four instantiations; its contracted 3-D f32 one differs by <= 2 ulp). */
STEVI_API int32_t stevi_sigma_filter(stevi_ctx* ctx, const stevi_array* mean, const stevi_array* img, int32_t h_radius,
						   int32_t v_radius, stevi_array* out);

/* Interpolation::averagePoolingDownsample<T_I, float> (interpolation/downsampling.h:67-118) with
 * DownSampleWindows(win_h, win_v): img (H,W) u8|f32 -> out (ceil(H/win_v), ceil(W/win_h)) f32, mean over the in-image
 * part of each window. */
STEVI_API int32_t stevi_avgpool_downsample(stevi_ctx* ctx, const stevi_array* img, int32_t win_h, int32_t win_v, stevi_array* out);

/* computeGuidedCV<matchFunc, ..., dDir, float> (correlation/hierarchical.h:73-233) fused with unfold +
 * getFeatureVolumeForMatchFunc of both images: img_l, img_r (H,W*) f32 (u8 for CC/SSD/SAD), guide (h_g,w_g) i32 ->
 * tcv (H,Wsource,2*radius+1) f32 (OffsetedCostVolume::truncated_cost_volume) and disp (H,Wsource) i32
 * (OffsetedCostVolume::disp_estimate). radius <= 4. STEVI_E_SHAPE for axes of fewer than 2 pixels (Q7). */
STEVI_API int32_t stevi_guided_cost_volume(stevi_ctx* ctx, int32_t matchfunc, const stevi_array* img_l, const stevi_array* img_r,
								 int32_t h_radius, int32_t v_radius, const stevi_array* guide, int32_t upscale_disp_radius,
								 int32_t direction, stevi_array* tcv, stevi_array* disp);

/* computeGuidedCV<matchFunc, float, float, dDir, float>(feature_vol_l, feature_vol_r, disp_guide, upscale_disp_radius) with the
 * reference's own operands (correlation/hierarchical.h:73-233): feat_l (H,Wl,F), feat_r (H,Wr,F) f32 feature volumes as the caller
 * prepared them (getFeatureVolumeForMatchFunc of an unfold, a FeatureVolumePyramid level, ..), guide (h_g,w_g) i32 ->
 * tcv (H,Wsource,2*radius+1) f32, disp (H,Wsource) i32. Sums in the reference's channel order. */
STEVI_API int32_t stevi_guided_cost_volume_features(stevi_ctx* ctx, int32_t matchfunc, const stevi_array* feat_l, const stevi_array* feat_r,
										  const stevi_array* guide, int32_t upscale_disp_radius, int32_t direction,
										  stevi_array* tcv, stevi_array* disp);

/* Row-band form of stevi_guided_cost_volume for frames split across GPUs: img_l / img_r hold image
 * rows img_row0 .. img_row0 + rows-1 of an image that is img_height rows tall, guide holds guide rows guide_row0 .. of a
 * guide that is guide_height rows tall, and tcv / disp receive the rows out_row0 .. out_row0 + tcv.shape[0]-1. The guide
 * is upsampled in the GLOBAL coordinates i*(guide_height-1)/(img_height-1) (hierarchical.h:116), so the band results are
 * identical to the rows of the whole-frame call. The image band must carry v_radius halo rows on its inner edges. */
typedef struct { int32_t img_row0, img_height, out_row0, guide_row0, guide_height; } stevi_band;
STEVI_API int32_t stevi_guided_cost_volume_band(stevi_ctx* ctx, int32_t matchfunc, const stevi_array* img_l, const stevi_array* img_r,
									  int32_t h_radius, int32_t v_radius, const stevi_array* guide, int32_t upscale_disp_radius,
									  int32_t direction, const stevi_band* band, stevi_array* tcv, stevi_array* disp);

/* hiearchicalTruncatedCostVolume<matchFunc, depth, ...> (sic; correlation/hierarchical.h:235-314). h_radii / v_radii:
 * depth+1 entries, coarsest level first, like the reference's std::array arguments (NULL = h_radius / v_radius at every
 * level, the overload at :297-314). img_l, img_r (H,W) f32 -> tcv (H,W,2*radius+1) f32, disp (H,W) i32. */
STEVI_API int32_t stevi_hierarchical(stevi_ctx* ctx, int32_t matchfunc, int32_t depth, const stevi_array* img_l,
						   const stevi_array* img_r, const int32_t* h_radii, const int32_t* v_radii, int32_t h_radius,
						   int32_t v_radius, int32_t disp_width, int32_t upscale_disp_radius, int32_t direction,
						   stevi_array* tcv, stevi_array* disp);

/* ---- fused device-resident pipeline ------------------------------------------------------------------------ */

/* census cost volume -> SGM -> WTA -> disparity in one call, the sequence of
 * test/benchmarks/benchmarkCrossCorrelationAlgorithms.cpp:289-293 with matchFunc = CENSUS and strategy = Cost.
 * img_l, img_r (H,W) u8|f32; disp (H,W) i32 (required; a dense u8 / u16 array is accepted for RIGHT_TO_LEFT when disp_width fits
 * the type: a quarter / half of the bytes for callers that keep disparities narrow). cv_out (H,W,D) u16 and sgm_out (H,W,D) f32 are optional
 * (NULL = keep the volume in the ctx workspace / do not materialise it at all when the semantics allow).
 * Results are bit-identical to calling stevi_cost_volume, stevi_sgm, stevi_wta, stevi_index_to_disp in turn. */
STEVI_API int32_t stevi_census_sgm_pipeline(stevi_ctx* ctx, const stevi_array* img_l, const stevi_array* img_r,
								  int32_t h_radius, int32_t v_radius, int32_t disp_width, int32_t direction,
								  int32_t n_directions, float P1, float P2, float Pout, const stevi_margins* margins,
								  int32_t semantics, stevi_array* cv_out, stevi_array* sgm_out, stevi_array* disp);

/* float cost volume -> WTA -> truncated cost volume (radius 1) -> subpixel refinement in one call: the dense local
 * matching sequence of test/benchmarks/benchmarkCrossCorrelationAlgorithms.cpp:93-95 followed by
 * truncatedCostVolume + refineDispCostInterpolation (examples/stereo-match's refinement step). img_l, img_r (H,W*) f32
 * (u8 for CC/SSD/SAD); refined (H,Wsource) f32 is required; cv_out (H,W,D) f32, idx_out (H,W) i32 (raw winner indices)
 * and tcv_out (H,W,3) f32 are optional (NULL = kept in the workspace / not materialised). The winner follows the
 * function's extraction strategy (Score for the correlations, Cost otherwise). Results are identical to calling
 * stevi_cost_volume, stevi_wta, stevi_truncated_cost_volume and stevi_refine_disp_cost_interp in turn. */
STEVI_API int32_t stevi_local_matching_pipeline(stevi_ctx* ctx, int32_t matchfunc, const stevi_array* img_l, const stevi_array* img_r,
									  int32_t h_radius, int32_t v_radius, int32_t disp_width, int32_t direction,
									  int32_t kernel, stevi_array* cv_out, stevi_array* idx_out, stevi_array* tcv_out,
									  stevi_array* refined);

/* ---- batches of independent pairs over several contexts / devices ---------------------------------------------------- */

/* One stereo pair of a batch: images and result as for stevi_census_sgm_pipeline (host or device memory; page-locked host memory
 * lets the copies overlap the kernels). */
typedef struct { const stevi_array* img_l; const stevi_array* img_r; stevi_array* disp; } stevi_pair;

/* stevi_census_sgm_pipeline (disparity map only) over n_pairs independent pairs, the loop of
 * test/benchmarks/benchmarkCostVolumeWorkflow.cpp:175-260 with the calls left in flight. ctxs: n_ctx contexts, each with its own
 * stream (stevi_ctx_create(device, NULL, ..)); several contexts on one device = that many pairs in flight there (3-4 hide the
 * copies), contexts on different devices run side by side, one host thread per device inside the call. Pair k goes to device
 * group k mod (number of devices among ctxs). Returns when every result is in place; the first failure (its text in the
 * failing context's stevi_last_error) stops the batch. No collective is involved: the pairs are independent. */
STEVI_API int32_t stevi_batch_census_sgm(stevi_ctx* const* ctxs, int32_t n_ctx, const stevi_pair* pairs, int32_t n_pairs,
							   int32_t h_radius, int32_t v_radius, int32_t disp_width, int32_t direction, int32_t n_directions,
							   float P1, float P2, float Pout, const stevi_margins* margins, int32_t semantics);

/* ---- an entry point that allocates nothing ----------------------------------------------------------------------------- */

/* No counterpart in the reference: added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The census cost volume of a pair, result identical to stevi_cost_volume(ctx, STEVI_MF_CENSUS, ..): img_l (H,Wl), img_r (H,Wr)
 * u8|f32 (same dtype), out (H,Wsource,D) u16|i32|f32 — ALL THREE IN DEVICE MEMORY (STEVI_DEVICE; a host operand is refused with
 * STEVI_E_ARG). The scratch memory of the stage, the census words of the two images, is a static __device__ array of
 * STEVI_NOALLOC_SCRATCH_BYTES inside the library instead of the context's workspace: every GPU that loads the library holds it
 * until the process exits, and the call neither grows the workspace nor calls cudaMalloc / cudaFree / any allocating routine.
 * Limit: round256(H*Wl*nw*4) + round256(H*Wr*nw*4) <= STEVI_NOALLOC_SCRATCH_BYTES with nw = ((2*h_radius+1)*(2*v_radius+1)-1)/32+1
 * words per pixel (1280x720 at 9x7 fits, 1920x1080 fits up to 5x5 windows = one word; 1920x1080 at 9x7 does not) — larger pairs get STEVI_E_UNSUPPORTED and a message stating both
 * figures. A context with per-kernel timing switched on is refused (timing creates events).
 * All calls, from every context and stream of the process, share the one array per GPU; it keeps the last call's contents (the
 * part a call uses is cleared first, on the context's stream). Calls that overlap on the GPU — different streams without an
 * ordering between them — race on it: serialise them. Work is enqueued on the context's stream like every other entry point. */
#define STEVI_NOALLOC_SCRATCH_BYTES (16u << 20)
STEVI_API int32_t stevi_census_cost_volume_noalloc(stevi_ctx* ctx, const stevi_array* img_l, const stevi_array* img_r,
										 int32_t h_radius, int32_t v_radius, int32_t disp_width, int32_t direction,
										 int32_t semantics, stevi_array* out);

#ifdef __cplusplus
}
#endif
#endif
