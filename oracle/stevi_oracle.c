/* TEST INFRASTRUCTURE — NOT PRODUCT CODE.
 *
 * Plain-C CPU restatement of the algorithms on LibStevi's dense-stereo hot path, used only as
 * the checker by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
 * The shipped library (libstevi_b200/csrc) never links, loads or calls anything in this file.
 *
 * Parity status: PINNED. Every function here is checked in tests/test_oracle_pinning.py against outputs of the
 * reference's own headers compiled unmodified (oracle/_ref/libstevi_ref.so, built by oracle/Makefile from
 * /root/reference) and against golden vectors generated from that library (tests/golden/, generator committed as
 * tests/golden/make_golden.py), plus Qt-free restatements of the reference's own unit tests.
 *
 * All buffers are dense, C-order (last index fastest). File:line citations are relative to /root/reference.
 * Compiled with -ffp-contract=off: the float operation ORDER below follows the reference's loops.
 */
#define _GNU_SOURCE /* qsort_r */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>

/* enums mirror correlation/matching_costs.h:38-53, correlation_base.h:31-45, cost_based_refinement.h:30-35 */
enum { MF_CC = 0, MF_NCC = 1, MF_SSD = 2, MF_SAD = 3, MF_ZCC = 4, MF_ZNCC = 5, MF_ZSSD = 6, MF_ZSAD = 7,
	   MF_HAMMING = 10, MF_CENSUS = 11 };
enum { STRAT_COST = 0, STRAT_SCORE = 1 };
enum { DIR_L2R = 0, DIR_R2L = 1 };
enum { K_EQUIANGULAR = 0, K_PARABOLA = 1, K_GAUSSIAN = 2 };

static int mf_zero_mean(int mf) { return mf == MF_ZCC || mf == MF_ZNCC || mf == MF_ZSSD || mf == MF_ZSAD; }
static int mf_normalized(int mf) { return mf == MF_NCC || mf == MF_ZNCC; }
/* extraction strategy per MatchingFunctionTraits (matching_costs.h:419-685): correlations are scores */
static int mf_strategy(int mf) { return (mf == MF_CC || mf == MF_NCC || mf == MF_ZCC || mf == MF_ZNCC) ? STRAT_SCORE : STRAT_COST; }

int orc_matchfunc_strategy(int mf) { return mf_strategy(mf); }

/* ---- unfold.h:248-291 (2-D image, auto padding, Rotate0): out(i,j,(2hr+1)*k+l) = img(i-vr+k, j-hr+l), 0 outside */
#define DEF_UNFOLD(NAME, T) \
int NAME(const T* img, int h, int w, int hr, int vr, T* out) { \
	int hs = 2 * hr + 1, vs = 2 * vr + 1, F = hs * vs; \
	for (int i = 0; i < h; i++) for (int j = 0; j < w; j++) \
		for (int k = 0; k < vs; k++) for (int l = 0; l < hs; l++) { \
			int ii = i - vr + k, jj = j - hr + l; \
			T v = (ii >= 0 && ii < h && jj >= 0 && jj < w) ? img[(size_t)ii * w + jj] : (T)0; \
			out[((size_t)i * w + j) * F + hs * k + l] = v; \
		} \
	return 0; \
}
DEF_UNFOLD(orc_unfold_f32, float)
DEF_UNFOLD(orc_unfold_u8, uint8_t)

/* ---- census.h:69-115 on an (h,w,F) feature volume: bit b of word n = (feat[0] > feat[1+32n+b]).
 * Q1: a word is stored only when all 32 bits were filled (census.h:103-108); trailing partial word stays 0
 * (zero-initialised allocation in the stand-in array type). nwords = (F-1)/32 + 1 (census.h:79). */
#define DEF_CENSUS_FEATS(NAME, T) \
static void NAME(const T* fv, int h, int w, int F, uint32_t* out) { \
	int nw = (F - 1) / 32 + 1; \
	memset(out, 0, (size_t)h * w * nw * sizeof(uint32_t)); \
	for (size_t p = 0; p < (size_t)h * w; p++) { \
		const T* f = fv + p * F; \
		T ref = f[0]; \
		uint32_t d = 0; int b = 0, ch = 0; \
		for (int c = 1; c < F; c++) { \
			uint32_t g = (ref > f[c]) ? 1u : 0u; \
			d |= g << b; b++; \
			if (b >= 32) { out[p * nw + ch] = d; ch++; d = 0; b = 0; } \
		} \
	} \
}
DEF_CENSUS_FEATS(census_feats_f32, float)
DEF_CENSUS_FEATS(census_feats_u8, uint8_t)

/* census.h:69-115 on a ready-made (h,w,F) float feature volume */
int orc_census_features_f32(const float* fv, int h, int w, int F, uint32_t* out) { census_feats_f32(fv, h, w, F, out); return 0; }

/* unfold.h:171-191 */
static int channel_from_cord(int k, int l, int ch, int hs, int vs, int nch, int orientation) {
	switch (orientation) {
	case 0: return nch * hs * k + nch * l + ch;
	case 1: return nch * vs * (hs - l - 1) + nch * k + ch;
	case 2: return nch * hs * (vs - k - 1) + nch * (hs - l - 1) + ch;
	default: return nch * vs * l + nch * (vs - k - 1) + ch;
	}
}

/* unfold.h:248-344 with explicit padding {left, top, right, bottom} (pad[0] < 0: automatic = the radii) and orientation;
 * c == 0: 2-D image. out (h', w', F*max(c,1)); the output size comes back in oh / ow (out may be NULL to query it). */
int orc_unfold_general_f32(const float* img, int h, int w, int c, int hr, int vr, const int* pad, int orientation, float* out, int* oh, int* ow) {
	int hs = 2 * hr + 1, vs = 2 * vr + 1, nch = c == 0 ? 1 : c;
	int pl = pad[0] < 0 ? hr : pad[0], pt = pad[0] < 0 ? vr : pad[1], pr = pad[0] < 0 ? hr : pad[2], pb = pad[0] < 0 ? vr : pad[3];
	int H = h - vs + pt + pb + 1, W = w - hs + pl + pr + 1, F = hs * vs * nch;
	*oh = H; *ow = W;
	if (!out) return 0;
	for (int k = 0; k < vs; k++) for (int l = 0; l < hs; l++) for (int ch = 0; ch < nch; ch++) {
		int cc = channel_from_cord(k, l, ch, hs, vs, nch, orientation);
		for (int i = 0; i < H; i++) for (int j = 0; j < W; j++) {
			int ii = i + k - pt, jj = j + l - pl;
			out[((size_t)i * W + j) * F + cc] = (ii >= 0 && ii < h && jj >= 0 && jj < w) ? img[((size_t)ii * w + jj) * nch + ch] : 0.f;
		}
	}
	return 0;
}

/* unfold.h:36-137 (UnFoldCompressor) + :346-471 (compressed unfold). Labels > 0, features in ascending label order,
 * weight = (float)(1./count), accumulation in index order (per feature: row-major over the mask). */
int orc_unfold_compressed_f32(const int* mask, int mh, int mw, const float* img, int h, int w, int c, const int* pad, float* out, int* oh, int* ow, int* nf) {
	int v_off = mh / 2, h_off = mw / 2, minH = 0, maxH = 0, minW = 0, maxW = 0, nlab = 0;
	int* labels = (int*)malloc(sizeof(int) * mh * mw);
	for (int i = 0; i < mh; i++) for (int j = 0; j < mw; j++) {
		int f = mask[i * mw + j];
		if (f <= 0) continue;
		if (i - v_off < minH) minH = i - v_off;
		if (i - v_off > maxH) maxH = i - v_off;
		if (j - h_off < minW) minW = j - h_off;
		if (j - h_off > maxW) maxW = j - h_off;
		int known = 0;
		for (int n = 0; n < nlab; n++) if (labels[n] == f) known = 1;
		if (!known) labels[nlab++] = f;
	}
	for (int a = 0; a < nlab; a++) for (int b = a + 1; b < nlab; b++) if (labels[b] < labels[a]) { int t = labels[a]; labels[a] = labels[b]; labels[b] = t; }
	int nch = c == 0 ? 1 : c, cw = maxW - minW + 1, chh = maxH - minH + 1, top = -minH, left = -minW;
	int pl = pad[0] < 0 ? -minW : pad[0], pt = pad[0] < 0 ? -minH : pad[1], pr = pad[0] < 0 ? maxW : pad[2], pb = pad[0] < 0 ? maxH : pad[3];
	int H = h - chh + pt + pb + 1, W = w - cw + pl + pr + 1, F = nlab * nch;
	*oh = H; *ow = W; *nf = nlab;
	if (!out) { free(labels); return 0; }
	memset(out, 0, sizeof(float) * (size_t)H * W * F);
	for (int f = 0; f < nlab; f++) {
		int count = 0;
		for (int n = 0; n < mh * mw; n++) if (mask[n] == labels[f]) count++;
		float wgt = (float)(1. / count);
		for (int mi = 0; mi < mh; mi++) for (int mj = 0; mj < mw; mj++) {
			if (mask[mi * mw + mj] != labels[f]) continue;
			int k = mi - v_off, l = mj - h_off;
			for (int ch = 0; ch < nch; ch++) for (int i = 0; i < H; i++) for (int j = 0; j < W; j++) {
				int ii = i + k + top - pt, jj = j + l + left - pl;
				float v = (ii >= 0 && ii < h && jj >= 0 && jj < w) ? img[((size_t)ii * w + jj) * nch + ch] : 0.f;
				out[((size_t)i * W + j) * F + ch * nlab + f] += wgt * v;
			}
		}
	}
	free(labels);
	return 0;
}

int orc_census_nwords(int hr, int vr) { int F = (2 * hr + 1) * (2 * vr + 1); return (F - 1) / 32 + 1; }

/* census.h:117-131 censusTransform2D = unfold + censusFeatures */
int orc_census_u8(const uint8_t* img, int h, int w, int hr, int vr, uint32_t* out) {
	int F = (2 * hr + 1) * (2 * vr + 1);
	if (F <= 1) return 1; /* census.h:75-77: empty result */
	uint8_t* fv = (uint8_t*)malloc((size_t)h * w * F);
	orc_unfold_u8(img, h, w, hr, vr, fv);
	census_feats_u8(fv, h, w, F, out);
	free(fv);
	return 0;
}
int orc_census_f32(const float* img, int h, int w, int hr, int vr, uint32_t* out) {
	int F = (2 * hr + 1) * (2 * vr + 1);
	if (F <= 1) return 1;
	float* fv = (float*)malloc((size_t)h * w * F * sizeof(float));
	orc_unfold_f32(img, h, w, hr, vr, fv);
	census_feats_f32(fv, h, w, F, out);
	free(fv);
	return 0;
}

/* ---- cross_correlations.h:645-722 for float images: in place on an (h,w,F) unfolded volume.
 * mean = (sum_c v) * (1/F)            correlation_base.h:1116-1130
 * norm = sqrtf(sum_c (v-mean)^2)      cross_correlations.h:61-104   (or sum v^2 for NCC, :149-191)
 * feat = (v-mean)/norm                cross_correlations.h:452      (no epsilon: flat patch -> NaN, Q6) */
static void features_inplace(int mf, float* fv, size_t npix, int F) {
	int zm = mf_zero_mean(mf), nm = mf_normalized(mf);
	if (!zm && !nm) return;
	float scale = 1.f / (float)F;
	for (size_t p = 0; p < npix; p++) {
		float* f = fv + p * F;
		float mean = 0.f;
		if (zm) {
			for (int c = 0; c < F; c++) mean += f[c];
			mean *= scale;
		}
		if (nm) {
			float n = 0.f;
			for (int c = 0; c < F; c++) { float t = zm ? (f[c] - mean) : f[c]; n += t * t; }
			n = sqrtf(n);
			for (int c = 0; c < F; c++) f[c] = (zm ? (f[c] - mean) : f[c]) / n;
		} else {
			for (int c = 0; c < F; c++) f[c] = f[c] - mean;
		}
	}
}

int orc_feature_volume_f32(int mf, const float* img, int h, int w, int hr, int vr, float* out) {
	int F = (2 * hr + 1) * (2 * vr + 1);
	orc_unfold_f32(img, h, w, hr, vr, out);
	features_inplace(mf, out, (size_t)h * w, F);
	return 0;
}

/* ---- matching_costs.h:59-78 / :100-116 / :136-156: sequential accumulation over channels in float */
static float compare_f32(int mf, const float* s, const float* t, int F) {
	float acc = 0.f;
	switch (mf) {
	case MF_CC: case MF_NCC: case MF_ZCC: case MF_ZNCC:
		for (int c = 0; c < F; c++) acc += s[c] * t[c];
		break;
	case MF_SSD: case MF_ZSSD:
		for (int c = 0; c < F; c++) { float d = s[c] - t[c]; acc += d * d; }
		break;
	default: /* SAD, ZSAD */
		for (int c = 0; c < F; c++) acc += fabsf(s[c] - t[c]);
		break;
	}
	return acc;
}

/* ---- cross_correlations.h:194-249 aggregateCost + :740-765 unfoldBasedCostVolume, float images.
 * RightToLeft: source = right image, target = left image, target column j + d (correlation_base.h:829-852,
 * cross_correlations.h:204,235); LeftToRight: source = left, target = right, column j - d.
 * Target features are 0 when the column is outside the TARGET image (valueOrAlt, :235).
 * out is (h, w_source, D) C-order. Returns 1 (reference returns an empty array) when heights differ (:751). */
int orc_cost_volume_f32(int mf, const float* l, const float* r, int h, int wl, int wr,
						int hr, int vr, int D, int dir, float* out) {
	int F = (2 * hr + 1) * (2 * vr + 1);
	float* fl = (float*)malloc((size_t)h * wl * F * sizeof(float));
	float* fr = (float*)malloc((size_t)h * wr * F * sizeof(float));
	orc_feature_volume_f32(mf, l, h, wl, hr, vr, fl);
	orc_feature_volume_f32(mf, r, h, wr, hr, vr, fr);
	const float* src = dir == DIR_R2L ? fr : fl; int ws = dir == DIR_R2L ? wr : wl;
	const float* tgt = dir == DIR_R2L ? fl : fr; int wt = dir == DIR_R2L ? wl : wr;
	int sgn = dir == DIR_R2L ? 1 : -1;
	float* zero = (float*)calloc((size_t)F, sizeof(float));
	#pragma omp parallel for
	for (int i = 0; i < h; i++) for (int j = 0; j < ws; j++) for (int d = 0; d < D; d++) {
		int jt = j + sgn * d;
		const float* t = (jt >= 0 && jt < wt) ? tgt + ((size_t)i * wt + jt) * F : zero;
		out[((size_t)i * ws + j) * D + d] = compare_f32(mf, src + ((size_t)i * ws + j) * F, t, F);
	}
	free(zero); free(fl); free(fr);
	return 0;
}

/* ---- feature-level forms ------------------------------------------------------------------------------------------------
 * cross_correlations.h:645-722 getFeatureVolumeForMatchFunc on a ready-made (h, w, F) float volume (copy, then in place) */
int orc_features_for_matchfunc_f32(int mf, const float* fv, int h, int w, int F, float* out) {
	memcpy(out, fv, (size_t)h * w * F * sizeof(float));
	features_inplace(mf, out, (size_t)h * w, F);
	return 0;
}

/* cross_correlations.h:194-249 aggregateCost and its searchOffset<1> overload (:251-308) on processed feature volumes
 * fl (h, wl, F), fr (h, wr, F): cell d compares the source vector with the target vector `lower + d` columns away
 * (idx2disp, correlation_base.h:288-409), zero features outside the target image. out (h, w_source, upper - lower + 1). */
int orc_aggregate_cost_f32(int mf, const float* fl, const float* fr, int h, int wl, int wr, int F, int lower, int upper, int dir, float* out) {
	const float* src = dir == DIR_R2L ? fr : fl; int ws = dir == DIR_R2L ? wr : wl;
	const float* tgt = dir == DIR_R2L ? fl : fr; int wt = dir == DIR_R2L ? wl : wr;
	int sgn = dir == DIR_R2L ? 1 : -1, D = upper - lower + 1;
	if (D < 1) return 1;
	float* zero = (float*)calloc((size_t)F, sizeof(float));
	#pragma omp parallel for
	for (int i = 0; i < h; i++) for (int j = 0; j < ws; j++) for (int d = 0; d < D; d++) {
		int jt = j + sgn * (d + lower);
		const float* t = (jt >= 0 && jt < wt) ? tgt + ((size_t)i * wt + jt) * F : zero;
		out[((size_t)i * ws + j) * D + d] = compare_f32(mf, src + ((size_t)i * ws + j) * F, t, F);
	}
	free(zero);
	return 0;
}

/* ---- cross_correlations.h:308-371 aggregateCost over searchOffset<2> + :794-822 unfoldBased2dDisparityCostVolume, float
 * images of equal size (the reference returns an empty array otherwise, :804-810). so = {lower0, upper0, lower1, upper1}:
 * axis 0 = rows, axis 1 = columns (correlation_base.h:299-315). The target pixel of cell (i, j, dh, dw) is
 * (i + dh + lower0, j + dw + lower1) for BOTH directions (:357) - the direction only picks which image is the source
 * (condImgRef) - and its features are 0 outside the target image (valueOrAlt). out (h, w, dH, dW) C-order; returns 1 for an
 * empty search window (:336-338). */
int orc_cost_volume_2d_f32(int mf, const float* l, const float* r, int h, int w, int hr, int vr, const int* so, int dir, float* out) {
	int dH = so[1] - so[0] + 1, dW = so[3] - so[2] + 1;
	if (dH <= 0 || dW <= 0) return 1;
	int F = (2 * hr + 1) * (2 * vr + 1);
	float* fl = (float*)malloc((size_t)h * w * F * sizeof(float));
	float* fr = (float*)malloc((size_t)h * w * F * sizeof(float));
	orc_feature_volume_f32(mf, l, h, w, hr, vr, fl);
	orc_feature_volume_f32(mf, r, h, w, hr, vr, fr);
	const float* src = dir == DIR_R2L ? fr : fl;
	const float* tgt = dir == DIR_R2L ? fl : fr;
	float* zero = (float*)calloc((size_t)F, sizeof(float));
	#pragma omp parallel for
	for (int i = 0; i < h; i++) for (int j = 0; j < w; j++) for (int a = 0; a < dH; a++) for (int b = 0; b < dW; b++) {
		int it = i + a + so[0], jt = j + b + so[2];
		const float* t = (it >= 0 && it < h && jt >= 0 && jt < w) ? tgt + ((size_t)it * w + jt) * F : zero;
		out[(((size_t)i * w + j) * dH + a) * dW + b] = compare_f32(mf, src + ((size_t)i * w + j) * F, t, F);
	}
	free(zero); free(fl); free(fr);
	return 0;
}

/* correlation_base.h:466-508 extractSelected2dIndex: the scan starts from cell (0,0) and visits it again, `<=` / `>=`, so
 * the LAST extremum in (d1, d2) row-major order wins and a NaN at (0,0) sticks. out (h, w, 2). */
int orc_wta_2d_f32(const float* v, int h, int w, int d1, int d2, int strategy, int32_t* out) {
	for (size_t p = 0; p < (size_t)h * w; p++) {
		const float* c = v + p * d1 * d2;
		float best = c[0]; int32_t b1 = 0, b2 = 0;
		for (int a = 0; a < d1; a++) for (int b = 0; b < d2; b++) {
			float x = c[a * d2 + b];
			if (strategy == STRAT_COST ? (x <= best) : (x >= best)) { best = x; b1 = a; b2 = b; }
		}
		out[2 * p] = b1; out[2 * p + 1] = b2;
	}
	return 0;
}

/* correlation_base.h:534-555 selected2dIndexToDisp: index + lower offset per axis */
int orc_index2d_to_disp(const int32_t* idx, int h, int w, const int* so, int32_t* out) {
	for (size_t p = 0; p < (size_t)h * w; p++) { out[2 * p] = idx[2 * p] + so[0]; out[2 * p + 1] = idx[2 * p + 1] + so[2]; }
	return 0;
}

/* Q2 (cross_correlations.h:235-236): every TARGET feature passes through `float t` before being narrowed back
 * to the feature type. For uint32 census words that is a round-to-nearest-even to 24 significant bits; a word
 * >= 0xFFFFFF80 rounds to 2^32 whose conversion back is undefined behaviour in C++ — the scalar x86-64 result
 * (cvttss2si on a 64-bit register, then truncation) is 0, which is what we document and reproduce. */
uint32_t orc_round_through_float(uint32_t w) {
	float t = (float)w;
	if (t >= 4294967296.0f) return 0u;
	return (uint32_t)t;
}

static int popc32(uint32_t x) { return __builtin_popcount(x); }

static int census_cv(const uint32_t* cl, const uint32_t* cr, int h, int wl, int wr, int nw, int D, int dir, uint16_t* out) {
	const uint32_t* src = dir == DIR_R2L ? cr : cl; int ws = dir == DIR_R2L ? wr : wl;
	const uint32_t* tgt = dir == DIR_R2L ? cl : cr; int wt = dir == DIR_R2L ? wl : wr;
	int sgn = dir == DIR_R2L ? 1 : -1;
	#pragma omp parallel for
	for (int i = 0; i < h; i++) for (int j = 0; j < ws; j++) for (int d = 0; d < D; d++) {
		int jt = j + sgn * d;
		uint16_t acc = 0; /* hamming_cv_t, matching_costs.h:234,252-263 */
		for (int n = 0; n < nw; n++) {
			uint32_t t = (jt >= 0 && jt < wt) ? orc_round_through_float(tgt[((size_t)i * wt + jt) * nw + n]) : 0u;
			acc = (uint16_t)(acc + popc32(src[((size_t)i * ws + j) * nw + n] ^ t));
		}
		out[((size_t)i * ws + j) * D + d] = acc;
	}
	return 0;
}

/* unfoldBasedCostVolume<CENSUS, uint8/float images, hamming_cv_t> */
int orc_census_cost_volume_u8(const uint8_t* l, const uint8_t* r, int h, int wl, int wr,
							  int hr, int vr, int D, int dir, uint16_t* out) {
	int nw = orc_census_nwords(hr, vr);
	uint32_t* cl = (uint32_t*)malloc((size_t)h * wl * nw * 4);
	uint32_t* cr = (uint32_t*)malloc((size_t)h * wr * nw * 4);
	orc_census_u8(l, h, wl, hr, vr, cl);
	orc_census_u8(r, h, wr, hr, vr, cr);
	census_cv(cl, cr, h, wl, wr, nw, D, dir, out);
	free(cl); free(cr);
	return 0;
}
int orc_census_cost_volume_f32(const float* l, const float* r, int h, int wl, int wr,
							   int hr, int vr, int D, int dir, uint16_t* out) {
	int nw = orc_census_nwords(hr, vr);
	uint32_t* cl = (uint32_t*)malloc((size_t)h * wl * nw * 4);
	uint32_t* cr = (uint32_t*)malloc((size_t)h * wr * nw * 4);
	orc_census_f32(l, h, wl, hr, vr, cl);
	orc_census_f32(r, h, wr, hr, vr, cr);
	census_cv(cl, cr, h, wl, wr, nw, D, dir, out);
	free(cl); free(cr);
	return 0;
}

/* ---- sgm.h:57-155 direction traits: {stepsVertical[2], stepsHorizontal[2]} alternating with the step parity */
static const int SGM_STEPS[16][4] = {
	/* Up2Down */ {1, 1, 0, 0}, /* Down2Up */ {-1, -1, 0, 0}, /* Left2Right */ {0, 0, 1, 1}, /* Right2Left */ {0, 0, -1, -1},
	/* UpLeft2DownRight */ {1, 1, 1, 1}, /* DownRight2UpLeft */ {-1, -1, -1, -1},
	/* UpRight2DownLeft */ {1, 1, -1, -1}, /* DownLeft2UpRight */ {-1, -1, 1, 1},
	/* UpLeft2Right */ {0, 1, 1, 1}, /* DownRight2Left */ {0, -1, -1, -1}, /* UpRight2Left */ {0, 1, -1, -1}, /* DownLeft2Right */ {0, -1, 1, 1},
	/* UpLeft2Down */ {1, 1, 0, 1}, /* DownRight2Up */ {-1, -1, 0, -1}, /* UpRight2Down */ {1, 1, 0, -1}, /* DownLeft2Up */ {-1, -1, 0, 1}};
/* call order of sgmCostVolume, sgm.h:379-401 */
static const int SGM_ORDER[16] = {0, 1, 2, 3, 4, 5, 6, 7, 12, 13, 14, 15, 8, 9, 10, 11};

/* sgm.h:186-311 traverseLine. cv (h,w,D) C-order float; sgm (h,w,D) accumulated in place.
 * The reference's inner O(D^2) loop over od is restated in O(D): subtracting/adding a constant is monotone in
 * IEEE arithmetic, so max_od (p[od] - P) == (max_od p[od]) - P exactly. Non-finite previous scores are skipped
 * (`std::isfinite(p_score)` tests at :224,:243,:264,:281). */
static void sgm_traverse_line(int dirIdx, int strategy, int si, int sj, const float* cv, float* sgm,
							  int h, int w, int D, float P1, float P2, float Pout, const int* mg, int corrected) {
	int left = mg[0], top = mg[1], right = mg[2], bottom = mg[3];
	const int* st = SGM_STEPS[dirIdx];
	float* prev = (float*)calloc((size_t)D, sizeof(float));
	float* act = (float*)malloc((size_t)D * sizeof(float));
	float* pre = (float*)malloc((size_t)(D + 2) * sizeof(float)); /* prefix extremum over finite p[0..k] */
	float* suf = (float*)malloc((size_t)(D + 2) * sizeof(float)); /* suffix extremum over finite p[k..D-1] */
	int c = 0;
	for (int i = si, j = sj; (i >= top && i < h - bottom) && (j >= left && j < w - right); i += st[c % 2], j += st[2 + c % 2], c++) {
		const float* cc = cv + ((size_t)i * w + j) * D;
		if (strategy == STRAT_SCORE) {
			float max_p = -INFINITY;
			for (int d = 0; d < D; d++) if (prev[d] > max_p && isfinite(prev[d])) max_p = prev[d];
			float run = -INFINITY;
			for (int d = 0; d < D; d++) { if (isfinite(prev[d]) && prev[d] > run) run = prev[d]; pre[d] = run; }
			run = -INFINITY;
			for (int d = D - 1; d >= 0; d--) { if (isfinite(prev[d]) && prev[d] > run) run = prev[d]; suf[d] = run; }
			for (int nd = 0; nd < D; nd++) {
				float m = -INFINITY, t;
				/* candidates in the order the reference meets them: od < nd-1 (P2), nd-1 (P1), nd, nd+1 (P1), od > nd+1 (P2) */
				if (nd - 2 >= 0) { t = pre[nd - 2] - P2; if (t > m && isfinite(t)) m = t; }
				if (nd - 1 >= 0) { t = prev[nd - 1] - P1; if (t > m && isfinite(t)) m = t; }
				t = prev[nd]; if (t > m && isfinite(t)) m = t;
				if (nd + 1 < D) { t = prev[nd + 1] - P1; if (t > m && isfinite(t)) m = t; }
				if (nd + 2 < D) { t = suf[nd + 2] - P2; if (t > m && isfinite(t)) m = t; }
				if (j + nd >= w) m -= Pout;
				act[nd] = cc[nd];
				if (isfinite(m) && isfinite(max_p)) act[nd] += m - max_p;
			}
		} else if (corrected) {
			/* CORRECTED semantics (not the reference's behaviour): `min_a_cost = p_score`, i.e. the textbook
			 * recurrence, the mirror image of the Score branch */
			float min_p = INFINITY;
			for (int d = 0; d < D; d++) if (prev[d] < min_p && isfinite(prev[d])) min_p = prev[d];
			float run = INFINITY;
			for (int d = 0; d < D; d++) { if (isfinite(prev[d]) && prev[d] < run) run = prev[d]; pre[d] = run; }
			run = INFINITY;
			for (int d = D - 1; d >= 0; d--) { if (isfinite(prev[d]) && prev[d] < run) run = prev[d]; suf[d] = run; }
			for (int nd = 0; nd < D; nd++) {
				float m = INFINITY, t;
				if (nd - 2 >= 0) { t = pre[nd - 2] + P2; if (t < m && isfinite(t)) m = t; }
				if (nd - 1 >= 0) { t = prev[nd - 1] + P1; if (t < m && isfinite(t)) m = t; }
				t = prev[nd]; if (t < m && isfinite(t)) m = t;
				if (nd + 1 < D) { t = prev[nd + 1] + P1; if (t < m && isfinite(t)) m = t; }
				if (nd + 2 < D) { t = suf[nd + 2] + P2; if (t < m && isfinite(t)) m = t; }
				if (j + nd >= w) m += Pout;
				act[nd] = cc[nd];
				if (isfinite(m) && isfinite(min_p)) act[nd] += m - min_p;
			}
		} else {
			float min_p = INFINITY;
			for (int d = 0; d < D; d++) if (prev[d] < min_p && isfinite(prev[d])) min_p = prev[d];
			/* Q3 (sgm.h:281-283): the Cost branch assigns min_a_cost = c_score (not p_score) as soon as one
			 * penalised previous score is finite and below the running value. */
			int any_finite = 0; /* p_score + P is finite iff p_score is (P1, P2 finite) */
			for (int od = 0; od < D; od++) if (isfinite(prev[od])) any_finite = 1;
			for (int nd = 0; nd < D; nd++) {
				float c_score = cc[nd];
				/* the od loop can only ever store c_score: the first finite penalised score is < +inf, and
				 * any later successful test stores c_score again */
				float min_a = any_finite ? c_score : INFINITY;
				if (j + nd >= w) min_a += Pout;
				act[nd] = c_score;
				if (isfinite(min_a) && isfinite(min_p)) act[nd] += min_a - min_p;
			}
		}
		float* s = sgm + ((size_t)i * w + j) * D;
		for (int d = 0; d < D; d++) s[d] += act[d] - cc[d];
		float* tmp = prev; prev = act; act = tmp;
	}
	free(prev); free(act); free(pre); free(suf);
}

/* sgm.h:162-184 + :313-356: start edges. 0 = NoStart, 1 = ZeroPos, 2 = EndPos. */
static int sgm_start_pos(int s0, int s1) {
	if (s0 == 0 && s1 == 0) return 0;
	if (s0 >= 0 && s1 >= 0) return 1;
	return 2;
}

static void sgm_add_direction(int dirIdx, int strategy, const float* cv, float* sgm, int h, int w, int D,
							  float P1, float P2, float Pout, const int* mg, int corrected) {
	int left = mg[0], top = mg[1], right = mg[2], bottom = mg[3];
	const int* st = SGM_STEPS[dirIdx];
	int colStart = sgm_start_pos(st[0], st[1]); /* from stepsVertical */
	int rowStart = sgm_start_pos(st[2], st[3]); /* from stepsHorizontal */
	int row_sj = -1;
	if (rowStart != 0) { /* lines starting on the left (or, Q4, one past the right) column: sgm.h:329-341 */
		/* CORRECTED: EndPos starts on the last valid index instead of one past it */
		int sj = (rowStart == 1) ? left : w - right - (corrected ? 1 : 0);
		row_sj = sj;
		/* Q8: lines of the eight extra directions of the 16-direction set (steps alternating between two moves) started
		 * on neighbouring pixels share pixels, so the reference's own `omp parallel for` over them races on sgm_cv;
		 * the sequential result is the defined one: those directions run on one thread here. */
		#pragma omp parallel for if (dirIdx < 8)
		for (int si = top; si < h - bottom; si++)
			sgm_traverse_line(dirIdx, strategy, si, sj, cv, sgm, h, w, D, P1, P2, Pout, mg, corrected);
	}
	if (colStart != 0) { /* lines starting on the top (or, Q4, one past the bottom) row: sgm.h:343-354 */
		int si = (colStart == 1) ? top : h - bottom - (corrected ? 1 : 0);
		#pragma omp parallel for if (dirIdx < 8)
		for (int sj = left; sj < w - right; sj++) {
			if (corrected && rowStart != 0 && sj == row_sj) continue; /* CORRECTED: the corner line is started once */
			sgm_traverse_line(dirIdx, strategy, si, sj, cv, sgm, h, w, D, P1, P2, Pout, mg, corrected);
		}
	}
}

/* sgm.h:360-404. margins = {left, top, right, bottom} (utils/margins.h:52). out (h,w,D) float.
 * corrected != 0: the STEVI_CORRECTED semantics of the product (our definition, not the reference's behaviour): textbook Cost
 * recurrence, every direction live, corner lines once; directions in the reference's order. */
int orc_sgm_f32_ex(const float* cv, int h, int w, int D, int ndir, int strategy,
				   float P1, float P2, float Pout, const int* margins, int corrected, float* out) {
	if (ndir != 4 && ndir != 8 && ndir != 16) return 2;
	memcpy(out, cv, (size_t)h * w * D * sizeof(float));
	for (int k = 0; k < ndir; k++)
		sgm_add_direction(SGM_ORDER[k], strategy, cv, out, h, w, D, P1, P2, Pout, margins, corrected);
	return 0;
}
int orc_sgm_f32(const float* cv, int h, int w, int D, int ndir, int strategy,
				float P1, float P2, float Pout, const int* margins, float* out) {
	return orc_sgm_f32_ex(cv, h, w, D, ndir, strategy, P1, P2, Pout, margins, 0, out);
}
int orc_sgm_u16(const uint16_t* cv, int h, int w, int D, int ndir, int strategy,
				float P1, float P2, float Pout, const int* margins, float* out) {
	size_t n = (size_t)h * w * D;
	float* f = (float*)malloc(n * sizeof(float)); /* static_cast<float>(cv_base.value(...)), sgm.h:233,272 */
	for (size_t k = 0; k < n; k++) f[k] = (float)cv[k];
	int r = orc_sgm_f32(f, h, w, D, ndir, strategy, P1, P2, Pout, margins, out);
	free(f);
	return r;
}

/* ---- correlation_base.h:427-464 WTA: `<=` / `>=` so the LAST extremum wins (Q5) */
#define DEF_WTA(NAME, T) \
int NAME(const T* v, int h, int w, int D, int strategy, int32_t* out) { \
	for (size_t p = 0; p < (size_t)h * w; p++) { \
		const T* c = v + p * D; \
		T best = c[0]; int32_t bd = 0; \
		for (int d = 1; d < D; d++) { \
			if (strategy == STRAT_COST) { if (c[d] <= best) { best = c[d]; bd = d; } } \
			else { if (c[d] >= best) { best = c[d]; bd = d; } } \
		} \
		out[p] = bd; \
	} \
	return 0; \
}
DEF_WTA(orc_wta_f32, float)
DEF_WTA(orc_wta_u16, uint16_t)

/* correlation_base.h:511-532 */
int orc_index_to_disp(const int32_t* idx, int h, int w, int dir, int offset, int32_t* out) {
	int sgn = dir == DIR_R2L ? 1 : -1;
	for (size_t p = 0; p < (size_t)h * w; p++) out[p] = sgn * idx[p] + offset;
	return 0;
}

/* correlation_base.h:557-577 */
int orc_selected_cost_f32(const float* cv, const int32_t* idx, int h, int w, int D, float* out) {
	for (size_t p = 0; p < (size_t)h * w; p++) out[p] = cv[p * D + (uint32_t)idx[p]];
	return 0;
}

/* correlation_base.h:579-674. tdir: 0 Same, 1 Reversed, 2 Both. out (h,w,2r+1) or (h,w,4r+1) for Both. */
int orc_truncated_cost_volume_f32(const float* cv, const int32_t* idx, int h, int w, int D,
								  int hr, int vr, int r, int dir, int tdir, float* out) {
	int T = tdir == 2 ? 4 * r + 1 : 2 * r + 1;
	int sgn = dir == DIR_R2L ? -1 : 1;
	for (int i = 0; i < h; i++) for (int j = 0; j < w; j++) {
		float* o = out + ((size_t)i * w + j) * T;
		int vbad = (i < vr) || (i + vr >= h);
		for (int d = 0; d <= 2 * r; d++) {
			int p = idx[(size_t)i * w + j] + d - r;
			int jp = j + sgn * (d - r);
			int pbad = (p < 0) || (p >= D);
			int same_bad = pbad || j < hr || j + p + hr >= w || vbad;
			if (tdir == 0) {
				o[d] = same_bad ? nanf("") : cv[((size_t)i * w + j) * D + p];
			} else if (tdir == 1) {
				int lo = jp < j ? jp : j, hi = jp > j ? jp : j;
				int bad = pbad || lo < hr || hi + hr >= w || vbad;
				o[d] = bad ? nanf("") : cv[((size_t)i * w + jp) * D + p];
			} else {
				int d_d = 2 * d, d_r = 2 * d + 1;
				if (d == r) jp = -1;
				if (d > r) { d_d -= 1; d_r -= 1; }
				int lo = jp < j ? jp : j, hi = jp > j ? jp : j;
				int rbad = pbad || lo < hr || hi + hr >= w || vbad;
				o[d_d] = same_bad ? nanf("") : cv[((size_t)i * w + j) * D + p];
				o[d_r] = rbad ? nanf("") : cv[((size_t)i * w + jp) * D + p];
			}
		}
	}
	return 0;
}

/* correlation_base.h:727-822: the tests of truncatedCostVolume as a 1/0 mask; `width` = disparity search width */
int orc_in_bound_domain_f32(const int32_t* idx, int h, int w, int width, int hr, int vr, int r, int dir, int tdir, float* out) {
	int T = tdir == 2 ? 4 * r + 1 : 2 * r + 1;
	int sgn = dir == DIR_R2L ? -1 : 1;
	for (int i = 0; i < h; i++) for (int j = 0; j < w; j++) {
		float* o = out + ((size_t)i * w + j) * T;
		int vbad = (i < vr) || (i + vr >= h);
		for (int d = -r; d <= r; d++) {
			int p = idx[(size_t)i * w + j] + d;
			int jp = j + sgn * d;
			int pbad = (p < 0) || (p >= width);
			int same_bad = pbad || j < hr || j + p + hr >= w || vbad;
			if (tdir == 0) { o[d + r] = same_bad ? 0.f : 1.f; continue; }
			if (tdir == 2 && d == 0) jp = -1;
			int lo = jp < j ? jp : j, hi = jp > j ? jp : j;
			int rbad = pbad || lo < hr || hi + hr >= w || vbad;
			if (tdir == 1) { o[d + r] = rbad ? 0.f : 1.f; continue; }
			int d_d = 2 * (d + r), d_r = 2 * (d + r) + 1;
			if (d > 0) { d_d -= 1; d_r -= 1; }
			if (d_r < T) o[d_r] = rbad ? 0.f : 1.f; /* r == 0 would write past the array in the reference */
			o[d_d] = same_bad ? 0.f : 1.f;
		}
	}
	return 0;
}

/* cost_based_refinement.h:43-69 */
float orc_refine_cost_triplet(int kernel, float cm1, float c0, float c1) {
	if (kernel == K_EQUIANGULAR) {
		float alpha = copysignf(1.f, c0 - cm1);
		alpha *= fmaxf(fabsf(c0 - cm1), fabsf(c1 - c0));
		return (c1 - cm1) / (2 * alpha);
	} else if (kernel == K_PARABOLA) {
		return (cm1 - c1) / (2 * (c1 - 2 * c0 + cm1));
	}
	return (logf(cm1) - logf(c1)) / (2 * (logf(c1) - 2 * logf(c0) + logf(cm1)));
}

/* cost_based_refinement.h:128-163. Returns 1 (reference: empty array) if the depth is even or < 3. */
int orc_refine_disp_cost_interp(const float* tcv, const int32_t* idx, int h, int w, int T, int kernel, float* out) {
	int r = (T - 1) / 2;
	if (r < 1 || 2 * r + 1 != T) return 1;
	for (size_t p = 0; p < (size_t)h * w; p++) {
		const float* c = tcv + p * T;
		out[p] = (float)idx[p] + orc_refine_cost_triplet(kernel, c[r - 1], c[r], c[r + 1]);
	}
	return 0;
}

/* correlation_base.h:677-725 truncatedBidirectionaCostVolume: the (2*r0+1) x (2*r1+1) cells around the 2-D winner, NaN where an
 * index leaves the search window (valueOrAlt, :718). The radii are uint8_t in the reference: 0 becomes 1 for axis 0 (:689-691)
 * and "same as axis 0" for axis 1 (:693-695). cv (h, w, d1, d2), idx (h, w, 2), out (h, w, 2*r0+1, 2*r1+1), all C-order. */
int orc_truncated_cost_volume_2d_f32(const float* cv, const int32_t* idx, int h, int w, int d1, int d2, int radius0, int radius1, float* out) {
	int r0 = radius0 & 255, r1 = radius1 & 255;
	if (r0 < 1) r0 = 1;
	if (r1 < 1) r1 = r0;
	int n0 = 2 * r0 + 1, n1 = 2 * r1 + 1;
	for (size_t p = 0; p < (size_t)h * w; p++) {
		const float* c = cv + p * d1 * d2;
		float* o = out + p * n0 * n1;
		for (int a = 0; a < n0; a++) for (int b = 0; b < n1; b++) {
			int p0 = idx[2 * p] + a - r0, p1 = idx[2 * p + 1] + b - r1;
			o[a * n1 + b] = (p0 >= 0 && p0 < d1 && p1 >= 0 && p1 < d2) ? c[p0 * d2 + p1] : nanf("");
		}
	}
	return 0;
}

/* cost_based_refinement.h:208-246: last extremum (`<=` / `>=`) along axis 0 (fixed column `col` of axis 1) or along axis 1 */
static int tcv_arg_best(const float* c, int n, int stride, int score) {
	float best = score ? -INFINITY : INFINITY;
	int arg = 0;
	for (int a = 0; a < n; a++) {
		float x = c[a * stride];
		if (score ? (x >= best) : (x <= best)) { best = x; arg = a; }
	}
	return arg;
}

/* cost_based_refinement.h:165-369 refineDisp2dCostInterpolation<kernel, isotropy>. tcv (h, w, n0, n1), raw (h, w, 2),
 * out (h, w, 2) C-order. Returns 1 (reference: empty array, :180-182) when a side is even or < 3. Whether the volume holds
 * scores is guessed from the CENTRE pixel of the image (:184-206): any of its four neighbours below the centre cell. The
 * isotropic estimate is one triplet per axis (:256-267); the anisotropic one (:269-353) fits the ridge of the per-row and
 * per-column optima through the three rows / columns around the centre and intersects the two lines. Estimates beyond one
 * cell, or NaN, are dropped (:356-360). */
int orc_refine_disp_2d_cost_interp(const float* tcv, const int32_t* raw, int h, int w, int n0, int n1, int kernel, int anisotropic, float* out) {
	int r0 = (n0 - 1) / 2, r1 = (n1 - 1) / 2;
	if (r0 < 1 || r1 < 1 || 2 * r0 + 1 != n0 || 2 * r1 + 1 != n1) return 1;
	int score = 0;
	if (h > 0 && w > 0) {
		const float* c = tcv + ((size_t)(h / 2) * w + (w / 2)) * n0 * n1;
		float v0 = c[r0 * n1 + r1];
		if (v0 > c[(r0 + 1) * n1 + r1]) score = 1;
		if (v0 > c[(r0 - 1) * n1 + r1]) score = 1;
		if (v0 > c[r0 * n1 + r1 + 1]) score = 1;
		if (v0 > c[r0 * n1 + r1 - 1]) score = 1;
	}
	#pragma omp parallel for
	for (int i = 0; i < h; i++) for (int j = 0; j < w; j++) {
		size_t p = (size_t)i * w + j;
		const float* c = tcv + p * n0 * n1;
#define TC(a, b) c[(a) * n1 + (b)]
		float delta0, delta1;
		float d0c = orc_refine_cost_triplet(kernel, TC(r0 - 1, r1), TC(r0, r1), TC(r0 + 1, r1));
		float d1c = orc_refine_cost_triplet(kernel, TC(r0, r1 - 1), TC(r0, r1), TC(r0, r1 + 1));
		if (!anisotropic) {
			delta0 = d0c; delta1 = d1c;
		} else {
			float delta0_0 = d0c, delta0_2 = d0c, delta1_0 = d1c, delta1_2 = d1c;
			int a = tcv_arg_best(c + (r1 - 1), n0, n1, score);
			if (a > 0 && a < n0 - 1) delta0_0 = (float)(a - r0) + orc_refine_cost_triplet(kernel, TC(a - 1, r1 - 1), TC(a, r1 - 1), TC(a + 1, r1 - 1));
			a = tcv_arg_best(c + (r1 + 1), n0, n1, score);
			if (a > 0 && a < n0 - 1) delta0_2 = (float)(a - r0) + orc_refine_cost_triplet(kernel, TC(a - 1, r1 + 1), TC(a, r1 + 1), TC(a + 1, r1 + 1));
			float a0 = (delta0_2 - delta0_0) / 2;
			float b0 = (delta0_0 + d0c + delta0_2) / 3;
			a = tcv_arg_best(c + (r0 - 1) * n1, n1, 1, score);
			if (a > 0 && a < n1 - 1) delta1_0 = (float)(a - r1) + orc_refine_cost_triplet(kernel, TC(r0 - 1, a - 1), TC(r0 - 1, a), TC(r0 - 1, a + 1));
			a = tcv_arg_best(c + (r0 + 1) * n1, n1, 1, score);
			if (a > 0 && a < n1 - 1) delta1_2 = (float)(a - r1) + orc_refine_cost_triplet(kernel, TC(r0 + 1, a - 1), TC(r0 + 1, a), TC(r0 + 1, a + 1));
			float a1 = (delta1_2 - delta1_0) / 2;
			float b1 = (delta1_0 + d1c + delta1_2) / 3;
			/* the reference's Release build (g++ -O3 on an FMA target: -ffp-contract=fast is the C++ default) contracts these
			 * three products; fmaf states that explicitly so the restatement is bit-identical to oracle/_ref */
			float den = fmaf(-a0, a1, 1.f);
			delta0 = fmaf(a0, b1, b0) / den;
			delta1 = fmaf(a1, b0, b1) / den;
		}
#undef TC
		if (fabsf(delta0) > 1 || fabsf(delta1) > 1 || isnan(delta0) || isnan(delta1)) { delta0 = 0; delta1 = 0; }
		out[2 * p] = (float)raw[2 * p] + delta0;
		out[2 * p + 1] = (float)raw[2 * p + 1] + delta1;
	}
	return 0;
}

/* cost_based_refinement.h:71-126 refineCostPatch + :371-433 refineDisp2dCostPatchInterpolation. PARITY UNPINNED AT THE BIT
 * LEVEL: the reference evaluates ((A^T A)^-1 A^T) L and a 2x2 inverse with Eigen, which is not in this image (the stand-in
 * headers are parse-level only), so the compiled reference cannot run this function. This restates the same estimator in
 * double precision with the normal equations solved in closed form (on the 3x3 grid they decouple into row, column and
 * diagonal sums); it is pinned by the reference's own known-answer test (testCostRefinement.cpp:168-227) and by a float64
 * numpy least-squares solve in tests/. kernel: K_PARABOLA or K_GAUSSIAN (logarithms of the costs). Returns 1 where the
 * reference returns an empty array (:386-388), 2 for the Equiangular kernel (static_assert :82). */
int orc_refine_disp_2d_cost_patch_interp(const float* tcv, const int32_t* raw, int h, int w, int n0, int n1, int kernel, float* out) {
	int r0 = (n0 - 1) / 2, r1 = (n1 - 1) / 2;
	if (r0 < 1 || r1 < 1 || 2 * r0 + 1 != n0 || 2 * r1 + 1 != n1) return 1;
	if (kernel != K_PARABOLA && kernel != K_GAUSSIAN) return 2;
	for (size_t p = 0; p < (size_t)h * w; p++) {
		const float* c = tcv + p * n0 * n1;
		double L[3][3];
		for (int a = 0; a < 3; a++) for (int b = 0; b < 3; b++) {
			float x = c[(r0 - 1 + a) * n1 + (r1 - 1 + b)];
			L[a][b] = kernel == K_GAUSSIAN ? (double)logf(x) : (double)x;
		}
		double rm = L[0][0] + L[0][1] + L[0][2], rz = L[1][0] + L[1][1] + L[1][2], rp = L[2][0] + L[2][1] + L[2][2];
		double cm = L[0][0] + L[1][0] + L[2][0], cz = L[0][1] + L[1][1] + L[2][1], cp = L[0][2] + L[1][2] + L[2][2];
		double qa = (rm + rp - 2 * rz) / 6, qg = (cm + cp - 2 * cz) / 6, qb = (L[0][0] + L[2][2] - L[0][2] - L[2][0]) / 4;
		double qd = (rp - rm) / 6, qe = (cp - cm) / 6, det = 4 * qa * qg - qb * qb;
		float delta0 = (float)((qb * qe - 2 * qg * qd) / det), delta1 = (float)((qb * qd - 2 * qa * qe) / det);
		if (fabsf(delta0) > 1 || fabsf(delta1) > 1 || isnan(delta0) || isnan(delta1)) { delta0 = 0; delta1 = 0; }
		out[2 * p] = (float)raw[2 * p] + delta0;
		out[2 * p + 1] = (float)raw[2 * p + 1] + delta1;
	}
	return 0;
}

/* guided_cost_filtering.h:102-176 variableCostVolumeAggregation<float, float>: every cell of the volume becomes the weighted
 * mean of its neighbours at the same disparity, with the pixel's own window of weights clipped to the image (:137-142); window
 * rows outer, columns inner (:151-161). `acc += w*cv` is a separate multiply and add: the reference's Release build does not
 * contract this one (checked against oracle/_ref, bit-identical). Returns 1 where the reference returns an empty array (:108-130). */
int orc_variable_cost_volume_aggregation_f32(const float* cv, int h, int w, int d, const float* win, int hw, int ww, int n0, int n1, float* out) {
	if (hw != h || ww != w || n0 <= 0 || n1 <= 0 || n0 % 2 != 1 || n1 % 2 != 1) return 1;
	int vr = n0 / 2, hr = n1 / 2;
	#pragma omp parallel for
	for (int i = 0; i < h; i++) {
		int lo_i = -vr > -i ? -vr : -i, hi_i = vr < h - 1 - i ? vr : h - 1 - i;
		for (int j = 0; j < w; j++) {
			int lo_j = -hr > -j ? -hr : -j, hi_j = hr < w - 1 - j ? hr : w - 1 - j;
			const float* wp = win + (((size_t)i * w + j) * n0 + vr) * n1 + hr;
			for (int c = 0; c < d; c++) {
				float acc = 0, weight = 0;
				for (int di = lo_i; di <= hi_i; di++) for (int dj = lo_j; dj <= hi_j; dj++) {
					float x = wp[di * n1 + dj];
					acc += x * cv[((size_t)(i + di) * w + (j + dj)) * d + c];
					weight += x;
				}
				out[((size_t)i * w + j) * d + c] = acc / weight;
			}
		}
	}
	return 0;
}

/* interpolation/downsampling.h:67-118 with DownSampleWindows(2): out (ceil(h/2), ceil(w/2)); the mean is over
 * the in-image part of each 2x2 window (offsets are 0 for window 2: hRem/2 == 0). Accumulates in float, in
 * row-major window order, then divides by the count. */
int orc_avgpool2_f32(const float* img, int h, int w, float* out) {
	int nh = (h + 1) / 2, nw = (w + 1) / 2;
	for (int i = 0; i < nh; i++) for (int j = 0; j < nw; j++) {
		float val = 0.f; int count = 0;
		for (int dv = 0; dv < 2; dv++) for (int dh = 0; dh < 2; dh++) {
			int pi = 2 * i + dv, pj = 2 * j + dh;
			if (pi < h && pj < w) { val += img[(size_t)pi * w + pj]; count++; }
		}
		val /= count;
		out[(size_t)i * nw + j] = val;
	}
	return 0;
}

/* ---- hierarchical.h:73-233 computeGuidedCV on float feature volumes (h,w,F) (source width ws, target width wt).
 * guide is (hg, wg) int32. tcv (h, ws, 2r+1) C-order; disp (h, ws). */
static void guided_cv(int mf, const float* fl, const float* fr, int h, int wl, int wr, int F,
					  const int32_t* guide, int hg, int wg, int r, int dir, float* tcv, int32_t* disp) {
	const float* src = dir == DIR_R2L ? fr : fl; int ws = dir == DIR_R2L ? wr : wl;
	const float* tgt = dir == DIR_R2L ? fl : fr; int wt = dir == DIR_R2L ? wl : wr;
	int dirSign = dir == DIR_R2L ? 1 : -1;
	int T = 2 * r + 1;
	int cost = mf_strategy(mf) == STRAT_COST;
	float* zero = (float*)calloc((size_t)F, sizeof(float));
	#pragma omp parallel for
	for (int i = 0; i < h; i++) {
		float v_pos = (float)(i * (hg - 1)) / (h - 1); /* Q7: int product, then float division (:116) */
		int v0 = (int)floorf(v_pos), v1 = (int)ceilf(v_pos);
		if (v0 == v1) v1 += 1;
		if (v1 == hg) { v0 -= 1; v1 -= 1; }
		for (int j = 0; j < ws; j++) {
			float h_pos = (float)(j * (wg - 1)) / (ws - 1);
			int h0 = (int)floorf(h_pos), h1 = (int)ceilf(h_pos);
			if (h0 == h1) h1 += 1;
			if (h1 == wg) { h0 -= 1; h1 -= 1; }
			float id = 0.f;
			id += (v_pos - v0) * (h_pos - h0) * guide[(size_t)v1 * wg + h1];
			id += (v1 - v_pos) * (h_pos - h0) * guide[(size_t)v0 * wg + h1];
			id += (v_pos - v0) * (h1 - h_pos) * guide[(size_t)v1 * wg + h0];
			id += (v1 - v_pos) * (h1 - h_pos) * guide[(size_t)v0 * wg + h0];
			id *= 2;
			int d0 = dirSign * (int)roundf(id);
			const float* s = src + ((size_t)i * ws + j) * F;
			float* o = tcv + ((size_t)i * ws + j) * T;
			float score = cost ? INFINITY : -INFINITY;
			int d_r = d0;
			for (int dd = -r; dd <= r; dd++) {
				int jt = j + d0 + dd;
				const float* t = (jt >= 0 && jt < wt) ? tgt + ((size_t)i * wt + jt) * F : zero;
				float cmp = compare_f32(mf, s, t, F);
				if (cost) { if (cmp < score) { score = cmp; d_r = d0 + dd; } } /* strict: FIRST extremum wins (:179,:184) */
				else { if (cmp > score) { score = cmp; d_r = d0 + dd; } }
			}
			disp[(size_t)i * ws + j] = dirSign * d_r;
			/* after the reference's shift-and-refill (:194-226) cell dd always holds the comparison against
			 * target column j + d_r + dirSign*(dd - r); each cell is a pure function of the inputs */
			for (int dd = 0; dd < T; dd++) {
				int jt = j + d_r + dirSign * (dd - r);
				const float* t = (jt >= 0 && jt < wt) ? tgt + ((size_t)i * wt + jt) * F : zero;
				o[dd] = compare_f32(mf, s, t, F);
			}
		}
	}
	free(zero);
}

int orc_guided_cv_f32(int mf, const float* l, const float* r, int h, int wl, int wr, int hr, int vr,
					  const int32_t* guide, int hg, int wg, int radius, int dir, float* tcv, int32_t* disp) {
	int F = (2 * hr + 1) * (2 * vr + 1);
	float* fl = (float*)malloc((size_t)h * wl * F * sizeof(float));
	float* fr = (float*)malloc((size_t)h * wr * F * sizeof(float));
	orc_feature_volume_f32(mf, l, h, wl, hr, vr, fl);
	orc_feature_volume_f32(mf, r, h, wr, hr, vr, fr);
	guided_cv(mf, fl, fr, h, wl, wr, F, guide, hg, wg, radius, dir, tcv, disp);
	free(fl); free(fr);
	return 0;
}

/* computeGuidedCV on caller-supplied feature volumes, the reference's own signature (hierarchical.h:73-233) */
int orc_guided_cv_features_f32(int mf, const float* fl, const float* fr, int h, int wl, int wr, int F, const int32_t* guide, int hg, int wg,
							   int radius, int dir, float* tcv, int32_t* disp) {
	guided_cv(mf, fl, fr, h, wl, wr, F, guide, hg, wg, radius, dir, tcv, disp);
	return 0;
}

/* ---- hierarchical.h:235-314 (same radii at every level). tcv (h,w,2*radius+1), disp (h,w) at full resolution. */
static int hier_level(int mf, int depth, const float* l, const float* r, int h, int w, int hr, int vr, int D,
					  int radius, int dir, float* tcv, int32_t* disp) {
	int nh = (h + 1) / 2, nw = (w + 1) / 2;
	float* dl = (float*)malloc((size_t)nh * nw * sizeof(float));
	float* dr = (float*)malloc((size_t)nh * nw * sizeof(float));
	int32_t* guide = (int32_t*)malloc((size_t)nh * nw * sizeof(int32_t));
	orc_avgpool2_f32(l, h, w, dl);
	orc_avgpool2_f32(r, h, w, dr);
	int Dn = (D + 1) / 2;
	if (depth == 1) {
		float* cv = (float*)malloc((size_t)nh * nw * Dn * sizeof(float));
		orc_cost_volume_f32(mf, dl, dr, nh, nw, nw, hr, vr, Dn, dir, cv);
		orc_wta_f32(cv, nh, nw, Dn, mf_strategy(mf), guide);
		free(cv);
	} else {
		float* ptcv = (float*)malloc((size_t)nh * nw * (2 * radius + 1) * sizeof(float));
		hier_level(mf, depth - 1, dl, dr, nh, nw, hr, vr, Dn, radius, dir, ptcv, guide);
		free(ptcv);
	}
	orc_guided_cv_f32(mf, l, r, h, w, w, hr, vr, guide, nh, nw, radius, dir, tcv, disp);
	free(dl); free(dr); free(guide);
	return 0;
}

int orc_hierarchical_f32(int mf, int depth, const float* l, const float* r, int h, int w,
						 int hr, int vr, int D, int radius, int dir, float* tcv, int32_t* disp) {
	if (depth < 1) return 2;
	return hier_level(mf, depth, l, r, h, w, hr, vr, D, radius, dir, tcv, disp);
}

/* ---- correlation_base.h:881-1004 meanFilter2D on (h,w) (c == 0) or (h,w,c) images: running row sums, then running column
 * sums of those, then a division by the box size (times c). Each running sum is a chain of float additions in the reference's
 * order: mH(i,j) = mH(i,j-1) + (in(i,j+hr) - in(i,j-hr-1)), the difference taken in the image's own type first (int for u8
 * images, :908), one += per channel for 3-D images (:972-974). Only the interior [vr, h-vr) x [hr, w-hr) of `out` is defined by
 * the reference (the rest of its array is never written); it is left untouched here. Returns 1 if the window does not fit. */
#define ORC_MEAN_FILTER(NAME, T, DIFF_T) \
int NAME(const T* in, int h, int w, int c, int hr, int vr, float* out) { \
	if (w < 2 * hr + 1 || h < 2 * vr + 1) return 1; \
	int nc = c > 0 ? c : 1; \
	float* mh = (float*)malloc((size_t)h * w * sizeof(float)); \
	_Pragma("omp parallel for") \
	for (int i = 0; i < h; i++) { \
		const T* row = in + (size_t)i * w * nc; \
		float m = 0.f; \
		for (int j = 0; j <= 2 * hr; j++) for (int k = 0; k < nc; k++) m += (float)row[j * nc + k]; \
		mh[(size_t)i * w + hr] = m; \
		for (int j = hr + 1; j < w - hr; j++) { \
			for (int k = 0; k < nc; k++) m += (float)((DIFF_T)row[(j + hr) * nc + k] - (DIFF_T)row[(j - hr - 1) * nc + k]); \
			mh[(size_t)i * w + j] = m; \
		} \
	} \
	float box = (float)((size_t)(2 * vr + 1) * (2 * hr + 1) * nc); \
	_Pragma("omp parallel for") \
	for (int j = hr; j < w - hr; j++) { \
		float m = 0.f; \
		for (int i = 0; i <= 2 * vr; i++) m += mh[(size_t)i * w + j]; \
		out[(size_t)vr * w + j] = m / box; \
		for (int i = vr + 1; i < h - vr; i++) { \
			m += mh[(size_t)(i + vr) * w + j] - mh[(size_t)(i - vr - 1) * w + j]; \
			out[(size_t)i * w + j] = m / box; \
		} \
	} \
	free(mh); \
	return 0; \
}
ORC_MEAN_FILTER(orc_mean_filter_2d_f32, float, float)
ORC_MEAN_FILTER(orc_mean_filter_2d_u8, uint8_t, int)

/* ---- correlation_base.h:1006-1077 sigmaFilter: sqrtf of the sum over the window (rows, then columns, then channels) of
 * (in - mean(i,j))^2, interior only. Product and sum are rounded separately, as the source reads. Whether a Release build of the
 * reference contracts `s += tmp*tmp` into a fused multiply-add is the compiler's choice per instantiation: g++ 13 -O3 on an FMA
 * target (oracle/_ref) leaves the 2-D f32, 2-D u8 and 3-D u8 instantiations uncontracted (this restatement is bit-identical to
 * them) and contracts the 3-D f32 one (differs by <= 2 ulp). tests/test_oracle_pinning.py holds both statements. */
#define ORC_SIGMA_FILTER(NAME, T) \
int NAME(const float* mean, const T* in, int h, int w, int c, int hr, int vr, float* out) { \
	if (w < 2 * hr + 1 || h < 2 * vr + 1) return 1; \
	int nc = c > 0 ? c : 1; \
	_Pragma("omp parallel for") \
	for (int i = vr; i < h - vr; i++) for (int j = hr; j < w - hr; j++) { \
		float s = 0.f, m = mean[(size_t)i * w + j]; \
		for (int k = -vr; k <= vr; k++) for (int l = -hr; l <= hr; l++) for (int q = 0; q < nc; q++) { \
			float tmp = (float)in[((size_t)(i + k) * w + (j + l)) * nc + q] - m; \
			s += tmp * tmp; \
		} \
		out[(size_t)i * w + j] = sqrtf(s); \
	} \
	return 0; \
}
ORC_SIGMA_FILTER(orc_sigma_filter_f32, float)
ORC_SIGMA_FILTER(orc_sigma_filter_u8, uint8_t)

/* ==== correlation/image_based_refinement.h: feature-domain subpixel refinement =================================================
 * The reference builds F x (K+1) Eigen matrices per pixel and solves with Eigen::FullPivHouseholderQR (Eigen 3.3, absent from this
 * image); the least-squares solves below go through oracle/small_qr.h, the same published algorithm in float. Everything else —
 * the affine-space reduction, the spherical fit, the weighted median, the border rules, the score comparisons — follows the
 * reference line by line in float. Pinned against the reference's own code compiled with the same stand-in solver
 * (oracle/_ref/libstevi_ref_ibr.so) and against its known-answer tests (tests/test_oracle_pinning.py). */
#include "small_qr.h"

int orc_qr_lstsq(int m, int n, const float* A, const float* b, float* x) { stevi_qr_lstsq(m, n, A, b, x); return 0; }

#define IBR_MAXC 9 /* columns of the largest system: Queen neighbourhood + centre */

static int ibr_family(int mf) { /* matching_costs.h:435-439, :475-479, :498-502, :538-542, :561-565, :584-588 */
	if (mf == MF_NCC || mf == MF_ZNCC) return 0;
	if (mf == MF_SSD || mf == MF_ZSSD) return 1;
	if (mf == MF_SAD || mf == MF_ZSAD) return 2;
	return -1;
}

/* l1optimization.h:33-60 (weighted_median_index) on the one-unknown overload of leastAbsoluteDifferences (:257-283) */
static int ibr_cmp_key(const void* a, const void* b, void* el) {
	const float* e = (const float*)el;
	float x = e[*(const int*)a], y = e[*(const int*)b];
	return x < y ? -1 : (y < x ? 1 : 0);
}
static float ibr_l1_one_unknown(const float* M, const float* rhs, int F) {
	float* t = (float*)malloc(sizeof(float) * 2 * F + sizeof(int) * F);
	float* w = t + F;
	int* key = (int*)(w + F);
	for (int i = 0; i < F; i++) { t[i] = 0.f; w[i] = 0.f; }
	for (int i = 0; i < F; i++) {
		float o = M[i];
		w[i] = fabsf(o);
		t[i] = rhs[i] / o;
		if (isnan(t[F - 1]) || isinf(t[F - 1])) { w[i] = 0.f; t[i] = -INFINITY; } /* (sic) the test reads t_breaks.back(), :273 */
	}
	for (int i = 0; i < F; i++) key[i] = i;
	float sum = 0.f;
	for (int i = 0; i < F; i++) sum += w[i];
	qsort_r(key, F, sizeof(int), ibr_cmp_key, t);
	float partial = 0.f, res = t[key[F - 1]];
	for (int i = 0; i < F; i++) {
		float tmp = 2 * w[key[i]];
		if (partial + tmp > sum) { res = t[key[i]]; break; }
		partial += tmp;
	}
	free(t);
	return res;
}

/* barycentricBestApproximation(A, b): A is F x n column-major (A[c + k*F]), origin of the affine space = last column
 * (affine_utils.h:44-57); coeffs gets n values. Returns 0 when the family has no such system (L1 with more than one unknown). */
static int ibr_best_approximation(int family, const float* A, const float* b, int F, int n, float* coeffs) {
	const int K = n - 1;
	float* M = (float*)malloc(sizeof(float) * (size_t)F * (K + 4));
	float* rhs = M + (size_t)F * K;
	float* bperp = rhs + F;
	float* aperp = bperp + F;
	const float* aK = A + (size_t)K * F;
	float alpha[IBR_MAXC], x[IBR_MAXC];
	for (int k = 0; k < K; k++) for (int c = 0; c < F; c++) M[c + (size_t)k * F] = A[c + (size_t)k * F] - aK[c];
	if (family == 1) { /* l2optimization.h:48-63 */
		for (int c = 0; c < F; c++) rhs[c] = b[c] - aK[c];
		stevi_qr_lstsq(F, K, M, rhs, alpha);
	} else if (family == 2) { /* l1optimization.h:293-310 */
		if (K != 1) { free(M); return 0; }
		for (int c = 0; c < F; c++) rhs[c] = b[c] - aK[c];
		alpha[0] = ibr_l1_one_unknown(M, rhs, F);
	} else { /* sphericaloptimization.h:35-63 */
		stevi_qr_lstsq(F, n, A, b, x);
		for (int c = 0; c < F; c++) { float s = 0.f; for (int k = 0; k < n; k++) s += A[c + (size_t)k * F] * x[k]; bperp[c] = s; }
		stevi_qr_lstsq(F, K, M, aK, x);
		for (int c = 0; c < F; c++) { float s = 0.f; for (int k = 0; k < K; k++) s += M[c + (size_t)k * F] * x[k]; aperp[c] = aK[c] - s; }
		float aa = 0.f, ab = 0.f;
		for (int c = 0; c < F; c++) aa += aperp[c] * aperp[c];
		for (int c = 0; c < F; c++) ab += aperp[c] * bperp[c];
		float g = aa / ab;
		for (int c = 0; c < F; c++) rhs[c] = g * bperp[c] - aK[c];
		stevi_qr_lstsq(F, K, M, rhs, alpha);
	}
	float sum = 0.f;
	for (int k = 0; k < K; k++) { coeffs[k] = alpha[k]; sum += alpha[k]; }
	coeffs[K] = 1 - sum; /* affine_utils.h:83-97 */
	free(M);
	return 1;
}

static float ibr_norm(const float* v, int F) { float s = 0.f; for (int c = 0; c < F; c++) s += v[c] * v[c]; return sqrtf(s); }
/* featureComparison of the (normalised when the function is) vectors s and t: image_based_refinement.h:253-306 */
static float ibr_compare(int mf, const float* s, const float* t, int F, float* tmp) {
	float* a = tmp; float* b = tmp + F;
	if (mf_normalized(mf)) {
		float ns = ibr_norm(s, F), nt = ibr_norm(t, F);
		for (int c = 0; c < F; c++) { a[c] = s[c] / ns; b[c] = t[c] / nt; }
	} else { memcpy(a, s, sizeof(float) * F); memcpy(b, t, sizeof(float) * F); }
	return compare_f32(mf, a, b, F);
}

/* image_based_refinement.h:184-355 */
int orc_refine_barycentric_disp(int mf, const float* fl, const float* fr, int h, int wl, int wr, int F, const int32_t* idx, int dir, float* out) {
	const int family = ibr_family(mf);
	if (family < 0) return 3;
	const float* src = dir ? fr : fl; const float* tgt = dir ? fl : fr;
	const int ws = dir ? wr : wl, wt = dir ? wl : wr, sign = dir ? 1 : -1, score_mode = mf_strategy(mf) == STRAT_SCORE;
	#pragma omp parallel for
	for (int i = 0; i < h; i++) {
		float* buf = (float*)malloc(sizeof(float) * F * 5);
		float* A = buf; float* interp = buf + 2 * F; float* tmp = buf + 3 * F;
		for (int j = 0; j < ws; j++) {
			int d = idx[(size_t)i * ws + j], jd = j + sign * d;
			if (jd < 1 || jd + 1 >= ws) { out[(size_t)i * ws + j] = (float)d; continue; }
			const float* s = src + ((size_t)i * ws + j) * F;
			const float* t0 = tgt + ((size_t)i * wt + jd) * F;
			float cp[2], cm[2];
			memcpy(A, t0, sizeof(float) * F); memcpy(A + F, t0 + F, sizeof(float) * F);
			ibr_best_approximation(family, A, s, F, 2, cp);
			float score = ibr_compare(mf, s, t0, F, tmp), delta = 0.f;
			float dplus = cp[1];
			if (dplus > 0 && dplus < 1) {
				for (int c = 0; c < F; c++) interp[c] = A[c] * cp[0] + A[F + c] * cp[1];
				float sc = ibr_compare(mf, s, interp, F, tmp);
				if (score_mode ? sc > score : sc < score) { score = sc; delta = dplus; }
			}
			memcpy(A, t0 - F, sizeof(float) * F); memcpy(A + F, t0, sizeof(float) * F);
			ibr_best_approximation(family, A, s, F, 2, cm);
			float dminus = cm[0];
			if (dminus > 0 && dminus < 1) {
				for (int c = 0; c < F; c++) interp[c] = A[c] * cm[0] + A[F + c] * cm[1];
				float sc = ibr_compare(mf, s, interp, F, tmp);
				if (score_mode ? sc > score : sc < score) { score = sc; delta = -dminus; }
			}
			out[(size_t)i * ws + j] = d + sign * delta;
		}
		free(buf);
	}
	return 0;
}

/* image_based_refinement.h:31-105 */
int orc_refine_barycentric_symmetric_disp(int mf, int radius, const float* fl, const float* fr, int h, int wl, int wr, int F,
										  const int32_t* idx, int width, int dir, float* out) {
	const int family = ibr_family(mf);
	if (family < 0 || family == 2 || radius < 1 || 2 * radius + 1 > IBR_MAXC) return 3;
	const float* src = dir ? fr : fl; const float* tgt = dir ? fl : fr;
	const int ws = dir ? wr : wl, wt = dir ? wl : wr, sign = dir ? 1 : -1, n = 2 * radius + 1;
	#pragma omp parallel for
	for (int i = 0; i < h; i++) {
		float* A = (float*)malloc(sizeof(float) * F * n);
		for (int j = 0; j < ws; j++) {
			int d = idx[(size_t)i * ws + j], jd = j + sign * d;
			float res = (float)d;
			if (!(j < 0 || j + 1 >= ws) && !(jd - radius < 0 || jd + 1 > ws - radius) && !(d == 0 || d + 1 >= width)) {
				float co[IBR_MAXC];
				for (int p = -radius; p <= radius; p++) memcpy(A + (size_t)(p + radius) * F, tgt + ((size_t)i * wt + jd + p) * F, sizeof(float) * F);
				ibr_best_approximation(family, A, src + ((size_t)i * ws + j) * F, F, n, co);
				float delta = 0.f;
				for (int p = -radius; p <= radius; p++) delta += co[p + radius] * (float)p;
				if (fabsf(delta) < 1) res = d + delta;
			}
			out[(size_t)i * ws + j] = res;
		}
		free(A);
	}
	return 0;
}

/* image_based_refinement.h:675-880 (sym 0) and :1068-1241 (sym 1); contiguity 0 Queen, 1 Rook, 2 Bishop (utils/contiguity.h:98-164) */
int orc_refine_barycentric_2d_disp(int mf, int contiguity, int sym, const float* fl, const float* fr, int h, int w, int F,
								   const int32_t* idx, const int* so, int dir, float* out) {
	static const int corner[3][3][2] = {{{1, 1}, {1, 0}, {0, 1}}, {{1, 0}, {0, 1}, {0, 0}}, {{1, 1}, {0, 0}, {0, 0}}};
	static const int ncorner[3] = {3, 2, 1};
	static const int dirs[3][8][2] = {{{1, 1}, {1, 0}, {1, -1}, {0, 1}, {0, -1}, {-1, 1}, {-1, 0}, {-1, -1}},
									  {{1, 0}, {0, 1}, {0, -1}, {-1, 0}, {0, 0}, {0, 0}, {0, 0}, {0, 0}},
									  {{1, 1}, {1, -1}, {-1, 1}, {-1, -1}, {0, 0}, {0, 0}, {0, 0}, {0, 0}}};
	static const int ndirs[3] = {8, 4, 4};
	const int family = ibr_family(mf);
	if (family < 0 || family == 2 || contiguity < 0 || contiguity > 2) return 3;
	const float* src = dir ? fr : fl; const float* tgt = dir ? fl : fr;
	const int score_mode = mf_strategy(mf) == STRAT_SCORE, K = sym ? ndirs[contiguity] : ncorner[contiguity], nsys = sym ? 1 : 4;
	#pragma omp parallel for
	for (int i = 0; i < h; i++) {
		float* buf = (float*)malloc(sizeof(float) * F * (IBR_MAXC + 3));
		float* A = buf; float* interp = buf + (size_t)IBR_MAXC * F; float* tmp = interp + F;
		for (int j = 0; j < w; j++) {
			int d0 = idx[((size_t)i * w + j) * 2], d1 = idx[((size_t)i * w + j) * 2 + 1], id = i + d0, jd = j + d1;
			float dd0 = 0.f, dd1 = 0.f;
			if (!(id < 1 || id + 1 >= h) && !(jd < 1 || jd + 1 >= w) && !(d0 < so[0] || d0 > so[1]) && !(d1 < so[2] || d1 > so[3])) {
				const float* s = src + ((size_t)i * w + j) * F;
				const float* t0 = tgt + ((size_t)id * w + jd) * F;
				float score = ibr_compare(mf, s, t0, F, tmp);
				for (int q = 0; q < nsys; q++) {
					int sx = sym ? 1 : (q < 2 ? 1 : -1), sy = sym ? 1 : ((q & 1) ? -1 : 1), di[IBR_MAXC], dj[IBR_MAXC];
					float co[IBR_MAXC];
					for (int k = 0; k < K; k++) {
						di[k] = (sym ? dirs[contiguity][k][0] : corner[contiguity][k][0]) * sx;
						dj[k] = (sym ? dirs[contiguity][k][1] : corner[contiguity][k][1]) * sy;
						memcpy(A + (size_t)k * F, tgt + ((size_t)(id + di[k]) * w + jd + dj[k]) * F, sizeof(float) * F);
					}
					memcpy(A + (size_t)K * F, t0, sizeof(float) * F);
					ibr_best_approximation(family, A, s, F, K + 1, co);
					float e0 = 0.f, e1 = 0.f;
					for (int k = 0; k < K; k++) { e0 += co[k] * di[k]; e1 += co[k] * dj[k]; }
					if (fabsf(e0) <= 1. && fabsf(e1) <= 1.) {
						for (int c = 0; c < F; c++) { float v = 0.f; for (int k = 0; k <= K; k++) v += A[c + (size_t)k * F] * co[k]; interp[c] = v; }
						float sc = ibr_compare(mf, s, interp, F, tmp);
						if (score_mode ? sc > score : sc < score) { dd0 = e0; dd1 = e1; score = sc; }
					}
				}
			}
			out[((size_t)i * w + j) * 2] = d0 + dd0;
			out[((size_t)i * w + j) * 2 + 1] = d1 + dd1;
		}
		free(buf);
	}
	return 0;
}

/* image_based_refinement.h:1591-1690 */
int orc_refine_cost_symmetric_disp(int mf, const float* fl, const float* fr, int h, int wl, int wr, int F, const int32_t* idx,
								   const float* cv, int D, int dir, float* out) {
	const float* src = dir ? fr : fl; const float* tgt = dir ? fl : fr;
	const int ws = dir ? wr : wl, wt = dir ? wl : wr, sign = dir ? 1 : -1;
	#pragma omp parallel for
	for (int i = 0; i < h; i++) {
		float* buf = (float*)malloc(sizeof(float) * F * 4);
		float* si = buf; float* tn = buf + F;
		for (int j = 0; j < ws; j++) {
			int d = idx[(size_t)i * ws + j], jd = j + sign * d;
			float delta = 0;
			if (j > 1 && j + 1 < ws && d > 0 && d + 1 < D) {
				const float* c = cv + ((size_t)i * ws + j) * D + d;
				float cm1 = c[-1], c0 = c[0], c1 = c[1];
				delta = (cm1 - c1) / (2 * (c1 - 2 * c0 + cm1));
				int sdir = delta > 0 ? -1 : 1;
				if (jd + 1 < ws && jd > 1) {
					const float* s = src + ((size_t)i * ws + j) * F;
					for (int k = 0; k < F; k++) si[k] = (float)(0.5 * s[k] + 0.5 * s[k + sdir * F]);
					if (mf_normalized(mf)) { float nn = ibr_norm(si, F); for (int k = 0; k < F; k++) si[k] /= nn; }
					float f[3];
					for (int q = -1; q <= 1; q++) {
						const float* t = tgt + ((size_t)i * wt + jd + q) * F;
						memcpy(tn, t, sizeof(float) * F);
						if (mf_normalized(mf)) { float nn = ibr_norm(tn, F); for (int k = 0; k < F; k++) tn[k] /= nn; }
						f[q + 1] = compare_f32(mf, si, tn, F);
					}
					float delta2 = (float)((f[0] - f[2]) / (2 * (f[2] - 2 * f[1] + f[0])) - sdir * 0.5);
					if (fabsf(delta2) < 1.) delta = (delta + delta2) / 2;
				}
			}
			out[(size_t)i * ws + j] = d + delta;
		}
		free(buf);
	}
	return 0;
}

/* ==== correlation/on_demand_cost_volume.h + correlation/patchmatch.h ===============================================================
 * fs (h,w,F), ft (ht,wt,F): feature volumes as the on-demand volume's constructor leaves them (already processed for the function).
 * search: {lo, up} (dims 1, columns) or {lo0, up0, lo1, up1} (dims 2: rows, columns). */
typedef struct { const float* fs; const float* ft; int h, w, ht, wt, F, dims, lo0, up0, lo1, up1, mf; } od_volume;

/* costValue, on_demand_cost_volume.h:105-181: 0 when the cell has no value */
static int od_cost(const od_volume* v, int i, int j, int d0, int d1, float* c) {
	if (v->dims == 2 && (d0 < v->lo0 || d0 > v->up0)) return 0;
	if (d1 < v->lo1 || d1 > v->up1) return 0;
	int ti = i + (v->dims == 2 ? d0 : 0), tj = j + d1;
	if (ti < 0 || ti >= v->ht || tj < 0 || tj >= v->wt) return 0;
	*c = compare_f32(v->mf, v->fs + ((size_t)i * v->w + j) * v->F, v->ft + ((size_t)ti * v->wt + tj) * v->F, v->F);
	return 1;
}

static void od_setup(od_volume* v, int mf, int dims, const float* fs, const float* ft, int h, int w, int ht, int wt, int F, const int* so) {
	v->fs = fs; v->ft = ft; v->h = h; v->w = w; v->ht = ht; v->wt = wt; v->F = F; v->dims = dims; v->mf = mf;
	v->lo0 = v->up0 = 0;
	if (dims == 1) { v->lo1 = so[0]; v->up1 = so[1]; } else { v->lo0 = so[0]; v->up0 = so[1]; v->lo1 = so[2]; v->up1 = so[3]; }
}

static float* od_processed(int mf, const float* f, size_t npix, int F, int passes) {
	float* p = (float*)malloc(sizeof(float) * npix * F);
	memcpy(p, f, sizeof(float) * npix * F);
	if (mf_zero_mean(mf) || mf_normalized(mf)) for (int k = 0; k < passes; k++) features_inplace(mf, p, npix, F);
	return p;
}

/* truncatedCostVolume(disp, radius), :187-310, incl. the index handed to costValue as a disparity (:224-241) */
int orc_on_demand_truncated_cost_volume(int mf, int dims, const float* fs, const float* ft, int h, int w, int ht, int wt, int F, const int* so,
										const int32_t* disp, int radius, float* out) {
	float* ps = od_processed(mf, fs, (size_t)h * w, F, 1);
	float* pt = od_processed(mf, ft, (size_t)ht * wt, F, 1);
	od_volume v;
	od_setup(&v, mf, dims, ps, pt, h, w, ht, wt, F, so);
	const int n = 2 * radius + 1;
	const float missing = mf_strategy(mf) == STRAT_COST ? 3.402823466e+38f : 1.17549435e-38f; /* matching_costs.h:706-713 */
	for (int i = 0; i < h; i++) for (int j = 0; j < w; j++) {
		if (dims == 1) {
			int d = disp[(size_t)i * w + j];
			for (int k = 0; k < n; k++) { float c; out[((size_t)i * w + j) * n + k] = od_cost(&v, i, j, 0, (d - v.lo1) + k - radius, &c) ? c : missing; }
		} else {
			int d0 = disp[((size_t)i * w + j) * 2], d1 = disp[((size_t)i * w + j) * 2 + 1];
			for (int a = 0; a < n; a++) for (int b = 0; b < n; b++) {
				float c;
				out[(((size_t)i * w + j) * n + a) * n + b] = od_cost(&v, i, j, (d0 - v.lo0) + a - radius, (d1 - v.lo1) + b - radius, &c) ? c : missing;
			}
		}
	}
	free(ps); free(pt);
	return 0;
}

/* patchMatchTestCost, patchmatch.h:166-231; cur / ok: the cost of the pixel's disparity (an empty optional compares below any value) */
static int pm_test(const od_volume* v, int32_t* sol, float* cur, unsigned char* ok, int i, int j, int d0, int d1) {
	float c;
	if (!od_cost(v, i, j, d0, d1, &c)) return 0;
	size_t px = (size_t)i * v->w + j;
	int score = mf_strategy(v->mf) == STRAT_SCORE;
	int keep = score ? (!ok[px] || c >= cur[px]) : (ok[px] && c <= cur[px]);
	if (!keep) return 0;
	if (v->dims == 2) { sol[px * 2] = d0; sol[px * 2 + 1] = d1; } else sol[px] = d1;
	cur[px] = c; ok[px] = 1;
	return 1;
}

/* patchMatch with nRandomSearch = 0 from a given initial map (patchmatch.h:371-446 propagate, :449-496 the loop, :498-556 the entry that
 * processes the features once more before the on-demand volume does): the deterministic part of the algorithm */
int orc_patch_match(int mf, int dims, const float* fs, const float* ft, int h, int w, int ht, int wt, int F, const int* so, int n_iter, int n_rand,
					const int32_t* init, int32_t* out) {
	if (n_rand != 0 || !init) return 3; /* the random draws are not restated: the reference seeds them from std::random_device */
	float* ps = od_processed(mf, fs, (size_t)h * w, F, 2);
	float* pt = od_processed(mf, ft, (size_t)ht * wt, F, 2);
	od_volume v;
	od_setup(&v, mf, dims, ps, pt, h, w, ht, wt, F, so);
	memcpy(out, init, sizeof(int32_t) * (size_t)h * w * dims);
	float* cur = (float*)malloc(sizeof(float) * (size_t)h * w);
	unsigned char* ok = (unsigned char*)malloc((size_t)h * w);
	for (int i = 0; i < h; i++) for (int j = 0; j < w; j++) {
		size_t px = (size_t)i * w + j;
		ok[px] = (unsigned char)od_cost(&v, i, j, dims == 2 ? out[px * 2] : 0, dims == 2 ? out[px * 2 + 1] : out[px], &cur[px]);
	}
	static const int inc[4][2] = {{1, 1}, {1, -1}, {-1, 1}, {-1, -1}}; /* utils/propagation_direction.h:65-86 */
	for (int it = 0; it < n_iter; it++) {
		int n = 0, di = inc[it % 4][0], dj = inc[it % 4][1];
		/* a backward scan ends before index 0 (initialAndFinalPos<-1> = {size-1, 0}, :45-60) */
		for (int i = 0; i < h; i++) for (int j = dj > 0 ? 0 : w - 1; dj > 0 ? j < w : j > 0; j += dj) {
			int pj = j - dj;
			if (pj < 0 || pj >= w) continue;
			size_t pp = (size_t)i * w + pj;
			n += pm_test(&v, out, cur, ok, i, j, dims == 2 ? out[pp * 2] : 0, dims == 2 ? out[pp * 2 + 1] : out[pp]);
		}
		for (int j = 0; j < w; j++) for (int i = di > 0 ? 0 : h - 1; di > 0 ? i < h : i > 0; i += di) {
			int pi = i - di;
			if (pi < 0 || pi >= h) continue;
			size_t pp = (size_t)pi * w + j;
			n += pm_test(&v, out, cur, ok, i, j, dims == 2 ? out[pp * 2] : 0, dims == 2 ? out[pp * 2 + 1] : out[pp]);
		}
		if (n == 0) break;
	}
	free(cur); free(ok); free(ps); free(pt);
	return 0;
}
