// Box-sum cost-volume kernel for the float matching functions (SURVEY.md §8(a) rows a4+a5+a6 fused; the unfolded
// F-channel volumes of a1 are never materialised).
//
// Reference semantics (file:line relative to the LibStevi tree):
//   unfold.h:248-291                 feature c = (2hr+1)k + l of pixel (i,j) = img(i-vr+k, j-hr+l), 0 outside the image
//   cross_correlations.h:416-462     ZNCC features (v - mean)/norm, :504-550 NCC v/norm, :570-594 ZCC/ZSSD v - mean
//   cross_correlations.h:194-249     CV(i,j,d) = cmp(src(i,j,.), tgt(i,j+s*d,.)); the WHOLE target vector is 0 when
//                                    j+s*d is outside the target image (:235)
//   matching_costs.h:59-156          cmp = sum s*t | sum (s-t)^2 | sum |s-t|
//
// With zero-padded images the window sums of a per-disparity "pair image" P_d(r,c) = op(s(r,c), t(r,c+s*d)) are box
// filters, so every cell costs O(1) instead of O(F):
//   CC/NCC/ZCC/ZNCC  = (box(s*t) - F*m_s*m_t) * a_s * a_t        a = 1/norm (1 when not normalised), m = 0 when not zero-mean
//   SSD/ZSSD         =  box((s-t)^2) - F*(m_s-m_t)^2
//   SAD              =  box(|s-t|)
// and the per-pixel window statistics (m, a, and the cost against the all-zero target vector) come from
// k_window_stats, which follows the reference's summation order. Cells whose target column is outside the image take
// the "zero target vector" value (0 for the correlations, sum of cmp(s_c, 0) for the costs).
//
// Work split: a thread owns 8 consecutive source columns x 4 consecutive disparities and walks down a band of R rows.
// The vertical window sums V[q][c] (NV = 16 columns incl. the +-hr halo, hr <= 4) live in registers and are updated per row
// with the entering and the leaving image row (2 FMA per V); the horizontal sums slide over the 8 output columns.
// Image rows come through L1/L2 as aligned 16-byte loads from zero-padded float copies (they are a few MB: L2-resident);
// stores are 2 x 16 bytes per (row, d) and thread — a half-warp writes 512 contiguous bytes of the reference's
// [H][D][W] layout.
#pragma once
#include "common.cuh"

namespace stevi {

enum { BOX_PROD = 0, BOX_SQ = 1, BOX_ABS = 2 };

struct BoxParams {
	const float* ps; const float* pt;   // padded images, pointing at pixel (0,0); pitch in floats (multiple of 4)
	const float* As; const float* Bs;   // source window statistics, same geometry as ps
	const float* At; const float* Bt;   // target window statistics, same geometry as pt
	int64_t pitch;
	int H, Ws, D, vr, R;
	float Ff;                           // number of features F as float
	float* out; int64_t o0, o1, o2;
	int vec_store;                      // o1 == 1, 16-byte aligned rows
};

template<int N> __device__ __forceinline__ void load_vec(float (&dst)[N], const float* __restrict__ p) {
#pragma unroll
	for (int k = 0; k < N / 4; k++) {
		float4 v = __ldg(reinterpret_cast<const float4*>(p) + k);
		dst[4 * k] = v.x; dst[4 * k + 1] = v.y; dst[4 * k + 2] = v.z; dst[4 * k + 3] = v.w;
	}
}

template<int OP, int NV, int SIGN, bool ADD>
__device__ __forceinline__ void box_accum_row(float (&V)[4][NV], const float* __restrict__ srow, const float* __restrict__ trow) {
	float s[NV], t[NV + 4];
	load_vec<NV>(s, srow);
	load_vec<NV + 4>(t, trow);
#pragma unroll
	for (int q = 0; q < 4; q++) {
#pragma unroll
		for (int c = 0; c < NV; c++) {
			const float tv = SIGN > 0 ? t[c + q] : t[c + 4 - q];
			if (OP == BOX_PROD) {
				V[q][c] = fmaf(ADD ? s[c] : -s[c], tv, V[q][c]);
			} else if (OP == BOX_SQ) {
				const float e = s[c] - tv;
				V[q][c] = fmaf(ADD ? e : -e, e, V[q][c]);
			} else {
				const float e = fabsf(s[c] - tv);
				V[q][c] = ADD ? V[q][c] + e : V[q][c] - e;
			}
		}
	}
}

template<int OP, int HR, int SIGN>
__global__ void __launch_bounds__(256, 2) k_cv_box(const BoxParams p) {
	static_assert(HR >= 1 && HR <= 4, "window half-width");
	constexpr int NV = 16;                  // V columns j0-4 .. j0+11
	constexpr int LH = 4;                   // left halo
	const int cg = threadIdx.x & 15, dg = threadIdx.x >> 4;
	const int j0 = blockIdx.x * 128 + cg * 8;
	const int d0 = blockIdx.y * 64 + dg * 4;
	const int i0 = blockIdx.z * p.R;
	if (d0 >= p.D || j0 >= p.Ws) return;
	// V column c <-> source column j0-LH+c; target column = source column + SIGN*(d0+q):
	//   SIGN>0: base j0-LH+d0, register c+q ; SIGN<0: base j0-LH-d0-4, register c+4-q
	const float* __restrict__ sb = p.ps + (j0 - LH);
	const float* __restrict__ tb = p.pt + (SIGN > 0 ? (j0 - LH + d0) : (j0 - LH - d0 - 4));
	float V[4][NV];
#pragma unroll
	for (int q = 0; q < 4; q++)
#pragma unroll
		for (int c = 0; c < NV; c++) V[q][c] = 0.f;
	for (int k = -p.vr; k <= p.vr; k++)
		box_accum_row<OP, NV, SIGN, true>(V, sb + (int64_t)(i0 + k) * p.pitch, tb + (int64_t)(i0 + k) * p.pitch);
	const int i_end = min(i0 + p.R, p.H);
	for (int i = i0; i < i_end; i++) {
		if (i > i0) {
			box_accum_row<OP, NV, SIGN, true>(V, sb + (int64_t)(i + p.vr) * p.pitch, tb + (int64_t)(i + p.vr) * p.pitch);
			box_accum_row<OP, NV, SIGN, false>(V, sb + (int64_t)(i - p.vr - 1) * p.pitch, tb + (int64_t)(i - p.vr - 1) * p.pitch);
		}
		// per-pixel statistics: source columns j0..j0+7; target columns j0+k+SIGN*(d0+q)
		float as[8], bs[8], at[12], bt[12];
		load_vec<8>(as, p.As + (int64_t)i * p.pitch + j0);
		load_vec<8>(bs, p.Bs + (int64_t)i * p.pitch + j0);
		const int64_t toff = (int64_t)i * p.pitch + (SIGN > 0 ? (j0 + d0) : (j0 - d0 - 4));
		if (OP != BOX_ABS) load_vec<12>(at, p.At + toff);
		load_vec<12>(bt, p.Bt + toff);
#pragma unroll
		for (int q = 0; q < 4; q++) {
			if (d0 + q >= p.D) break;
			float o[8];
			float h = V[q][LH - HR];
#pragma unroll
			for (int c = LH - HR + 1; c <= LH + HR; c++) h += V[q][c];
#pragma unroll
			for (int k = 0; k < 8; k++) {
				if (k > 0) h = (h + V[q][LH + HR + k]) - V[q][LH - HR + k - 1];
				const int ti = SIGN > 0 ? k + q : k + 4 - q;
				float v;
				if (OP == BOX_PROD) {
					v = fmaf(-as[k], at[ti], h);      // as = F*m_s, at = m_t
					v = (v * bs[k]) * bt[ti];         // bs = 1/n_s (NaN when n_s == 0), bt = 1/n_t, 0 outside the image
				} else if (OP == BOX_SQ) {
					const float e = as[k] - at[ti];   // m_s - m_t
					v = fmaf(-p.Ff * e, e, h);
					v = bt[ti] != 0.f ? v : bs[k];    // bs = cost against the zero vector
				} else {
					v = bt[ti] != 0.f ? h : bs[k];
				}
				o[k] = v;
			}
			if (p.vec_store && j0 + 8 <= p.Ws) {
				float4* dst = reinterpret_cast<float4*>(p.out + (int64_t)i * p.o0 + (int64_t)(d0 + q) * p.o2 + j0);
				__stcs(dst, make_float4(o[0], o[1], o[2], o[3]));
				__stcs(dst + 1, make_float4(o[4], o[5], o[6], o[7]));
			} else {
#pragma unroll
				for (int k = 0; k < 8; k++)
					if (j0 + k < p.Ws) p.out[(int64_t)i * p.o0 + (int64_t)(j0 + k) * p.o1 + (int64_t)(d0 + q) * p.o2] = o[k];
			}
		}
	}
}

template<int OP, int HR>
static int32_t launch_cv_box_hr(stevi_ctx* c, const BoxParams& p, int sign, dim3 grid) {
	if (sign > 0) { STEVI_LAUNCH(c, (k_cv_box<OP, HR, 1>), grid, 256, 0, p); }
	else { STEVI_LAUNCH(c, (k_cv_box<OP, HR, -1>), grid, 256, 0, p); }
	return STEVI_OK;
}

template<int OP>
int32_t launch_cv_box(stevi_ctx* c, const BoxParams& p, int hr, int sign) {
	dim3 grid(cdiv(p.Ws, 128), cdiv(p.D, 64), cdiv(p.H, p.R));
	switch (hr) {
	case 1: return launch_cv_box_hr<OP, 1>(c, p, sign, grid);
	case 2: return launch_cv_box_hr<OP, 2>(c, p, sign, grid);
	case 3: return launch_cv_box_hr<OP, 3>(c, p, sign, grid);
	case 4: return launch_cv_box_hr<OP, 4>(c, p, sign, grid);
	default: return fail(c, STEVI_E_UNSUPPORTED, "box cost volume: h_radius %d", hr);
	}
}

// explicit instantiations live in costvol_box_{prod,sq,abs}.cu so that they compile in parallel
int32_t launch_cv_box_prod(stevi_ctx* c, const BoxParams& p, int hr, int sign);
int32_t launch_cv_box_sq(stevi_ctx* c, const BoxParams& p, int hr, int sign);
int32_t launch_cv_box_abs(stevi_ctx* c, const BoxParams& p, int hr, int sign);

} // namespace stevi
