// Winner neighbourhood extraction and cost-based subpixel refinement.
//
//   correlation_base.h:557-577   selectedCost: out(i,j) = cv(i,j,idx(i,j))
//   correlation_base.h:579-674   truncatedCostVolume<T_CV,dir,sdir>: the 2r+1 (Same, Reversed) or 4r+1 (Both) costs around the
//                                winner, NaN where the disparity is out of range or the matching window leaves the image
//   correlation_base.h:727-822   extractInBoundDomain: the same tests as a 0/1 mask
//   cost_based_refinement.h:43-69, :128-163   refineCostTriplet / refineDispCostInterpolation: idx + delta(c-1, c0, c+1)
//
//   correlation_base.h:677-725   truncatedBidirectionaCostVolume: the (2r0+1) x (2r1+1) cells around a 2-D winner
//   cost_based_refinement.h:165-369   refineDisp2dCostInterpolation: one triplet per axis (isotropic) or the intersection of
//                                the two ridge lines fitted through the neighbouring rows / columns (anisotropic)
//
// All of these are one gather per pixel: thread per pixel, pixels of a warp adjacent along the fastest pixel axis.
#include "common.cuh"
#include "kernels.cuh"

namespace stevi {

namespace {

struct Tcv {
	const float* cv; int64_t c0, c1, c2;
	const int32_t* idx; int64_t i0, i1;
	int H, W, D, hr, vr, r, sgn, mode; // sgn = -1 for RightToLeft (correlation_base.h:618), mode = stevi_tcv_direction
	float* out; int64_t o0, o1, o2;
	int mask; // 1: write the in-bound mask (1/0) instead of cost / NaN; D is then the search width
};

__global__ void __launch_bounds__(128) k_truncated_cv(const Tcv p) {
	const int j = blockIdx.x * 128 + threadIdx.x, i = blockIdx.y;
	if (j >= p.W) return;
	const int sel = p.idx[i * p.i0 + j * p.i1];
	const bool vbad = i < p.vr || i + p.vr >= p.H;
	const float bad = p.mask ? 0.f : __int_as_float(0x7fc00000);
	float* o = p.out + i * p.o0 + j * p.o1;
	for (int d = 0; d <= 2 * p.r; d++) {
		const int q = sel + d - p.r;
		int jp = j + p.sgn * (d - p.r);
		const bool pbad = q < 0 || q >= p.D;
		const bool same_bad = pbad || j < p.hr || j + q + p.hr >= p.W || vbad;
		if (p.mode == STEVI_TCV_SAME) {
			o[d * p.o2] = same_bad ? bad : (p.mask ? 1.f : p.cv[i * p.c0 + j * p.c1 + q * p.c2]);
			continue;
		}
		int d_d = 2 * d, d_r = 2 * d + 1;
		if (p.mode == STEVI_TCV_BOTH) {
			if (d == p.r) jp = -1;
			if (d > p.r) { d_d -= 1; d_r -= 1; }
		}
		const bool rbad = pbad || min(jp, j) < p.hr || max(jp, j) + p.hr >= p.W || vbad;
		const float rv = rbad ? bad : (p.mask ? 1.f : p.cv[i * p.c0 + (int64_t)jp * p.c1 + q * p.c2]);
		if (p.mode == STEVI_TCV_REVERSED) { o[d * p.o2] = rv; continue; }
		// Both: at d == r the reversed slot 2r+1 is written (always "bad": jp = -1) and then overwritten by d = r+1
		o[d_d * p.o2] = same_bad ? bad : (p.mask ? 1.f : p.cv[i * p.c0 + j * p.c1 + q * p.c2]);
		if (d != p.r) o[d_r * p.o2] = rv;
	}
}

template<typename T>
__global__ void __launch_bounds__(128) k_selected_cost(const T* __restrict__ cv, int64_t c0, int64_t c1, int64_t c2,
													   const int32_t* __restrict__ idx, int64_t i0, int64_t i1, int H, int W,
													   T* __restrict__ out, int64_t o0, int64_t o1) {
	const int j = blockIdx.x * 128 + threadIdx.x, i = blockIdx.y;
	if (j >= W) return;
	out[i * o0 + j * o1] = cv[i * c0 + j * c1 + (int64_t)idx[i * i0 + j * i1] * c2];
}

// cost_based_refinement.h:43-69
__device__ __forceinline__ float refine_triplet(int kernel, float cm1, float c0, float c1) {
	if (kernel == STEVI_EQUIANGULAR) {
		float alpha = copysignf(1.f, __fsub_rn(c0, cm1));
		alpha = __fmul_rn(alpha, fmaxf(fabsf(__fsub_rn(c0, cm1)), fabsf(__fsub_rn(c1, c0))));
		return __fdiv_rn(__fsub_rn(c1, cm1), __fmul_rn(2.f, alpha));
	}
	if (kernel == STEVI_PARABOLA)
		return __fdiv_rn(__fsub_rn(cm1, c1), __fmul_rn(2.f, __fadd_rn(__fsub_rn(c1, __fmul_rn(2.f, c0)), cm1)));
	const float lm = logf(cm1), l0 = logf(c0), l1 = logf(c1);
	return __fdiv_rn(__fsub_rn(lm, l1), __fmul_rn(2.f, __fadd_rn(__fsub_rn(l1, __fmul_rn(2.f, l0)), lm)));
}

__global__ void __launch_bounds__(128) k_refine_interp(const float* __restrict__ tcv, int64_t t0, int64_t t1, int64_t t2,
													   const int32_t* __restrict__ idx, int64_t i0, int64_t i1, int H, int W, int r, int kernel,
													   float* __restrict__ out, int64_t o0, int64_t o1) {
	const int j = blockIdx.x * 128 + threadIdx.x, i = blockIdx.y;
	if (j >= W) return;
	const float* c = tcv + i * t0 + j * t1;
	out[i * o0 + j * o1] = __fadd_rn((float)idx[i * i0 + j * i1], refine_triplet(kernel, c[(r - 1) * t2], c[r * t2], c[(r + 1) * t2]));
}

// truncatedCostVolume<Same, r = 1> + refineDispCostInterpolation straight from the cost volume: 3 gathers per pixel
__global__ void __launch_bounds__(128) k_refine_from_cv(const Tcv p, int kernel, float* __restrict__ out, int64_t r0, int64_t r1) {
	const int j = blockIdx.x * 128 + threadIdx.x, i = blockIdx.y;
	if (j >= p.W) return;
	const int sel = p.idx[i * p.i0 + j * p.i1];
	const bool vbad = i < p.vr || i + p.vr >= p.H;
	float c[3];
#pragma unroll
	for (int d = 0; d < 3; d++) {
		const int q = sel + d - 1;
		const bool bad = q < 0 || q >= p.D || j < p.hr || j + q + p.hr >= p.W || vbad;
		c[d] = bad ? __int_as_float(0x7fc00000) : p.cv[i * p.c0 + j * p.c1 + q * p.c2];
		if (p.out) p.out[i * p.o0 + j * p.o1 + d * p.o2] = c[d];
	}
	out[i * r0 + j * r1] = __fadd_rn((float)sel, refine_triplet(kernel, c[0], c[1], c[2]));
}

// correlation_base.h:677-725: thread per (pixel, axis-0 cell): the n1 cells of one row of the neighbourhood are adjacent in
// the reference's output layout; NaN where an index leaves the search window (valueOrAlt, :718)
struct Tcv2 {
	const float* cv; int64_t c0, c1, c2, c3;
	const int32_t* idx; int64_t i0, i1, i2;
	int H, W, D1, D2, r0, r1;
	float* out; int64_t o0, o1, o2, o3;
};

__global__ void __launch_bounds__(128) k_truncated_cv_2d(const Tcv2 p) {
	const int j = blockIdx.x * 128 + threadIdx.x, i = blockIdx.y, a = blockIdx.z;
	if (j >= p.W) return;
	const int32_t* s = p.idx + i * p.i0 + j * p.i1;
	const int p0 = s[0] + a - p.r0, s1 = s[p.i2] - p.r1;
	const bool bad0 = p0 < 0 || p0 >= p.D1;
	const float* c = p.cv + i * p.c0 + j * p.c1 + (int64_t)p0 * p.c2;
	float* o = p.out + i * p.o0 + j * p.o1 + a * p.o2;
	for (int b = 0; b <= 2 * p.r1; b++) {
		const int p1 = s1 + b;
		o[b * p.o3] = (bad0 || p1 < 0 || p1 >= p.D2) ? __int_as_float(0x7fc00000) : c[(int64_t)p1 * p.c3];
	}
}

// cost_based_refinement.h:208-246: last extremum along one axis of the neighbourhood
template<bool SCORE>
__device__ __forceinline__ int tcv_arg_best(const float* c, int n, int64_t stride) {
	float best = SCORE ? -INFINITY : INFINITY;
	int arg = 0;
	for (int a = 0; a < n; a++) {
		const float x = c[a * stride];
		if (SCORE ? (x >= best) : (x <= best)) { best = x; arg = a; }
	}
	return arg;
}

// cost_based_refinement.h:165-369, thread per pixel. Whether the volume holds scores is read off the image's centre pixel
// (:184-206) by every thread (five broadcast loads). The three products of the line intersection (:350-351) are fused
// multiply-adds, as the reference's Release build contracts them.
template<bool ANISO>
__global__ void __launch_bounds__(128) k_refine_2d(const float* __restrict__ tcv, int64_t t0, int64_t t1, int64_t t2, int64_t t3,
												   const int32_t* __restrict__ raw, int64_t i0, int64_t i1, int64_t i2, int H, int W,
												   int n0, int n1, int kernel, float* __restrict__ out, int64_t o0, int64_t o1, int64_t o2) {
	const int j = blockIdx.x * 128 + threadIdx.x, i = blockIdx.y;
	if (j >= W) return;
	const int r0 = (n0 - 1) / 2, r1 = (n1 - 1) / 2;
	const float* c = tcv + i * t0 + j * t1;
#define TC(a, b) c[(a) * t2 + (b) * t3]
	const float cc = TC(r0, r1);
	const float d0c = refine_triplet(kernel, TC(r0 - 1, r1), cc, TC(r0 + 1, r1));
	const float d1c = refine_triplet(kernel, TC(r0, r1 - 1), cc, TC(r0, r1 + 1));
	float delta0 = d0c, delta1 = d1c;
	if (ANISO) {
		const float* m = tcv + (H / 2) * t0 + (W / 2) * t1;
		const float v0 = m[r0 * t2 + r1 * t3];
		const bool score = v0 > m[(r0 + 1) * t2 + r1 * t3] || v0 > m[(r0 - 1) * t2 + r1 * t3] || v0 > m[r0 * t2 + (r1 + 1) * t3] ||
						   v0 > m[r0 * t2 + (r1 - 1) * t3];
		float d00 = d0c, d02 = d0c, d10 = d1c, d12 = d1c;
		int a = score ? tcv_arg_best<true>(c + (r1 - 1) * t3, n0, t2) : tcv_arg_best<false>(c + (r1 - 1) * t3, n0, t2);
		if (a > 0 && a < n0 - 1) d00 = __fadd_rn((float)(a - r0), refine_triplet(kernel, TC(a - 1, r1 - 1), TC(a, r1 - 1), TC(a + 1, r1 - 1)));
		a = score ? tcv_arg_best<true>(c + (r1 + 1) * t3, n0, t2) : tcv_arg_best<false>(c + (r1 + 1) * t3, n0, t2);
		if (a > 0 && a < n0 - 1) d02 = __fadd_rn((float)(a - r0), refine_triplet(kernel, TC(a - 1, r1 + 1), TC(a, r1 + 1), TC(a + 1, r1 + 1)));
		const float a0 = __fdiv_rn(__fsub_rn(d02, d00), 2.f);
		const float b0 = __fdiv_rn(__fadd_rn(__fadd_rn(d00, d0c), d02), 3.f);
		a = score ? tcv_arg_best<true>(c + (r0 - 1) * t2, n1, t3) : tcv_arg_best<false>(c + (r0 - 1) * t2, n1, t3);
		if (a > 0 && a < n1 - 1) d10 = __fadd_rn((float)(a - r1), refine_triplet(kernel, TC(r0 - 1, a - 1), TC(r0 - 1, a), TC(r0 - 1, a + 1)));
		a = score ? tcv_arg_best<true>(c + (r0 + 1) * t2, n1, t3) : tcv_arg_best<false>(c + (r0 + 1) * t2, n1, t3);
		if (a > 0 && a < n1 - 1) d12 = __fadd_rn((float)(a - r1), refine_triplet(kernel, TC(r0 + 1, a - 1), TC(r0 + 1, a), TC(r0 + 1, a + 1)));
		const float a1 = __fdiv_rn(__fsub_rn(d12, d10), 2.f);
		const float b1 = __fdiv_rn(__fadd_rn(__fadd_rn(d10, d1c), d12), 3.f);
		const float den = __fmaf_rn(-a0, a1, 1.f);
		delta0 = __fdiv_rn(__fmaf_rn(a0, b1, b0), den);
		delta1 = __fdiv_rn(__fmaf_rn(a1, b0, b1), den);
	}
#undef TC
	// :356-360 (a NaN fails both comparisons of !(<= 1))
	if (!(fabsf(delta0) <= 1.f) || !(fabsf(delta1) <= 1.f)) { delta0 = 0.f; delta1 = 0.f; }
	const int32_t* s = raw + i * i0 + j * i1;
	float* o = out + i * o0 + j * o1;
	o[0] = __fadd_rn((float)s[0], delta0);
	o[o2] = __fadd_rn((float)s[i2], delta1);
}

// cost_based_refinement.h:71-126 refineCostPatch + :371-433 refineDisp2dCostPatchInterpolation: least-squares quadric
// c(v,h) = a v^2 + b v h + g h^2 + d v + e h + f through the 3x3 cells around the centre, then its stationary point
// [2a b; b 2g]^-1 (-d, -e). On the 3x3 grid the normal equations (A^T A)^-1 A^T L of the reference decouple, so the six
// coefficients are plain row / column / diagonal sums. Thread per pixel.
__global__ void __launch_bounds__(128) k_refine_2d_patch(const float* __restrict__ tcv, int64_t t0, int64_t t1, int64_t t2, int64_t t3,
														 const int32_t* __restrict__ raw, int64_t i0, int64_t i1, int64_t i2, int H, int W,
														 int r0, int r1, int gaussian, float* __restrict__ out, int64_t o0, int64_t o1, int64_t o2) {
	const int j = blockIdx.x * 128 + threadIdx.x, i = blockIdx.y;
	if (j >= W) return;
	const float* c = tcv + i * t0 + j * t1 + (r0 - 1) * t2 + (r1 - 1) * t3;
	float L[3][3];
#pragma unroll
	for (int a = 0; a < 3; a++)
#pragma unroll
		for (int b = 0; b < 3; b++) { const float x = c[a * t2 + b * t3]; L[a][b] = gaussian ? logf(x) : x; }
	const float rm = L[0][0] + L[0][1] + L[0][2], rz = L[1][0] + L[1][1] + L[1][2], rp = L[2][0] + L[2][1] + L[2][2];
	const float cm = L[0][0] + L[1][0] + L[2][0], cz = L[0][1] + L[1][1] + L[2][1], cp = L[0][2] + L[1][2] + L[2][2];
	const float qa = (rm + rp - 2.f * rz) / 6.f, qg = (cm + cp - 2.f * cz) / 6.f;
	const float qb = (L[0][0] + L[2][2] - L[0][2] - L[2][0]) / 4.f;
	const float qd = (rp - rm) / 6.f, qe = (cp - cm) / 6.f;
	const float det = 4.f * qa * qg - qb * qb;
	float delta0 = (qb * qe - 2.f * qg * qd) / det;
	float delta1 = (qb * qd - 2.f * qa * qe) / det;
	if (!(fabsf(delta0) <= 1.f) || !(fabsf(delta1) <= 1.f)) { delta0 = 0.f; delta1 = 0.f; }   // :421-425
	const int32_t* s = raw + i * i0 + j * i1;
	float* o = out + i * o0 + j * o1;
	o[0] = __fadd_rn((float)s[0], delta0);
	o[o2] = __fadd_rn((float)s[i2], delta1);
}

int32_t check_idx(stevi_ctx* c, const View& idx, int H, int W, const char* who) {
	if (idx.ndim != 2 || idx.dtype != STEVI_I32 || idx.shape[0] != H || idx.shape[1] != W)
		return fail(c, STEVI_E_SHAPE, "%s: index map must be (%d,%d) i32", who, H, W);
	if (H > 65535) return fail(c, STEVI_E_UNSUPPORTED, "%s: more than 65535 rows", who);
	return STEVI_OK;
}

} // namespace

int32_t truncated_cv_dev(stevi_ctx* c, const View* cv, const View& idx, int width, int hr, int vr, int r, int direction, int tdir,
						 const View& out) {
	const bool mask = cv == nullptr;
	const int H = idx.shape[0], W = idx.shape[1];
	if (!mask && (cv->ndim != 3 || cv->dtype != STEVI_F32)) return fail(c, STEVI_E_DTYPE, "truncated_cost_volume: cv must be (H,W,D) f32");
	if (!mask && (cv->shape[0] != H || cv->shape[1] != W)) return fail(c, STEVI_E_SHAPE, "truncated_cost_volume: cv and index map disagree");
	STEVI_TRY(check_idx(c, idx, H, W, "truncated_cost_volume"));
	if (r < 0 || hr < 0 || vr < 0 || tdir < 0 || tdir > 2) return fail(c, STEVI_E_ARG, "truncated_cost_volume: bad radius / direction");
	const int T = tdir == STEVI_TCV_BOTH ? 4 * r + 1 : 2 * r + 1;
	if (out.ndim != 3 || out.dtype != STEVI_F32 || out.shape[0] != H || out.shape[1] != W || out.shape[2] != T)
		return fail(c, STEVI_E_SHAPE, "truncated_cost_volume: out must be (%d,%d,%d) f32", H, W, T);
	if (H == 0 || W == 0) return STEVI_OK;
	Tcv p;
	p.cv = mask ? nullptr : cv->as<float>();
	p.c0 = mask ? 0 : cv->st[0]; p.c1 = mask ? 0 : cv->st[1]; p.c2 = mask ? 0 : cv->st[2];
	p.idx = idx.as<int32_t>(); p.i0 = idx.st[0]; p.i1 = idx.st[1];
	p.H = H; p.W = W; p.D = mask ? width : cv->shape[2]; p.hr = hr; p.vr = vr; p.r = r;
	p.sgn = direction == STEVI_RIGHT_TO_LEFT ? -1 : 1; p.mode = tdir;
	p.out = out.as<float>(); p.o0 = out.st[0]; p.o1 = out.st[1]; p.o2 = out.st[2];
	p.mask = mask ? 1 : 0;
	dim3 grid(cdiv(W, 128), H);
	STEVI_LAUNCH(c, k_truncated_cv, grid, 128, 0, p);
	return STEVI_OK;
}

int32_t selected_cost_dev(stevi_ctx* c, const View& cv, const View& idx, const View& out) {
	if (cv.ndim != 3) return fail(c, STEVI_E_SHAPE, "selected_cost: cv must be (H,W,D)");
	const int H = cv.shape[0], W = cv.shape[1];
	STEVI_TRY(check_idx(c, idx, H, W, "selected_cost"));
	if (out.ndim != 2 || out.dtype != cv.dtype || out.shape[0] != H || out.shape[1] != W)
		return fail(c, STEVI_E_SHAPE, "selected_cost: out must be (H,W) of the volume's type");
	if (H == 0 || W == 0) return STEVI_OK;
	dim3 grid(cdiv(W, 128), H);
	switch (cv.dtype) {
	case STEVI_F32: case STEVI_I32: case STEVI_U32:
		STEVI_LAUNCH(c, k_selected_cost<uint32_t>, grid, 128, 0, cv.as<uint32_t>(), cv.st[0], cv.st[1], cv.st[2], idx.as<int32_t>(), idx.st[0], idx.st[1],
					 H, W, out.as<uint32_t>(), out.st[0], out.st[1]);
		return STEVI_OK;
	case STEVI_U16:
		STEVI_LAUNCH(c, k_selected_cost<uint16_t>, grid, 128, 0, cv.as<uint16_t>(), cv.st[0], cv.st[1], cv.st[2], idx.as<int32_t>(), idx.st[0], idx.st[1],
					 H, W, out.as<uint16_t>(), out.st[0], out.st[1]);
		return STEVI_OK;
	default: return fail(c, STEVI_E_DTYPE, "selected_cost: volume dtype");
	}
}

int32_t refine_interp_dev(stevi_ctx* c, const View& tcv, const View& idx, int kernel, const View& out) {
	if (tcv.ndim != 3 || tcv.dtype != STEVI_F32) return fail(c, STEVI_E_DTYPE, "refine: truncated cost volume must be (H,W,T) f32");
	const int H = tcv.shape[0], W = tcv.shape[1], T = tcv.shape[2];
	const int r = (T - 1) / 2;
	// cost_based_refinement.h:141-143: an even or too small depth gives an empty array
	if (r < 1 || 2 * r + 1 != T) return fail(c, STEVI_E_SHAPE, "refine: depth %d is even or < 3 (cost_based_refinement.h:141)", T);
	if (kernel < 0 || kernel > 2) return fail(c, STEVI_E_ARG, "refine: unknown interpolation kernel");
	STEVI_TRY(check_idx(c, idx, H, W, "refine"));
	if (out.ndim != 2 || out.dtype != STEVI_F32 || out.shape[0] != H || out.shape[1] != W) return fail(c, STEVI_E_SHAPE, "refine: out must be (H,W) f32");
	if (H == 0 || W == 0) return STEVI_OK;
	dim3 grid(cdiv(W, 128), H);
	STEVI_LAUNCH(c, k_refine_interp, grid, 128, 0, tcv.as<float>(), tcv.st[0], tcv.st[1], tcv.st[2], idx.as<int32_t>(), idx.st[0], idx.st[1],
				 H, W, r, kernel, out.as<float>(), out.st[0], out.st[1]);
	return STEVI_OK;
}

int32_t truncated_cv2d_dev(stevi_ctx* c, const View& cv, const View& idx, int radius0, int radius1, const View& out) {
	if (cv.ndim != 4 || cv.dtype != STEVI_F32) return fail(c, STEVI_E_DTYPE, "truncated_cost_volume_2d: cv must be (H,W,D1,D2) f32");
	const int H = cv.shape[0], W = cv.shape[1];
	if (idx.ndim != 3 || idx.dtype != STEVI_I32 || idx.shape[0] != H || idx.shape[1] != W || idx.shape[2] != 2)
		return fail(c, STEVI_E_SHAPE, "truncated_cost_volume_2d: index map must be (%d,%d,2) i32", H, W);
	if (radius0 < 0 || radius0 > 255 || radius1 < 0 || radius1 > 255) return fail(c, STEVI_E_ARG, "truncated_cost_volume_2d: radii are uint8_t in the reference");
	// correlation_base.h:689-695
	const int r0 = radius0 < 1 ? 1 : radius0, r1 = radius1 < 1 ? r0 : radius1;
	if (out.ndim != 4 || out.dtype != STEVI_F32 || out.shape[0] != H || out.shape[1] != W || out.shape[2] != 2 * r0 + 1 || out.shape[3] != 2 * r1 + 1)
		return fail(c, STEVI_E_SHAPE, "truncated_cost_volume_2d: out must be (%d,%d,%d,%d) f32", H, W, 2 * r0 + 1, 2 * r1 + 1);
	if (H > 65535) return fail(c, STEVI_E_UNSUPPORTED, "truncated_cost_volume_2d: more than 65535 rows");
	if (H == 0 || W == 0) return STEVI_OK;
	Tcv2 p;
	p.cv = cv.as<float>(); p.c0 = cv.st[0]; p.c1 = cv.st[1]; p.c2 = cv.st[2]; p.c3 = cv.st[3];
	p.idx = idx.as<int32_t>(); p.i0 = idx.st[0]; p.i1 = idx.st[1]; p.i2 = idx.st[2];
	p.H = H; p.W = W; p.D1 = cv.shape[2]; p.D2 = cv.shape[3]; p.r0 = r0; p.r1 = r1;
	p.out = out.as<float>(); p.o0 = out.st[0]; p.o1 = out.st[1]; p.o2 = out.st[2]; p.o3 = out.st[3];
	dim3 grid(cdiv(W, 128), H, 2 * r0 + 1);
	STEVI_LAUNCH(c, k_truncated_cv_2d, grid, 128, 0, p);
	return STEVI_OK;
}

int32_t refine2d_interp_dev(stevi_ctx* c, const View& tcv, const View& raw, int kernel, int isotropy, const View& out) {
	if (tcv.ndim != 4 || tcv.dtype != STEVI_F32) return fail(c, STEVI_E_DTYPE, "refine_2d: truncated cost volume must be (H,W,T0,T1) f32");
	const int H = tcv.shape[0], W = tcv.shape[1], n0 = tcv.shape[2], n1 = tcv.shape[3];
	const int r0 = (n0 - 1) / 2, r1 = (n1 - 1) / 2;
	// cost_based_refinement.h:180-182: an even or too small side gives an empty array
	if (r0 < 1 || r1 < 1 || 2 * r0 + 1 != n0 || 2 * r1 + 1 != n1)
		return fail(c, STEVI_E_SHAPE, "refine_2d: sides %d x %d must be odd and >= 3 (cost_based_refinement.h:180)", n0, n1);
	if (kernel < 0 || kernel > 2) return fail(c, STEVI_E_ARG, "refine_2d: unknown interpolation kernel");
	if (isotropy != STEVI_ISOTROPIC && isotropy != STEVI_ANISOTROPIC) return fail(c, STEVI_E_ARG, "refine_2d: unknown isotropy hypothesis");
	if (raw.ndim != 3 || raw.dtype != STEVI_I32 || raw.shape[0] != H || raw.shape[1] != W || raw.shape[2] != 2)
		return fail(c, STEVI_E_SHAPE, "refine_2d: raw disparity must be (%d,%d,2) i32", H, W);
	if (out.ndim != 3 || out.dtype != STEVI_F32 || out.shape[0] != H || out.shape[1] != W || out.shape[2] != 2)
		return fail(c, STEVI_E_SHAPE, "refine_2d: out must be (%d,%d,2) f32", H, W);
	if (H > 65535) return fail(c, STEVI_E_UNSUPPORTED, "refine_2d: more than 65535 rows");
	if (H == 0 || W == 0) return STEVI_OK;
	dim3 grid(cdiv(W, 128), H);
	if (isotropy == STEVI_ANISOTROPIC) {
		STEVI_LAUNCH(c, k_refine_2d<true>, grid, 128, 0, tcv.as<float>(), tcv.st[0], tcv.st[1], tcv.st[2], tcv.st[3], raw.as<int32_t>(), raw.st[0],
					 raw.st[1], raw.st[2], H, W, n0, n1, kernel, out.as<float>(), out.st[0], out.st[1], out.st[2]);
	} else {
		STEVI_LAUNCH(c, k_refine_2d<false>, grid, 128, 0, tcv.as<float>(), tcv.st[0], tcv.st[1], tcv.st[2], tcv.st[3], raw.as<int32_t>(), raw.st[0],
					 raw.st[1], raw.st[2], H, W, n0, n1, kernel, out.as<float>(), out.st[0], out.st[1], out.st[2]);
	}
	return STEVI_OK;
}

int32_t refine2d_patch_dev(stevi_ctx* c, const View& tcv, const View& raw, int kernel, const View& out) {
	if (tcv.ndim != 4 || tcv.dtype != STEVI_F32) return fail(c, STEVI_E_DTYPE, "refine_2d_patch: truncated cost volume must be (H,W,T0,T1) f32");
	const int H = tcv.shape[0], W = tcv.shape[1], n0 = tcv.shape[2], n1 = tcv.shape[3];
	const int r0 = (n0 - 1) / 2, r1 = (n1 - 1) / 2;
	if (r0 < 1 || r1 < 1 || 2 * r0 + 1 != n0 || 2 * r1 + 1 != n1)   // cost_based_refinement.h:386-388
		return fail(c, STEVI_E_SHAPE, "refine_2d_patch: sides %d x %d must be odd and >= 3 (cost_based_refinement.h:386)", n0, n1);
	// :82: the reference static_asserts Parabola or Gaussian
	if (kernel != STEVI_PARABOLA && kernel != STEVI_GAUSSIAN) return fail(c, STEVI_E_ARG, "refine_2d_patch: Parabola or Gaussian kernel (cost_based_refinement.h:82)");
	if (raw.ndim != 3 || raw.dtype != STEVI_I32 || raw.shape[0] != H || raw.shape[1] != W || raw.shape[2] != 2)
		return fail(c, STEVI_E_SHAPE, "refine_2d_patch: raw disparity must be (%d,%d,2) i32", H, W);
	if (out.ndim != 3 || out.dtype != STEVI_F32 || out.shape[0] != H || out.shape[1] != W || out.shape[2] != 2)
		return fail(c, STEVI_E_SHAPE, "refine_2d_patch: out must be (%d,%d,2) f32", H, W);
	if (H > 65535) return fail(c, STEVI_E_UNSUPPORTED, "refine_2d_patch: more than 65535 rows");
	if (H == 0 || W == 0) return STEVI_OK;
	dim3 grid(cdiv(W, 128), H);
	STEVI_LAUNCH(c, k_refine_2d_patch, grid, 128, 0, tcv.as<float>(), tcv.st[0], tcv.st[1], tcv.st[2], tcv.st[3], raw.as<int32_t>(), raw.st[0],
				 raw.st[1], raw.st[2], H, W, r0, r1, kernel == STEVI_GAUSSIAN ? 1 : 0, out.as<float>(), out.st[0], out.st[1], out.st[2]);
	return STEVI_OK;
}

int32_t refine_from_cv_dev(stevi_ctx* c, const View& cv, const View& idx, int hr, int vr, int kernel, const View* tcv_out, const View& out) {
	if (cv.ndim != 3 || cv.dtype != STEVI_F32) return fail(c, STEVI_E_DTYPE, "refine: cost volume must be (H,W,D) f32");
	const int H = cv.shape[0], W = cv.shape[1];
	STEVI_TRY(check_idx(c, idx, H, W, "refine"));
	if (kernel < 0 || kernel > 2) return fail(c, STEVI_E_ARG, "refine: unknown interpolation kernel");
	if (out.ndim != 2 || out.dtype != STEVI_F32 || out.shape[0] != H || out.shape[1] != W) return fail(c, STEVI_E_SHAPE, "refine: out must be (H,W) f32");
	if (tcv_out && (tcv_out->ndim != 3 || tcv_out->dtype != STEVI_F32 || tcv_out->shape[0] != H || tcv_out->shape[1] != W || tcv_out->shape[2] != 3))
		return fail(c, STEVI_E_SHAPE, "refine: tcv_out must be (H,W,3) f32");
	if (H == 0 || W == 0) return STEVI_OK;
	Tcv p;
	p.cv = cv.as<float>(); p.c0 = cv.st[0]; p.c1 = cv.st[1]; p.c2 = cv.st[2];
	p.idx = idx.as<int32_t>(); p.i0 = idx.st[0]; p.i1 = idx.st[1];
	p.H = H; p.W = W; p.D = cv.shape[2]; p.hr = hr; p.vr = vr; p.r = 1; p.sgn = 0; p.mode = STEVI_TCV_SAME; p.mask = 0;
	p.out = tcv_out ? tcv_out->as<float>() : nullptr;
	p.o0 = tcv_out ? tcv_out->st[0] : 0; p.o1 = tcv_out ? tcv_out->st[1] : 0; p.o2 = tcv_out ? tcv_out->st[2] : 0;
	dim3 grid(cdiv(W, 128), H);
	STEVI_LAUNCH(c, k_refine_from_cv, grid, 128, 0, p, kernel, out.as<float>(), out.st[0], out.st[1]);
	return STEVI_OK;
}

} // namespace stevi
