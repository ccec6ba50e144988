// Winner neighbourhood extraction and cost-based subpixel refinement.
//
//   correlation_base.h:557-577   selectedCost: out(i,j) = cv(i,j,idx(i,j))
//   correlation_base.h:579-674   truncatedCostVolume<T_CV,dir,sdir>: the 2r+1 (Same, Reversed) or 4r+1 (Both) costs around the
//                                winner, NaN where the disparity is out of range or the matching window leaves the image
//   correlation_base.h:727-822   extractInBoundDomain: the same tests as a 0/1 mask
//   cost_based_refinement.h:43-69, :128-163   refineCostTriplet / refineDispCostInterpolation: idx + delta(c-1, c0, c+1)
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
