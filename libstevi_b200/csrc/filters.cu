// Box filters of correlation/correlation_base.h: meanFilter2D (:881-1004) and sigmaFilter (:1006-1077), on (H,W) or (H,W,C)
// u8 | f32 images, results (H,W) f32.
//
// meanFilter2D is two passes of RUNNING sums in float — along each row mH(i,j) = mH(i,j-1) + (in(i,j+hr) - in(i,j-hr-1)) (one +=
// per channel for 3-D images), then along each column of mH — followed by a division by the box size. Every value therefore
// carries the rounding of the whole chain before it, and the only way to get the reference's bits is to walk the chains in its
// order: H independent chains of W steps, then W chains of H steps. The kernels keep the chains sequential and make everything
// around them parallel:
//   k_mean_rows   block = 32 rows, two warps: warp 1 fetches the next 32 steps of all 32 rows (coalesced along the row) and leaves
//                 their differences in shared memory, warp 0 (lane = row) walks the previous 32 steps and writes a 32x32 tile of mH
//                 back coalesced. Multi-channel rows are walked as the flattened sequence (j, c).
//   k_mean_cols   thread = column (coalesced by construction), 16 steps of operands prefetched ahead of the chain; stores
//                 mean = sum / box (the reference divides after the sums are complete, :928-934: same value).
// The reference never writes the border of its result (outside [vr, H-vr) x [hr, W-hr): uninitialised memory); it is set to 0 here.
//
// sigmaFilter: thread per interior pixel, the window read through L1 (neighbouring threads read neighbouring addresses), the sum
// taken rows, then columns, then channels with separately rounded product and sum, sqrtf correctly rounded.
#include "common.cuh"
#include "kernels.cuh"

namespace stevi {

namespace {

template<typename T> __device__ __forceinline__ float ldf(const T* p) { return (float)__ldg(p); }

struct RowsParams {
	const void* in; int64_t s0, s1, s2;
	int H, W, C, hr;
	float* mh;   // dense (H, W)
};

template<typename T>
__global__ void __launch_bounds__(64) k_mean_rows(const RowsParams p) {
	__shared__ float delta[2][32][33];   // [buffer][row][step]
	__shared__ float tile[32][33];       // mH values of the 32 steps just walked
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const int r0 = blockIdx.x * 32;
	const T* in = static_cast<const T*>(p.in);
	const int C = p.C, hr = p.hr;
	// the chain of a row: mH(i,hr) = sum of the first 2hr+1 pixels, then steps f = 0 .. n_steps-1, step f adds channel f % C of
	// column j = hr + 1 + f / C; mH(i,j) is complete after the last channel of j
	const int64_t n_steps = (int64_t)(p.W - 2 * hr - 1) * C;
	const int n_tiles = (int)((n_steps + 31) / 32);
	auto fetch = [&](int t, int buf) {   // warp 1: differences of steps [32t, 32t+32) of the 32 rows
		const int64_t f = (int64_t)t * 32 + lane;
		const bool live = f < n_steps;
		const int j = hr + 1 + (int)(f / C), c = (int)(f % C);
		const T* lead = in + (int64_t)(j + hr) * p.s1 + (int64_t)c * p.s2;
		const T* trail = in + (int64_t)(j - hr - 1) * p.s1 + (int64_t)c * p.s2;
#pragma unroll 8
		for (int r = 0; r < 32; r++) {
			const int i = r0 + r;
			float d = 0.f;
			// the difference is taken in the image's own type first (:908): exact for u8, the same float subtraction for f32
			if (live && i < p.H) d = __fsub_rn(ldf(lead + i * p.s0), ldf(trail + i * p.s0));
			delta[buf][r][lane] = d;
		}
	};
	float m = 0.f;
	const int row = r0 + lane;
	if (warp == 0) {
		if (row < p.H) {
			const T* q = in + row * p.s0;
			for (int j = 0; j <= 2 * hr; j++)
				for (int c = 0; c < C; c++) m = __fadd_rn(m, ldf(q + (int64_t)j * p.s1 + (int64_t)c * p.s2));
			p.mh[(int64_t)row * p.W + hr] = m;
		}
	} else if (n_tiles > 0) fetch(0, 0);
	__syncthreads();
	for (int t = 0; t < n_tiles; t++) {
		const int buf = t & 1;
		if (warp == 1) {
			if (t + 1 < n_tiles) fetch(t + 1, buf ^ 1);
		} else {
			const int64_t f0 = (int64_t)t * 32;
#pragma unroll 8
			for (int x = 0; x < 32; x++) {
				m = __fadd_rn(m, delta[buf][lane][x]);
				tile[lane][x] = m;
			}
			__syncwarp();
			// coalesced write-back: lane = step; a step closes a column when it is the column's last channel
			const int64_t f = f0 + lane;
			const bool closes = f < n_steps && (int)(f % C) == C - 1;
			const int j = hr + 1 + (int)(f / C);
			for (int r = 0; r < 32; r++) {
				const int i = r0 + r;
				if (closes && i < p.H) p.mh[(int64_t)i * p.W + j] = tile[r][lane];
			}
			__syncwarp();
		}
		__syncthreads();
	}
}

struct ColsParams {
	const float* mh;
	float* out; int64_t o0, o1;
	int H, W, hr, vr;
	float box;
};

__global__ void __launch_bounds__(32) k_mean_cols(const ColsParams p) {
	const int j = blockIdx.x * 32 + threadIdx.x;
	if (j >= p.W) return;
	float* o = p.out + (int64_t)j * p.o1;
	const int H = p.H, vr = p.vr;
	if (j < p.hr || j >= p.W - p.hr) {   // border columns
		for (int i = 0; i < H; i++) o[(int64_t)i * p.o0] = 0.f;
		return;
	}
	for (int i = 0; i < vr; i++) { o[(int64_t)i * p.o0] = 0.f; o[(int64_t)(H - 1 - i) * p.o0] = 0.f; }
	const float* q = p.mh + j;
	const int64_t W = p.W;
	float m = 0.f;
	for (int i = 0; i <= 2 * vr; i++) m = __fadd_rn(m, q[i * W]);
	o[(int64_t)vr * p.o0] = __fdiv_rn(m, p.box);
	constexpr int B = 16;
	float d[B], nd[B];
	const int first = vr + 1, last = H - vr;   // steps i in [first, last)
	auto load = [&](int i0, float (&v)[B]) {
#pragma unroll
		for (int k = 0; k < B; k++) {
			const int i = i0 + k;
			v[k] = i < last ? __fsub_rn(q[(int64_t)(i + vr) * W], q[(int64_t)(i - vr - 1) * W]) : 0.f;
		}
	};
	load(first, d);
	for (int i0 = first; i0 < last; i0 += B) {
		load(i0 + B, nd);
#pragma unroll
		for (int k = 0; k < B; k++) {
			const int i = i0 + k;
			if (i < last) {
				m = __fadd_rn(m, d[k]);
				o[(int64_t)i * p.o0] = __fdiv_rn(m, p.box);
			}
		}
#pragma unroll
		for (int k = 0; k < B; k++) d[k] = nd[k];
	}
}

struct SigmaParams {
	const void* in; int64_t s0, s1, s2;
	const float* mean; int64_t m0, m1;
	float* out; int64_t o0, o1;
	int H, W, C, hr, vr;
};

template<typename T>
__global__ void __launch_bounds__(256) k_sigma_filter(const SigmaParams p) {
	const int j = blockIdx.x * 32 + threadIdx.x, i = blockIdx.y * 8 + threadIdx.y;
	if (i >= p.H || j >= p.W) return;
	float* o = p.out + i * p.o0 + j * p.o1;
	if (i < p.vr || i >= p.H - p.vr || j < p.hr || j >= p.W - p.hr) { *o = 0.f; return; }
	const float m = __ldg(p.mean + i * p.m0 + j * p.m1);
	const T* in = static_cast<const T*>(p.in);
	float s = 0.f;
	for (int k = -p.vr; k <= p.vr; k++) {
		const T* row = in + (int64_t)(i + k) * p.s0;
		for (int l = -p.hr; l <= p.hr; l++) {
			const T* px = row + (int64_t)(j + l) * p.s1;
			for (int c = 0; c < p.C; c++) {
				const float tmp = __fsub_rn(ldf(px + (int64_t)c * p.s2), m);
				s = __fadd_rn(s, __fmul_rn(tmp, tmp));
			}
		}
	}
	*o = __fsqrt_rn(s);
}

int32_t image_check(stevi_ctx* c, const char* who, const View& img, int hr, int vr, const View& out, int& C) {
	if ((img.ndim != 2 && img.ndim != 3) || (img.dtype != STEVI_U8 && img.dtype != STEVI_F32))
		return fail(c, STEVI_E_DTYPE, "%s: the image must be (H,W) or (H,W,C) u8|f32", who);
	if (hr < 0 || vr < 0 || hr > 255 || vr > 255) return fail(c, STEVI_E_ARG, "%s: radius out of range (uint8_t in the reference)", who);
	C = img.ndim == 3 ? img.shape[2] : 1;
	if (out.ndim != 2 || out.dtype != STEVI_F32 || out.shape[0] != img.shape[0] || out.shape[1] != img.shape[1])
		return fail(c, STEVI_E_SHAPE, "%s: out must be (%d,%d) f32", who, img.shape[0], img.shape[1]);
	// the reference indexes column h_radius and row v_radius unconditionally (:899, :916): a window that does not fit is out of
	// bounds there; refused here
	if (img.shape[1] < 2 * hr + 1 || img.shape[0] < 2 * vr + 1 || C < 1)
		return fail(c, STEVI_E_SHAPE, "%s: the %dx%d window does not fit the (%d,%d) image", who, 2 * hr + 1, 2 * vr + 1, img.shape[0], img.shape[1]);
	return STEVI_OK;
}

} // namespace

int32_t mean_filter_dev(stevi_ctx* c, const View& img, int hr, int vr, const View& out) {
	int C;
	STEVI_TRY(image_check(c, "mean_filter_2d", img, hr, vr, out, C));
	const int H = img.shape[0], W = img.shape[1];
	float* mh = nullptr;
	STEVI_TRY(arena_alloc_t(c, (size_t)H * W, &mh));
	RowsParams r;
	r.in = img.p; r.s0 = img.st[0]; r.s1 = img.st[1]; r.s2 = img.ndim == 3 ? img.st[2] : 0;
	r.H = H; r.W = W; r.C = C; r.hr = hr; r.mh = mh;
	if (img.dtype == STEVI_U8) STEVI_LAUNCH(c, k_mean_rows<uint8_t>, cdiv(H, 32), 64, 0, r);
	else STEVI_LAUNCH(c, k_mean_rows<float>, cdiv(H, 32), 64, 0, r);
	ColsParams q;
	q.mh = mh; q.out = out.as<float>(); q.o0 = out.st[0]; q.o1 = out.st[1];
	q.H = H; q.W = W; q.hr = hr; q.vr = vr;
	q.box = (float)((size_t)(2 * vr + 1) * (size_t)(2 * hr + 1) * (size_t)(img.ndim == 3 ? C : 1));   // size_t -> float (:932, :998)
	STEVI_LAUNCH(c, k_mean_cols, cdiv(W, 32), 32, 0, q);
	return STEVI_OK;
}

int32_t sigma_filter_dev(stevi_ctx* c, const View& mean, const View& img, int hr, int vr, const View& out) {
	int C;
	STEVI_TRY(image_check(c, "sigma_filter", img, hr, vr, out, C));
	if (mean.ndim != 2 || mean.dtype != STEVI_F32 || mean.shape[0] != img.shape[0] || mean.shape[1] != img.shape[1])
		return fail(c, STEVI_E_SHAPE, "sigma_filter: mean must be (%d,%d) f32", img.shape[0], img.shape[1]);
	const int H = img.shape[0], W = img.shape[1];
	if (cdiv(H, 8) > 65535) return fail(c, STEVI_E_UNSUPPORTED, "sigma_filter: more than 524280 rows");
	SigmaParams p;
	p.in = img.p; p.s0 = img.st[0]; p.s1 = img.st[1]; p.s2 = img.ndim == 3 ? img.st[2] : 0;
	p.mean = mean.as<float>(); p.m0 = mean.st[0]; p.m1 = mean.st[1];
	p.out = out.as<float>(); p.o0 = out.st[0]; p.o1 = out.st[1];
	p.H = H; p.W = W; p.C = C; p.hr = hr; p.vr = vr;
	dim3 grid(cdiv(W, 32), cdiv(H, 8)), block(32, 8);
	if (img.dtype == STEVI_U8) STEVI_LAUNCH(c, k_sigma_filter<uint8_t>, grid, block, 0, p);
	else STEVI_LAUNCH(c, k_sigma_filter<float>, grid, block, 0, p);
	return STEVI_OK;
}

} // namespace stevi
