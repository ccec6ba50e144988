// Box filters of correlation/correlation_base.h: meanFilter2D (:881-1004) and sigmaFilter (:1006-1077), on (H,W) or (H,W,C)
// u8 | f32 images, results (H,W) f32.
//
// meanFilter2D is two passes of RUNNING sums in float — along each row mH(i,j) = mH(i,j-1) + (in(i,j+hr) - in(i,j-hr-1)) (one +=
// per channel for 3-D images), then along each column of mH — followed by a division by the box size. Every value therefore
// carries the rounding of the whole chain before it, and the only way to get the reference's bits is to walk the chains in its
// order: H independent chains of W steps, then W chains of H steps. The kernels keep the chains sequential and make everything
// around them parallel:
//   k_mean_chain  a block owns 32 chains; one warp walks them (lane = chain, one dependent addition per step, operands and
//                 results in shared memory) while twelve helper warps load the operands three tiles ahead, coalesced, and write
//                 the results back, coalesced (the row pass transposes through shared memory). Multi-channel rows are walked as
//                 the flattened sequence (j, c). The column pass stores mean = sum / box (the reference divides after the
//                 sums are complete, :928-934: same value).
// The reference never writes the border of its result (outside [vr, H-vr) x [hr, W-hr): uninitialised memory); it is set to 0 here.
//
// sigmaFilter: thread per interior pixel, the window read through L1 (neighbouring threads read neighbouring addresses), the sum
// taken rows, then columns, then channels with separately rounded product and sum, sqrtf correctly rounded.
#include "common.cuh"
#include "kernels.cuh"

namespace stevi {

namespace {

template<typename T> __device__ __forceinline__ float ldf(const T* p) { return (float)__ldg(p); }

// One kernel for both passes. A block owns 32 LINES (image rows for the row pass, columns of mH for the column pass) and walks
// them in tiles of 32 STEPS. Warp 0 is the chain: lane = line, one dependent float addition per step, operands and results in
// shared memory. Warps 1..12 are helpers in three groups of four; group g handles the tiles t = g (mod 3): it loads the two
// operands of every step of its tile three tiles ahead of the chain (coalesced: lane = step along the row for the row pass,
// lane = column for the column pass; the loads stay in flight, in registers, for three chain tiles), leaves their difference
// in shared memory one tile ahead, and writes the chain's results of the tile just finished back to global memory (the row pass
// transposes through shared memory, so both sides of it are coalesced; the column pass divides by the box size on the way).
struct ChainParams {
	const void* in; int64_t s0, s1, s2;   // row pass: the image; column pass: mH as (H, W, 1) with s1 = 1
	float* out; int64_t o0, o1;           // row pass: mH (dense); column pass: the mean
	int H, W, C, hr, vr;
	float box;
};

template<typename T, bool ROWS>
__global__ void __launch_bounds__(416) k_mean_chain(const ChainParams p) {
	__shared__ float delta[2][32][33];
	__shared__ float sums[2][32][33];
	const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const int l0 = blockIdx.x * 32;                 // first line of the block
	const T* in = static_cast<const T*>(p.in);
	const int C = ROWS ? p.C : 1, hr = p.hr, vr = p.vr;
	const int n_lines = ROWS ? p.H : p.W;
	// steps: row pass f = (column - hr - 1) * C + channel; column pass s = row - vr - 1
	const int64_t n_steps = ROWS ? (int64_t)(p.W - 2 * hr - 1) * C : (int64_t)(p.H - 2 * vr - 1);
	const int n_tiles = (int)((n_steps + 31) / 32);
	const int group = (warp - 1) >> 2, q = (warp - 1) & 3;   // helpers
	float lead[8], trail[8];

	// operands of tile t for the 8 (row pass: lines; column pass: steps) this helper warp owns
	auto issue = [&](int t) {
		if (t >= n_tiles) return;
		if (ROWS) {
			const int64_t f = (int64_t)t * 32 + lane;
			const bool live = f < n_steps;
			const int j = hr + 1 + (int)(f / C), c = (int)(f - (int64_t)(j - hr - 1) * C);
			const T* a = in + (int64_t)(j + hr) * p.s1 + (int64_t)c * p.s2;
			const T* b = in + (int64_t)(j - hr - 1) * p.s1 + (int64_t)c * p.s2;
#pragma unroll
			for (int k = 0; k < 8; k++) {
				const int i = l0 + q * 8 + k;
				const bool ok = live && i < p.H;
				lead[k] = ok ? ldf(a + i * p.s0) : 0.f;
				trail[k] = ok ? ldf(b + i * p.s0) : 0.f;
			}
		} else {
			const int j = l0 + lane;
			const bool live = j >= hr && j < p.W - hr;
#pragma unroll
			for (int k = 0; k < 8; k++) {
				const int64_t s = (int64_t)t * 32 + q * 8 + k;
				const bool ok = live && s < n_steps;
				lead[k] = ok ? ldf(in + (s + 2 * vr + 1) * p.s0 + j) : 0.f;
				trail[k] = ok ? ldf(in + s * p.s0 + j) : 0.f;
			}
		}
	};
	// the difference is taken in the image's own type first (:908, :924): exact for u8, the same float subtraction for f32
	auto land = [&](int t) {
		if (t >= n_tiles) return;
#pragma unroll
		for (int k = 0; k < 8; k++) delta[t & 1][q * 8 + k][lane] = __fsub_rn(lead[k], trail[k]);
	};
	// results of tile t: sums[t & 1] holds, row pass [line][step], column pass [step][line]
	auto write_back = [&](int t) {
		if (ROWS) {
			const int64_t f = (int64_t)t * 32 + lane;
			const int j = hr + 1 + (int)(f / C);
			const bool closes = f < n_steps && f - (int64_t)(j - hr - 1) * C == C - 1;   // the column's last channel
#pragma unroll
			for (int k = 0; k < 8; k++) {
				const int i = l0 + q * 8 + k;
				if (closes && i < p.H) p.out[(int64_t)i * p.o0 + j] = sums[t & 1][q * 8 + k][lane];
			}
		} else {
			const int j = l0 + lane;
			const bool live = j >= hr && j < p.W - hr;
#pragma unroll
			for (int k = 0; k < 8; k++) {
				const int64_t s = (int64_t)t * 32 + q * 8 + k;
				if (j < p.W && s < n_steps) p.out[(s + vr + 1) * p.o0 + j * p.o1] = live ? __fdiv_rn(sums[t & 1][q * 8 + k][lane], p.box) : 0.f;
			}
		}
	};
	float m = 0.f;
	if (warp == 0) {
		// the head of the chain: the first full window (:899-903, :916-920)
		const int line = l0 + lane;
		if (ROWS) {
			if (line < p.H) {
				const T* row = in + line * p.s0;
				for (int j = 0; j <= 2 * hr; j++)
					for (int c = 0; c < C; c++) m = __fadd_rn(m, ldf(row + (int64_t)j * p.s1 + (int64_t)c * p.s2));
				p.out[(int64_t)line * p.o0 + hr] = m;
			}
		} else if (line < p.W) {
			const bool live = line >= hr && line < p.W - hr;
			if (live) for (int i = 0; i <= 2 * vr; i++) m = __fadd_rn(m, ldf(in + i * p.s0 + line));
			p.out[(int64_t)vr * p.o0 + line * p.o1] = live ? __fdiv_rn(m, p.box) : 0.f;
		}
	} else {
		if (!ROWS && l0 + lane < p.W)   // rows above and below the interior: never written by the reference, 0 here
			for (int i = warp - 1; i < vr; i += 12) {
				p.out[(int64_t)i * p.o0 + (l0 + lane) * p.o1] = 0.f;
				p.out[(int64_t)(p.H - 1 - i) * p.o0 + (l0 + lane) * p.o1] = 0.f;
			}
		issue(group);
		if (group == 0) { land(0); issue(3); }
	}
	__syncthreads();
	for (int t = 0; t < n_tiles; t++) {
		if (warp == 0) {
			const int b = t & 1;
#pragma unroll
			for (int x = 0; x < 32; x++) {
				m = __fadd_rn(m, ROWS ? delta[b][lane][x] : delta[b][x][lane]);
				if (ROWS) sums[b][lane][x] = m; else sums[b][x][lane] = m;
			}
		} else {
			if (group == (t + 1) % 3) { land(t + 1); issue(t + 4); }
			else if (group == (t + 2) % 3 && t > 0) write_back(t - 1);
		}
		__syncthreads();
	}
	if (warp != 0 && group == 0 && n_tiles > 0) write_back(n_tiles - 1);
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
	const T* row = static_cast<const T*>(p.in) + (int64_t)(i - p.vr) * p.s0 + (int64_t)(j - p.hr) * p.s1;
	const int nh = 2 * p.hr + 1, nv = 2 * p.vr + 1;
	float s = 0.f;
	for (int k = 0; k < nv; k++, row += p.s0) {
		const T* px = row;
		if (p.C == 1) {
#pragma unroll 4
			for (int l = 0; l < nh; l++, px += p.s1) {
				const float tmp = __fsub_rn(ldf(px), m);
				s = __fadd_rn(s, __fmul_rn(tmp, tmp));
			}
		} else {
			for (int l = 0; l < nh; l++, px += p.s1) {
				const T* pc = px;
				for (int c = 0; c < p.C; c++, pc += p.s2) {
					const float tmp = __fsub_rn(ldf(pc), m);
					s = __fadd_rn(s, __fmul_rn(tmp, tmp));
				}
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
	ChainParams r;
	r.in = img.p; r.s0 = img.st[0]; r.s1 = img.st[1]; r.s2 = img.ndim == 3 ? img.st[2] : 0;
	r.out = mh; r.o0 = W; r.o1 = 1;
	r.H = H; r.W = W; r.C = C; r.hr = hr; r.vr = vr; r.box = 1.f;
	if (img.dtype == STEVI_U8) STEVI_LAUNCH(c, (k_mean_chain<uint8_t, true>), cdiv(H, 32), 416, 0, r);
	else STEVI_LAUNCH(c, (k_mean_chain<float, true>), cdiv(H, 32), 416, 0, r);
	ChainParams q = r;
	q.in = mh; q.s0 = W; q.s1 = 1; q.s2 = 0; q.C = 1;
	q.out = out.as<float>(); q.o0 = out.st[0]; q.o1 = out.st[1];
	q.box = (float)((size_t)(2 * vr + 1) * (size_t)(2 * hr + 1) * (size_t)(img.ndim == 3 ? C : 1));   // size_t -> float (:932, :998)
	STEVI_LAUNCH(c, (k_mean_chain<float, false>), cdiv(W, 32), 416, 0, q);
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
