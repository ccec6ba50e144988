// Guided cost-volume aggregation.
//
//   guided_cost_filtering.h:102-176   variableCostVolumeAggregation: out(i,j,c) = sum_w w(i,j,di,dj) cv(i+di,j+dj,c) / sum_w w(i,j,di,dj)
//                                     over the in-image part of the pixel's own (2v+1) x (2h+1) window of weights
//
// The weights come from computeAggregationWindows (:26-100), which evaluates two caller-supplied functors per window cell:
// that is the caller's code and stays on the caller's side; this file is the volume-sized pass.
//
// Thread per pixel, a block per (row, 128 columns, slice of the disparity axis). A thread keeps its pixel's weights in
// registers (compile-time radii 1..3) and walks its disparities; for the reference's volume layout {W*D, 1, W} the loads of
// a warp are 128 contiguous bytes for every window cell. Sums follow the reference's order (window rows outer, columns
// inner, a rounded multiply and a rounded add per cell: its Release build does not contract `acc += w*cv`), so the result is
// bit-identical.
#include <cstdlib>
#include <type_traits>
#include "common.cuh"
#include "kernels.cuh"

namespace stevi {

namespace {

struct Agg {
	const float* cv; int64_t c0, c1, c2;
	const float* win; int64_t w0, w1, w2, w3;
	float* out; int64_t o0, o1, o2;
	int H, W, D, vr, hr, chunk, vec_out;
};

template<int VR, int HR>
__global__ void __launch_bounds__(128) k_var_aggregate(const Agg p) {
	const int j = blockIdx.x * 128 + threadIdx.x, i = blockIdx.y;
	if (j >= p.W) return;
	const int lo_i = max(-VR, -i), hi_i = min(VR, p.H - 1 - i);
	const int lo_j = max(-HR, -j), hi_j = min(HR, p.W - 1 - j);
	float w[2 * VR + 1][2 * HR + 1];
	float weight = 0.f;
	const float* wp = p.win + i * p.w0 + j * p.w1;
#pragma unroll
	for (int di = -VR; di <= VR; di++)
#pragma unroll
		for (int dj = -HR; dj <= HR; dj++) {
			const bool in = di >= lo_i && di <= hi_i && dj >= lo_j && dj <= hi_j;
			const float x = in ? wp[(di + VR) * p.w2 + (dj + HR) * p.w3] : 0.f;
			w[di + VR][dj + HR] = x;
			if (in) weight = __fadd_rn(weight, x);
		}
	const bool inner = lo_i == -VR && hi_i == VR && lo_j == -HR && hi_j == HR;
	const int c_begin = blockIdx.z * p.chunk, c_end = min(p.D, c_begin + p.chunk);
	const float* src = p.cv + i * p.c0 + j * p.c1;
	float* dst = p.out + i * p.o0 + j * p.o1;
	if (inner && p.c1 == 1) {
		// contiguous columns: one pointer per window row advanced by the disparity stride, column offsets as immediates
		const float* rowp[2 * VR + 1];
#pragma unroll
		for (int a = 0; a <= 2 * VR; a++) rowp[a] = src + (int64_t)(a - VR) * p.c0 + (int64_t)c_begin * p.c2;
		float* o = dst + (int64_t)c_begin * p.o2;
		for (int c = c_begin; c < c_end; c++) {
			float acc = 0.f;
#pragma unroll
			for (int a = 0; a <= 2 * VR; a++) {
#pragma unroll
				for (int dj = -HR; dj <= HR; dj++) acc = __fadd_rn(acc, __fmul_rn(w[a][dj + HR], __ldg(rowp[a] + dj)));
				rowp[a] += p.c2;
			}
			*o = __fdiv_rn(acc, weight);
			o += p.o2;
		}
	} else if (inner) {
		for (int c = c_begin; c < c_end; c++) {
			const float* s = src + c * p.c2;
			float acc = 0.f;
#pragma unroll
			for (int di = -VR; di <= VR; di++)
#pragma unroll
				for (int dj = -HR; dj <= HR; dj++) acc = __fadd_rn(acc, __fmul_rn(w[di + VR][dj + HR], s[di * p.c0 + dj * p.c1]));
			dst[c * p.o2] = __fdiv_rn(acc, weight);
		}
	} else {
		for (int c = c_begin; c < c_end; c++) {
			const float* s = src + c * p.c2;
			float acc = 0.f;
#pragma unroll
			for (int di = -VR; di <= VR; di++)
#pragma unroll
				for (int dj = -HR; dj <= HR; dj++)
					if (di >= lo_i && di <= hi_i && dj >= lo_j && dj <= hi_j) acc = __fadd_rn(acc, __fmul_rn(w[di + VR][dj + HR], s[di * p.c0 + dj * p.c1]));
			dst[c * p.o2] = __fdiv_rn(acc, weight);
		}
	}
}

// any radii, one pixel: the weights are re-read per disparity (they stay in L1)
__device__ __forceinline__ void aggregate_pixel(const Agg& p, int i, int j, int c_begin, int c_end) {
	const int lo_i = max(-p.vr, -i), hi_i = min(p.vr, p.H - 1 - i);
	const int lo_j = max(-p.hr, -j), hi_j = min(p.hr, p.W - 1 - j);
	const float* wp = p.win + i * p.w0 + j * p.w1 + p.vr * p.w2 + p.hr * p.w3;
	float weight = 0.f;
	for (int di = lo_i; di <= hi_i; di++)
		for (int dj = lo_j; dj <= hi_j; dj++) weight = __fadd_rn(weight, wp[di * p.w2 + dj * p.w3]);
	for (int c = c_begin; c < c_end; c++) {
		const float* s = p.cv + i * p.c0 + j * p.c1 + c * p.c2;
		float acc = 0.f;
		for (int di = lo_i; di <= hi_i; di++)
			for (int dj = lo_j; dj <= hi_j; dj++) acc = __fadd_rn(acc, __fmul_rn(wp[di * p.w2 + dj * p.w3], s[di * p.c0 + dj * p.c1]));
		p.out[i * p.o0 + j * p.o1 + c * p.o2] = __fdiv_rn(acc, weight);
	}
}

__global__ void __launch_bounds__(128) k_var_aggregate_any(const Agg p) {
	const int j = blockIdx.x * 128 + threadIdx.x, i = blockIdx.y;
	if (j >= p.W) return;
	const int c_begin = blockIdx.z * p.chunk;
	aggregate_pixel(p, i, j, c_begin, min(p.D, c_begin + p.chunk));
}

// Volumes whose columns are contiguous and 16-byte aligned (the reference's layout {W*D, 1, W} with W % 4 == 0): a thread owns
// four adjacent columns, keeps their 4 x (2R+1)^2 weights in registers and reads each window row of a disparity as three
// aligned 16-byte loads (columns j-4 .. j+7 cover the four windows for R <= 3): 3 wide loads per window row for 4 results
// instead of 4 (2R+1) scalar ones, and the result leaves as one 16-byte store. Pixels whose window (or whose 16-byte loads)
// would leave the image - the first and last R rows, the first and last four columns - are left to k_var_aggregate_border,
// a thread per (border pixel, disparity): done by their owner threads they would run 4 x 192 clipped windows in sequence on
// one lane and hold half of the blocks of every row back (measured: 2.5x on the whole launch). Same summation order per
// pixel as above.
template<int R>
__global__ void __launch_bounds__(128) k_var_aggregate_x4(const Agg p) {
	const int j = (blockIdx.x * 128 + threadIdx.x) * 4, i = blockIdx.y;
	if (j >= p.W) return;
	const int c_begin = blockIdx.z * p.chunk, c_end = min(p.D, c_begin + p.chunk);
	if (i < R || i + R >= p.H || j < 4 || j + 8 > p.W) return;   // border pixels: k_var_aggregate_border
	constexpr int N = 2 * R + 1;
	float w[4][N][N], weight[4];
#pragma unroll
	for (int q = 0; q < 4; q++) {
		const float* wp = p.win + i * p.w0 + (j + q) * p.w1;
		weight[q] = 0.f;
#pragma unroll
		for (int a = 0; a < N; a++)
#pragma unroll
			for (int b = 0; b < N; b++) {
				w[q][a][b] = wp[a * p.w2 + b * p.w3];
				weight[q] = __fadd_rn(weight[q], w[q][a][b]);
			}
	}
	// one pointer per window row, advanced by the disparity stride: no 64-bit stride arithmetic inside the loop
	const float4* rowp[N];
#pragma unroll
	for (int a = 0; a < N; a++) rowp[a] = reinterpret_cast<const float4*>(p.cv + (int64_t)(i - R + a) * p.c0 + (int64_t)c_begin * p.c2 + (j - 4));
	float* dst = p.out + i * p.o0 + j * p.o1 + (int64_t)c_begin * p.o2;
	for (int c = c_begin; c < c_end; c++) {
		float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
		for (int a = 0; a < N; a++) {
			const float4 x0 = __ldg(rowp[a]), x1 = __ldg(rowp[a] + 1), x2 = __ldg(rowp[a] + 2);
			rowp[a] = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(rowp[a]) + p.c2);
			const float v[12] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w, x2.x, x2.y, x2.z, x2.w};
#pragma unroll
			for (int b = 0; b < N; b++)
#pragma unroll
				for (int q = 0; q < 4; q++) acc[q] = __fadd_rn(acc[q], __fmul_rn(w[q][a][b], v[4 + q + b - R]));
		}
		float4 r;
		r.x = __fdiv_rn(acc[0], weight[0]); r.y = __fdiv_rn(acc[1], weight[1]);
		r.z = __fdiv_rn(acc[2], weight[2]); r.w = __fdiv_rn(acc[3], weight[3]);
		if (p.vec_out) *reinterpret_cast<float4*>(dst) = r;
		else { dst[0] = r.x; dst[p.o1] = r.y; dst[2 * p.o1] = r.z; dst[3 * p.o1] = r.w; }
		dst += p.o2;
	}
}


// Windows of 5 x 5 and 7 x 7 on volumes with contiguous columns: the d-plane tile a block of 8 rows x 16*CPT columns needs (its
// pixels plus R on every side, zero outside the image) goes through shared memory, double-buffered along the disparity axis — the
// plane of d+1 is fetched into registers before the arithmetic of d and stored after it, one barrier per disparity. A thread owns
// CPT adjacent columns of one row with their CPT x (2R+1)^2 weights in registers and reads each window row of the tile as 16-byte
// (CPT = 4) or 8-byte (CPT = 2) shared-memory loads. What this removes against k_var_aggregate_x4 / k_var_aggregate: the 5- / 7-fold
// re-read of every plane row through L2 by the blocks of neighbouring rows (one block per row there) whose latency 12 resident
// warps could not hide; the tile is read 1.6 - 2.1 times. Weights of window cells outside the image are 0 and their tile
// values 0: the sum takes the same terms in the same order as the reference's clipped loop plus exact zeros. No border launch.
template<int R, int CPT, bool VEC>
__global__ void __launch_bounds__(128, 3) k_var_aggregate_tile(const Agg p) {
	// VEC (16-byte aligned rows, W % 4 == 0): the tile starts four columns left of the block so that it is made of whole 16-byte
	// units, fetched and stored as float4 (1 - 2 per thread and plane instead of 5 - 7 scalars)
	constexpr int N = 2 * R + 1, TXN = 16, TYN = 8;
	constexpr int ORG = VEC ? 4 : R, SHIFT = ORG - R;
	constexpr int TC = TXN * CPT, TR = TYN, TW = TC + 2 * ORG, TH = TR + 2 * R;
	constexpr int PITCH = (TW + 3) / 4 * 4;
	constexpr int EL = VEC ? TH * (TW / 4) : TH * TW;       // elements (float4 or float) of a plane's tile
	constexpr int K = (EL + 127) / 128;
	constexpr int LW = CPT == 4 ? 4 : 2;                     // floats per shared-memory load
	constexpr int NV = (CPT + 2 * R + SHIFT + LW - 1) / LW * LW;
	__shared__ __align__(16) float tile[2][TH * PITCH];
	const int tx = threadIdx.x % TXN, ty = threadIdx.x / TXN;
	const int j0 = blockIdx.x * TC, i0 = blockIdx.y * TR;
	const int j = j0 + tx * CPT, i = i0 + ty;
	const int c_begin = blockIdx.z * p.chunk, c_end = min(p.D, c_begin + p.chunk);
	// the elements of the tile this thread moves: a pointer into the first plane (nullptr outside the image) and the place in the tile
	const float* src[K];
	int sidx[K];
#pragma unroll
	for (int k = 0; k < K; k++) {
		const int e = threadIdx.x + k * 128;
		constexpr int PER_ROW = VEC ? TW / 4 : TW;
		const int r = e / PER_ROW, t = (e - r * PER_ROW) * (VEC ? 4 : 1);
		const int gi = i0 - R + r, gj = j0 - ORG + t;
		const bool in = e < EL && gi >= 0 && gi < p.H && gj >= 0 && gj + (VEC ? 4 : 1) <= p.W;
		src[k] = in ? p.cv + (int64_t)gi * p.c0 + gj + (int64_t)c_begin * p.c2 : nullptr;
		sidx[k] = e < EL ? r * PITCH + t : -1;
	}
	float w[CPT][N][N], weight[CPT];
#pragma unroll
	for (int q = 0; q < CPT; q++) {
		const bool pix = i < p.H && j + q < p.W;
		const float* wp = p.win + (int64_t)(pix ? i : 0) * p.w0 + (int64_t)(pix ? j + q : 0) * p.w1;
		weight[q] = 0.f;
#pragma unroll
		for (int a = 0; a < N; a++)
#pragma unroll
			for (int b = 0; b < N; b++) {
				const int gi = i + a - R, gj = j + q + b - R;
				const bool in = pix && gi >= 0 && gi < p.H && gj >= 0 && gj < p.W;
				const float x = in ? wp[a * p.w2 + b * p.w3] : 0.f;
				w[q][a][b] = x;
				if (in) weight[q] = __fadd_rn(weight[q], x);
			}
	}
	typedef typename std::conditional<VEC, float4, float>::type El;
	auto fetch = [](const float* q) -> El {
		if constexpr (VEC) return q ? __ldg(reinterpret_cast<const float4*>(q)) : make_float4(0.f, 0.f, 0.f, 0.f);
		else return q ? __ldg(q) : 0.f;
	};
#pragma unroll
	for (int k = 0; k < K; k++)
		if (sidx[k] >= 0) *reinterpret_cast<El*>(&tile[0][sidx[k]]) = fetch(src[k]);
	__syncthreads();
	float* dst = p.out + (int64_t)min(i, p.H - 1) * p.o0 + (int64_t)min(j, p.W - 1) * p.o1 + (int64_t)c_begin * p.o2;
	const bool vec = p.vec_out && i < p.H && j + CPT <= p.W;
	for (int c = c_begin; c < c_end; c++) {
		const int buf = (c - c_begin) & 1;
		const bool more = c + 1 < c_end;
		El nx[K];
#pragma unroll
		for (int k = 0; k < K; k++) {
			if (src[k]) src[k] += p.c2;
			nx[k] = fetch(more ? src[k] : nullptr);
		}
		float acc[CPT];
#pragma unroll
		for (int q = 0; q < CPT; q++) acc[q] = 0.f;
		const float* trow = &tile[buf][ty * PITCH + tx * CPT];
#pragma unroll
		for (int a = 0; a < N; a++) {
			float v[NV];
			if (LW == 4) {
#pragma unroll
				for (int u = 0; u < NV / 4; u++) {
					const float4 x = *reinterpret_cast<const float4*>(trow + a * PITCH + 4 * u);
					v[4 * u] = x.x; v[4 * u + 1] = x.y; v[4 * u + 2] = x.z; v[4 * u + 3] = x.w;
				}
			} else {
#pragma unroll
				for (int u = 0; u < NV / 2; u++) {
					const float2 x = *reinterpret_cast<const float2*>(trow + a * PITCH + 2 * u);
					v[2 * u] = x.x; v[2 * u + 1] = x.y;
				}
			}
#pragma unroll
			for (int b = 0; b < N; b++)
#pragma unroll
				for (int q = 0; q < CPT; q++) acc[q] = __fadd_rn(acc[q], __fmul_rn(w[q][a][b], v[SHIFT + q + b]));
		}
		float res[CPT];
#pragma unroll
		for (int q = 0; q < CPT; q++) res[q] = __fdiv_rn(acc[q], weight[q]);
		if (vec) {
			if (CPT == 4) *reinterpret_cast<float4*>(dst) = make_float4(res[0], res[1], res[2], res[3]);
			else *reinterpret_cast<float2*>(dst) = make_float2(res[0], res[1]);
		} else if (i < p.H) {
#pragma unroll
			for (int q = 0; q < CPT; q++)
				if (j + q < p.W) dst[q * p.o1] = res[q];
		}
		dst += p.o2;
		if (more) {
#pragma unroll
			for (int k = 0; k < K; k++)
				if (sidx[k] >= 0) *reinterpret_cast<El*>(&tile[buf ^ 1][sidx[k]]) = nx[k];
		}
		__syncthreads();
	}
}

// the frame of pixels k_var_aggregate_x4 skips: rows [0, top) and [bot, H) in full, columns [0, 4) and [W - 4, W) of the rows
// in between; thread per (pixel, disparity)
__global__ void __launch_bounds__(128) k_var_aggregate_border(const Agg p, int top, int bot) {
	const int n_full = (top + (p.H - bot)) * p.W, n = n_full + (bot - top) * 8;
	const int k = blockIdx.x * 128 + threadIdx.x, c = blockIdx.y;
	if (k >= n) return;
	int i, j;
	if (k < n_full) {
		i = k / p.W; j = k - i * p.W;
		if (i >= top) i += bot - top;
	} else {
		const int m = k - n_full;
		i = top + m / 8; j = m % 8;
		if (j >= 4) j += p.W - 8;
	}
	aggregate_pixel(p, i, j, c, c + 1);
}

} // namespace

int32_t var_aggregate_dev(stevi_ctx* c, const View& cv, const View& win, const View& out) {
	if (cv.ndim != 3 || cv.dtype != STEVI_F32) return fail(c, STEVI_E_DTYPE, "variable_cost_volume_aggregation: cv must be (H,W,D) f32");
	if (win.ndim != 4 || win.dtype != STEVI_F32) return fail(c, STEVI_E_DTYPE, "variable_cost_volume_aggregation: windows must be (H,W,2v+1,2h+1) f32");
	const int H = cv.shape[0], W = cv.shape[1], D = cv.shape[2];
	// guided_cost_filtering.h:108-130: any of these gives an empty array
	if (win.shape[0] != H || win.shape[1] != W || win.shape[2] <= 0 || win.shape[3] <= 0 || win.shape[2] % 2 != 1 || win.shape[3] % 2 != 1)
		return fail(c, STEVI_E_SHAPE, "variable_cost_volume_aggregation: windows (%d,%d,%d,%d) do not fit a (%d,%d,.) volume with odd window sides "
					"(guided_cost_filtering.h:108-130)", win.shape[0], win.shape[1], win.shape[2], win.shape[3], H, W);
	if (out.ndim != 3 || out.dtype != STEVI_F32 || out.shape[0] != H || out.shape[1] != W || out.shape[2] != D)
		return fail(c, STEVI_E_SHAPE, "variable_cost_volume_aggregation: out must be (%d,%d,%d) f32", H, W, D);
	if (H > 65535) return fail(c, STEVI_E_UNSUPPORTED, "variable_cost_volume_aggregation: more than 65535 rows");
	if (H == 0 || W == 0 || D == 0) return STEVI_OK;
	Agg p;
	p.cv = cv.as<float>(); p.c0 = cv.st[0]; p.c1 = cv.st[1]; p.c2 = cv.st[2];
	p.win = win.as<float>(); p.w0 = win.st[0]; p.w1 = win.st[1]; p.w2 = win.st[2]; p.w3 = win.st[3];
	p.out = out.as<float>(); p.o0 = out.st[0]; p.o1 = out.st[1]; p.o2 = out.st[2];
	p.H = H; p.W = W; p.D = D; p.vr = win.shape[2] / 2; p.hr = win.shape[3] / 2;
	const int key = p.vr == p.hr ? p.vr : -1;
	auto aligned = [](const void* q, int64_t a, int64_t b) { return (uintptr_t)q % 16 == 0 && a % 4 == 0 && b % 4 == 0; };
	static const bool no_x4 = getenv("STEVI_AGG_NO_X4") != nullptr;   // measurement aid: the one-column-per-thread kernels only
	const bool x4 = !no_x4 && (key == 1 || key == 2) && cv.st[1] == 1 && aligned(cv.p, cv.st[0], cv.st[2]) && W >= 16 && W % 4 == 0;
	p.vec_out = out.st[1] == 1 && aligned(out.p, out.st[0], out.st[2]) ? 1 : 0;
	static const bool no_tile = getenv("STEVI_AGG_NO_TILE") != nullptr;   // measurement aid: the kernels without the shared-memory tile
	if (!no_tile && (key == 2 || key == 3) && cv.st[1] == 1) {
		const int tc = key == 2 ? 64 : 32;
		const int tiles = cdiv(W, tc) * cdiv(H, 8);
		int slices = std::max(1, std::min(cdiv(D, 16), cdiv(c->sm_count * 3 * 4, tiles)));
		p.chunk = cdiv(D, slices);
		slices = cdiv(D, p.chunk);
		// 16-byte (8-byte) stores of a thread's four (two) results
		p.vec_out = out.st[1] == 1 && (uintptr_t)out.p % 16 == 0 && out.st[0] % 4 == 0 && out.st[2] % 4 == 0 ? 1 : 0;
		dim3 grid(cdiv(W, tc), cdiv(H, 8), slices);
		const bool vec_in = aligned(cv.p, cv.st[0], cv.st[2]) && W % 4 == 0;
		if (key == 2) { if (vec_in) STEVI_LAUNCH(c, (k_var_aggregate_tile<2, 4, true>), grid, 128, 0, p); else STEVI_LAUNCH(c, (k_var_aggregate_tile<2, 4, false>), grid, 128, 0, p); }
		else { if (vec_in) STEVI_LAUNCH(c, (k_var_aggregate_tile<3, 2, true>), grid, 128, 0, p); else STEVI_LAUNCH(c, (k_var_aggregate_tile<3, 2, false>), grid, 128, 0, p); }
		return STEVI_OK;
	}
	// slices of the disparity axis: enough blocks for a few waves over the SMs, each thread amortising its weight loads
	const int col_blocks = x4 ? cdiv(cdiv(W, 4), 128) : cdiv(W, 128);
	const int tiles = col_blocks * H;
	int slices = std::max(1, std::min(cdiv(D, 16), cdiv(c->sm_count * 16 * 4, tiles)));
	p.chunk = cdiv(D, slices);
	slices = cdiv(D, p.chunk);
	dim3 grid(col_blocks, H, slices);
	if (x4) {
		if (key == 1) STEVI_LAUNCH(c, (k_var_aggregate_x4<1>), grid, 128, 0, p);
		else STEVI_LAUNCH(c, (k_var_aggregate_x4<2>), grid, 128, 0, p);
		const int top = std::min(key, H), bot = std::max(H - key, top);
		const int n = (top + (H - bot)) * W + (bot - top) * 8;
		if (D > 65535) return fail(c, STEVI_E_UNSUPPORTED, "variable_cost_volume_aggregation: more than 65535 disparities");
		dim3 gb(cdiv(n, 128), D);
		STEVI_LAUNCH(c, k_var_aggregate_border, gb, 128, 0, p, top, bot);
		return STEVI_OK;
	}
	switch (key) {
	case 1: STEVI_LAUNCH(c, (k_var_aggregate<1, 1>), grid, 128, 0, p); break;
	case 2: STEVI_LAUNCH(c, (k_var_aggregate<2, 2>), grid, 128, 0, p); break;
	case 3: STEVI_LAUNCH(c, (k_var_aggregate<3, 3>), grid, 128, 0, p); break;
	default: STEVI_LAUNCH(c, k_var_aggregate_any, grid, 128, 0, p); break;
	}
	return STEVI_OK;
}

} // namespace stevi
