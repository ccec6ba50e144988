// Semi-global matching path aggregation (correlation/sgm.h:57-404 of the reference).
//
// Two implementations live here.
//
// (1) COLLAPSED — COMPAT semantics, Cost strategy, integer (uint16) cost volume, integer-valued Pout.
//     As written (Q3, sgm.h:281-283) the reference's Cost recurrence stores `c_score` instead of the penalised
//     previous score, so along a line only the scalar m(p) = min_d L(p,d) is handed from pixel to pixel:
//         L(p,d) = C + ((C + Pout*[j+d >= W]) - m(prev)),   m(p) = g(p) - m(prev),   g(p) = min_d (2C + Pout*[j+d >= W])
//         S(p,d) = C + sum over the live line passes r through p of ((C + Pout*[..]) - m_r(prev))
//                = (1 + R_p)*C + R_p*Pout*[j+d >= W] - T_p,      R_p = #passes, T_p = sum_r m_r(prev)
//     Every quantity is an integer; as long as all magnitudes stay below 2^24 the reference's float arithmetic is
//     exact and order independent, so the integer evaluation below is bit-identical to it. The bound is checked
//     (a priori when the caller knows the largest cost, else from the measured max |m|); if it does not hold the
//     general implementation (2) is used instead. Volume traffic: one read of C for g (skipped when the cost-volume
//     kernel already emitted g), one read of C and one write of S.
//
// (2) GENERAL — Score strategy (a genuine P1/P2 recurrence, sgm.h:218-256), float volumes, CORRECTED semantics:
//     one warp per line, lanes strided over d. See the second half of this file.
//
// Live line sets (Q4, sgm.h:321-354): lines start at ZeroPos (top row / left column) or at EndPos which, as written,
// is one past the last valid index, so those lines traverse nothing; the (top,left) corner line of a diagonal is
// started by both loops and therefore counted twice. CORRECTED starts EndPos lines on the last valid index and starts
// corner lines once.
#include "common.cuh"
#include "kernels.cuh"

namespace stevi {

// sgm.h:57-155 directionTraits: {stepsVertical[0], stepsVertical[1], stepsHorizontal[0], stepsHorizontal[1]}
__constant__ int c_sgm_steps[16][4] = {
	{1, 1, 0, 0}, {-1, -1, 0, 0}, {0, 0, 1, 1}, {0, 0, -1, -1},
	{1, 1, 1, 1}, {-1, -1, -1, -1}, {1, 1, -1, -1}, {-1, -1, 1, 1},
	{0, 1, 1, 1}, {0, -1, -1, -1}, {0, 1, -1, -1}, {0, -1, 1, 1},
	{1, 1, 0, 1}, {-1, -1, 0, -1}, {1, 1, 0, -1}, {-1, -1, 0, 1}};
// call order of sgmCostVolume, sgm.h:379-401
__constant__ int c_sgm_order[16] = {0, 1, 2, 3, 4, 5, 6, 7, 12, 13, 14, 15, 8, 9, 10, 11};
static const int h_sgm_steps[16][4] = {
	{1, 1, 0, 0}, {-1, -1, 0, 0}, {0, 0, 1, 1}, {0, 0, -1, -1},
	{1, 1, 1, 1}, {-1, -1, -1, -1}, {1, 1, -1, -1}, {-1, -1, 1, 1},
	{0, 1, 1, 1}, {0, -1, -1, -1}, {0, 1, -1, -1}, {0, -1, 1, 1},
	{1, 1, 0, 1}, {-1, -1, 0, -1}, {1, 1, 0, -1}, {-1, -1, 0, 1}};
static const int h_sgm_order[16] = {0, 1, 2, 3, 4, 5, 6, 7, 12, 13, 14, 15, 8, 9, 10, 11};

struct SgmGeom { int H, W, top, bottom, left, right, corrected; };

// sgm.h:162-184: 0 = NoStart, 1 = ZeroPos, 2 = EndPos
__host__ __device__ __forceinline__ int sgm_start_pos(int s0, int s1) {
	if (s0 == 0 && s1 == 0) return 0;
	if (s0 >= 0 && s1 >= 0) return 1;
	return 2;
}

struct SgmLines { int n_row, n_col, sj_row, si_col; };

// sgm.h:313-356 addDirectionalCost: which lines a direction starts. kind 0 = lines starting on a column (one per row,
// `rowStart`), kind 1 = lines starting on a row (one per column, `colStart`).
__host__ __device__ __forceinline__ SgmLines sgm_lines(const int* st, const SgmGeom& g) {
	SgmLines l;
	const int colStart = sgm_start_pos(st[0], st[1]);
	const int rowStart = sgm_start_pos(st[2], st[3]);
	const int nr = g.H - g.bottom - g.top, nc = g.W - g.right - g.left;
	l.n_row = (rowStart != 0 && nr > 0) ? nr : 0;
	l.n_col = (colStart != 0 && nc > 0) ? nc : 0;
	l.sj_row = rowStart == 1 ? g.left : g.W - g.right - (g.corrected ? 1 : 0);
	l.si_col = colStart == 1 ? g.top : g.H - g.bottom - (g.corrected ? 1 : 0);
	return l;
}

__device__ __forceinline__ bool sgm_inside(int i, int j, const SgmGeom& g) {
	return i >= g.top && i < g.H - g.bottom && j >= g.left && j < g.W - g.right;
}

// ------------------------------------------------------------------------------------------------------------------
// (1) collapsed implementation
// ------------------------------------------------------------------------------------------------------------------

// g(p) = min_d (2*C(p,d) + pout*[j+d >= W]) for volumes with a pixel axis of stride 1: thread = 8 consecutive pixels
__global__ void __launch_bounds__(128) k_sgm_gmap_planes(const uint16_t* __restrict__ cv, int64_t sa, int64_t sd, int na, int nb, int D,
														 int W, int b_is_j, int pout, int32_t* __restrict__ g) {
	const int b0 = (blockIdx.x * 128 + threadIdx.x) * 8;
	const int a = blockIdx.y;
	if (b0 >= nb) return;
	const uint16_t* p = cv + a * sa + b0;
	int gm[8];
#pragma unroll
	for (int k = 0; k < 8; k++) gm[k] = 0x7fffffff;
	if (b0 + 8 <= nb) {
#pragma unroll 8
		for (int d = 0; d < D; d++) {
			const uint4 r = __ldg(reinterpret_cast<const uint4*>(p + d * sd));
			const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
			for (int k = 0; k < 8; k++) {
				const int cst = (w[k >> 1] >> ((k & 1) * 16)) & 0xffff;
				const int j = b_is_j ? b0 + k : a;
				gm[k] = min(gm[k], 2 * cst + (j + d >= W ? pout : 0));
			}
		}
	} else {
		for (int k = 0; b0 + k < nb; k++) {
			const int j = b_is_j ? b0 + k : a;
			for (int d = 0; d < D; d++) gm[k] = min(gm[k], 2 * (int)p[k + d * sd] + (j + d >= W ? pout : 0));
		}
	}
#pragma unroll
	for (int k = 0; k < 8; k++) if (b0 + k < nb) {
		const int i = b_is_j ? a : b0 + k, j = b_is_j ? b0 + k : a;
		g[(int64_t)i * W + j] = gm[k];
	}
}

// any strides: a warp per pixel, lanes strided over d (coalesced when d is the fastest axis)
__global__ void __launch_bounds__(256) k_sgm_gmap_generic(const uint16_t* __restrict__ cv, int64_t s0, int64_t s1, int64_t sd, int H, int W, int D,
														  int pout, int32_t* __restrict__ g) {
	const int lane = threadIdx.x & 31;
	const int64_t pix = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
	if (pix >= (int64_t)H * W) return;
	const int i = (int)(pix / W), j = (int)(pix - (int64_t)i * W);
	const uint16_t* p = cv + i * s0 + j * s1;
	int gm = 0x7fffffff;
	for (int d = lane; d < D; d += 32) gm = min(gm, 2 * (int)p[d * sd] + (j + d >= W ? pout : 0));
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) gm = min(gm, __shfl_xor_sync(0xffffffffu, gm, o));
	if (lane == 0) g[pix] = gm;
}

// Scalar recurrences m(p) = g(p) - m(prev) along every live line: T(p) += m(prev), R(p) += 1 per pass; integer atomics,
// so the result does not depend on the schedule. Three kernels share the work:
//   k_sgm_pathsum_sweep  vertical and diagonal directions: the lanes of a warp sit on 32 adjacent lines at the same image
//                        row (every load and atomic of a warp touches one contiguous segment), the warps of a block
//                        on 32 row segments of those lines, stitched through the linearity of the recurrence;
//   k_sgm_pathsum_rows   horizontal directions: a warp per row, 32 columns per step through a signed warp scan
//                        (m_k = (-1)^k * sum_{t<=k} (-1)^t g_t);
//   k_sgm_pathsum        the generic line walker below, kept for the eight extra directions of the 16-direction set.
// Generic walker: one thread per line, loads of g batched 8 ahead (they do not depend on the recurrence).
// blockIdx.y + slot0 = direction slot in the reference's call order.
__global__ void __launch_bounds__(128) k_sgm_pathsum(const int32_t* __restrict__ g, int32_t* __restrict__ T, int32_t* __restrict__ R,
													 int32_t* __restrict__ max_abs_m, const SgmGeom geom, int slot0) {
	const int dir = c_sgm_order[slot0 + blockIdx.y];
	const int* st = c_sgm_steps[dir];
	const SgmLines L = sgm_lines(st, geom);
	const int line = blockIdx.x * blockDim.x + threadIdx.x;
	int mabs = 0;
	if (line < L.n_row + L.n_col) {
		int i, j;
		if (line < L.n_row) { i = geom.top + line; j = L.sj_row; }
		else { i = L.si_col; j = geom.left + (line - L.n_row); }
		// CORRECTED: a corner line started by both loops is walked once
		const bool dup = geom.corrected && line >= L.n_row && L.n_row > 0 && j == L.sj_row;
		if (!dup) {
			const int sv0 = st[0], sv1 = st[1], sh0 = st[2], sh1 = st[3];
			int c = 0, m_prev = 0;
			while (sgm_inside(i, j, geom)) {
				int pos[8], gv[8];
#pragma unroll
				for (int u = 0; u < 8; u++) {
					const bool ok = sgm_inside(i, j, geom);
					pos[u] = ok ? i * geom.W + j : -1;
					if (ok) { i += (c & 1) ? sv1 : sv0; j += (c & 1) ? sh1 : sh0; c++; }
				}
#pragma unroll
				for (int u = 0; u < 8; u++) gv[u] = pos[u] >= 0 ? __ldg(g + pos[u]) : 0;
#pragma unroll
				for (int u = 0; u < 8; u++) if (pos[u] >= 0) {
					atomicAdd(T + pos[u], m_prev);
					atomicAdd(R + pos[u], 1);
					m_prev = gv[u] - m_prev;
					mabs = max(mabs, abs(m_prev));
				}
			}
		}
	}
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) mabs = max(mabs, __shfl_xor_sync(0xffffffffu, mabs, o));
	if ((threadIdx.x & 31) == 0 && mabs > 0) atomicMax(max_abs_m, mabs);
}

// number of times the reference walks a line, from which of its two start loops are live (sgm.h:329-354, Q4)
__device__ __forceinline__ bool sgm_start_live(int start_pos, int corrected) { return start_pos == 1 || (start_pos == 2 && corrected); }

// Vertical and diagonal directions. The recurrence m_k = g_k - m_{k-1} is linear, m_k = sum_{t<=k} (-1)^(k-t) g_t, so a
// line does not have to be walked by one thread: a block owns 32 adjacent lines (lanes) and splits the rows into
// SWEEP_SEGS segments (warps). Phase A: every warp forms the alternating sum of its segment with a zero carry-in;
// phase B: the carry into a segment is folded from the sums of the segments before it (shared memory);
// phase C: the segment is replayed with the true carry and accumulates T += w*m_prev, R += w. The serial chain is one
// segment long (68 rows at 1080p) instead of the image height; integer arithmetic, so the split changes nothing.
constexpr int SWEEP_SEGS = 16;   // (measured at 1080p, one launch with the row walks: 32 segments 44 us, 16: 41 us, 8: 42 us)

// (bx, by: the block's place in the grid of the stand-alone launch; the fused launch below hands them in)
__device__ __forceinline__ void pathsum_sweep_body(const int32_t* __restrict__ g, int32_t* __restrict__ T, int32_t* __restrict__ R,
												   int32_t* __restrict__ max_abs_m, const SgmGeom& geom, int bx, int by) {
	__shared__ int seg_sum[SWEEP_SEGS][32];
	__shared__ int seg_cnt[SWEEP_SEGS][32];
	const int sweep_dirs[6] = {0, 1, 4, 5, 6, 7};
	const int dir = sweep_dirs[by];
	const int sv = c_sgm_steps[dir][0], sh = c_sgm_steps[dir][2];
	const bool col_live = sgm_start_live(sgm_start_pos(sv, sv), geom.corrected);
	const bool row_live = sh != 0 && sgm_start_live(sgm_start_pos(sh, sh), geom.corrected);
	const int nrows = geom.H - geom.top - geom.bottom, ncols = geom.W - geom.left - geom.right;
	if (nrows <= 0 || ncols <= 0 || (!col_live && !row_live)) return;
	const int nlines = ncols + (sh != 0 ? nrows - 1 : 0);
	const int lane = threadIdx.x & 31, seg = threadIdx.x >> 5;
	const int t = bx * 32 + lane;
	const int jlo = geom.left, jhi = geom.W - geom.right - 1;
	// k counts rows in traversal order; the line sits on column j0 + sh*k and lives on rows kf <= k < kl
	int j0 = 0, kf = 0x7fffffff, kl = 0, w = 0;
	if (t < nlines) {
		j0 = jlo + t - (sh > 0 ? nrows - 1 : 0);
		kf = sh > 0 ? max(0, jlo - j0) : (sh < 0 ? max(0, j0 - jhi) : 0);
		kl = sh > 0 ? min(nrows, jhi + 1 - j0) : (sh < 0 ? min(nrows, j0 - jlo + 1) : nrows);
		if (kf == 0) {
			w = col_live ? 1 : 0;
			const bool corner = sh != 0 && j0 == (sh > 0 ? jlo : jhi);
			if (corner && row_live) w = geom.corrected ? 1 : w + 1;
		} else w = row_live ? 1 : 0;
		if (w == 0 || kf >= kl) { kf = 0x7fffffff; kl = 0; }
	}
	const int seg_len = (nrows + SWEEP_SEGS - 1) / SWEEP_SEGS;
	const int ka = max(seg * seg_len, kf), kb = min((seg + 1) * seg_len, kl);   // this thread's rows of its line
	const int k_lo = __reduce_min_sync(0xffffffffu, ka), k_hi = __reduce_max_sync(0xffffffffu, kb);
	auto row_of = [&](int k) { return sv > 0 ? geom.top + k : geom.H - geom.bottom - 1 - k; };
	// phase A: m at the end of the segment for a zero carry-in
	int m = 0;
	for (int k0 = k_lo; k0 < k_hi; k0 += 16) {
		int gv[16];
#pragma unroll
		for (int u = 0; u < 16; u++) {
			const int k = k0 + u;
			gv[u] = (k >= ka && k < kb) ? __ldg(g + (int64_t)row_of(k) * geom.W + j0 + sh * k) : 0;
		}
#pragma unroll
		for (int u = 0; u < 16; u++) {
			const int k = k0 + u;
			if (k >= ka && k < kb) m = gv[u] - m;
		}
	}
	seg_sum[seg][lane] = m;
	seg_cnt[seg][lane] = max(kb - ka, 0);
	__syncthreads();
	// phase B: carry into this segment = m at the end of the previous one, with the carries before it folded in
	int carry = 0;
	for (int s = 0; s < seg; s++) {
		const int n = seg_cnt[s][lane];
		carry = seg_sum[s][lane] + ((n & 1) ? -carry : carry);
	}
	// phase C: replay with the true carry
	int m_prev = carry, mabs = 0;
	for (int k0 = k_lo; k0 < k_hi; k0 += 16) {
		int gv[16];
#pragma unroll
		for (int u = 0; u < 16; u++) {
			const int k = k0 + u;
			gv[u] = (k >= ka && k < kb) ? __ldg(g + (int64_t)row_of(k) * geom.W + j0 + sh * k) : 0;
		}
#pragma unroll
		for (int u = 0; u < 16; u++) {
			const int k = k0 + u;
			if (k >= ka && k < kb) {
				const int64_t pos = (int64_t)row_of(k) * geom.W + j0 + sh * k;
				atomicAdd(T + pos, w * m_prev);
				atomicAdd(R + pos, w);
				m_prev = gv[u] - m_prev;
				mabs = max(mabs, abs(m_prev));
			}
		}
	}
	mabs = __reduce_max_sync(0xffffffffu, mabs);
	if (lane == 0 && mabs > 0) atomicMax(max_abs_m, mabs);
}

__global__ void __launch_bounds__(32 * SWEEP_SEGS) k_sgm_pathsum_sweep(const int32_t* __restrict__ g, int32_t* __restrict__ T, int32_t* __restrict__ R,
																		int32_t* __restrict__ max_abs_m, const SgmGeom geom) {
	pathsum_sweep_body(g, T, R, max_abs_m, geom, blockIdx.x, blockIdx.y);
}

// row: the warp's row inside the margin rectangle; by: 0 Left2Right, 1 Right2Left
__device__ __forceinline__ void pathsum_rows_body(const int32_t* __restrict__ g, int32_t* __restrict__ T, int32_t* __restrict__ R,
												  int32_t* __restrict__ max_abs_m, const SgmGeom& geom, int row, int by) {
	const int dir = 2 + by; // Left2Right, Right2Left
	const int sh = c_sgm_steps[dir][2];
	if (!sgm_start_live(sgm_start_pos(sh, sh), geom.corrected)) return;
	const int nrows = geom.H - geom.top - geom.bottom, ncols = geom.W - geom.left - geom.right;
	const int lane = threadIdx.x & 31;
	if (row >= nrows || ncols <= 0) return;
	const int64_t base = (int64_t)(geom.top + row) * geom.W;
	int carry = 0, mabs = 0;
	for (int c0 = 0; c0 < ncols; c0 += 128) {
		int gv[4];
#pragma unroll
		for (int u = 0; u < 4; u++) {
			const int kk = c0 + u * 32 + lane;
			const int j = sh > 0 ? geom.left + kk : geom.W - geom.right - 1 - kk;
			gv[u] = kk < ncols ? __ldg(g + base + j) : 0;
		}
#pragma unroll
		for (int u = 0; u < 4; u++) {
			const int kk = c0 + u * 32 + lane;
			const int h = (kk & 1) ? -gv[u] : gv[u];
			int P = h;
#pragma unroll
			for (int o = 1; o < 32; o <<= 1) { const int n = __shfl_up_sync(0xffffffffu, P, o); if (lane >= o) P += n; }
			P += carry;
			if (kk < ncols) {
				const int Pprev = P - h;
				const int m_prev = (kk & 1) ? Pprev : -Pprev; // m_{k-1} = (-1)^(k-1) * P_{k-1}; zero for k = 0
				const int j = sh > 0 ? geom.left + kk : geom.W - geom.right - 1 - kk;
				atomicAdd(T + base + j, m_prev);
				atomicAdd(R + base + j, 1);
				mabs = max(mabs, abs(P));
			}
			carry = __shfl_sync(0xffffffffu, P, 31);
		}
	}
	mabs = __reduce_max_sync(0xffffffffu, mabs);
	if (lane == 0 && mabs > 0) atomicMax(max_abs_m, mabs);
}

__global__ void __launch_bounds__(128) k_sgm_pathsum_rows(const int32_t* __restrict__ g, int32_t* __restrict__ T, int32_t* __restrict__ R,
														  int32_t* __restrict__ max_abs_m, const SgmGeom geom) {
	pathsum_rows_body(g, T, R, max_abs_m, geom, blockIdx.x * 4 + (threadIdx.x >> 5), blockIdx.y);
}

// Both in one launch (round 2c): the row walks are latency-bound chains (issue slots 19 % busy, 22 us by themselves at 1080p) and the
// sweeps are not; side by side the walks hide behind the sweeps. The row blocks come first in the grid so that they start first.
__global__ void __launch_bounds__(32 * SWEEP_SEGS) k_sgm_pathsum_both(const int32_t* __restrict__ g, int32_t* __restrict__ T, int32_t* __restrict__ R,
																	   int32_t* __restrict__ max_abs_m, const SgmGeom geom, int n_rows_x, int n_rows_blocks,
																	   int n_sweep_x) {
	int b = blockIdx.x;
	if (b < n_rows_blocks) {   // (warp-uniform; no barrier on this side)
		pathsum_rows_body(g, T, R, max_abs_m, geom, (b % n_rows_x) * SWEEP_SEGS + (threadIdx.x >> 5), b / n_rows_x);
		return;
	}
	b -= n_rows_blocks;
	pathsum_sweep_body(g, T, R, max_abs_m, geom, b % n_sweep_x, b / n_sweep_x);
}

// S = (1+R)*C + R*pout*[j+d >= W] - T on the reference's [H][D][W] layout (column stride 1 in and out): a thread owns 4
// consecutive columns; per disparity one 8-byte load of costs and one 16-byte store of floats. With WTA the running
// arg-min (`<=`: last minimum wins, correlation_base.h:446) is kept along the way, so the aggregated volume is not read again.
// blockIdx.z splits the disparity range so that the grid is many short blocks (no tail wave).
// STORE = false: the caller only wants the winners (disparity-only pipeline): the same S is evaluated, nothing is stored.
template<bool WTA, bool STORE, int COLS>
__global__ void __launch_bounds__(256) k_sgm_apply_rows(const uint16_t* __restrict__ cv, int64_t c0, int64_t cd,
														float* __restrict__ S, int64_t s0, int64_t sd,
														const int32_t* __restrict__ T, const int32_t* __restrict__ R,
														int H, int W, int D, int dchunk, int pout,
														unsigned long long* __restrict__ keys,
														int32_t* __restrict__ dout, int64_t do0, int64_t do1, int dsign, int doffset) {
	// COLS (4 or 8) columns per thread: per disparity a warp loads 32*COLS*2 contiguous bytes of costs and stores
	// 32*COLS*4 contiguous bytes of floats (every sector it touches is written whole by one instruction)
	static_assert(COLS == 4 || COLS == 8, "columns per thread");
	using LoadT = typename std::conditional<COLS == 8, uint4, uint2>::type;
	const int j = (blockIdx.x * 256 + threadIdx.x) * COLS;
	const int i = blockIdx.y;
	if (j >= W) return; // W is a multiple of 8 on this path
	const int d_begin = blockIdx.z * dchunk, d_end = min(D, d_begin + dchunk);
	int r1[COLS], rp[COLS], t[COLS];
#pragma unroll
	for (int v = 0; v < COLS / 4; v++) {
		const int4 rv = __ldg(reinterpret_cast<const int4*>(R + (int64_t)i * W + j) + v);
		const int4 tv = __ldg(reinterpret_cast<const int4*>(T + (int64_t)i * W + j) + v);
		const int rr[4] = {rv.x, rv.y, rv.z, rv.w}, tt[4] = {tv.x, tv.y, tv.z, tv.w};
#pragma unroll
		for (int k = 0; k < 4; k++) { r1[4 * v + k] = 1 + rr[k]; rp[4 * v + k] = rr[k] * pout; t[4 * v + k] = tt[k]; }
	}
	const uint16_t* pc = cv + i * c0 + j;
	float* ps = S + i * s0 + j;
	int best[COLS], bd[COLS];
	if (WTA) {
#pragma unroll
		for (int k = 0; k < COLS; k++) { best[k] = 0x7fffffff; bd[k] = 0; }
	}
	// explicit two-phase loop: all loads of a batch are issued before the first store (the compiler does not move the
	// cache-hinted loads across the stores by itself), which keeps 64 bytes per thread in flight
	constexpr int U = 32 / COLS;
	for (int d0 = d_begin; d0 < d_end; d0 += U) {
		LoadT r[U];
#pragma unroll
		for (int u = 0; u < U; u++) {
			if (d0 + u < d_end) r[u] = __ldcs(reinterpret_cast<const LoadT*>(pc + (d0 + u) * cd));
			else r[u] = LoadT();
		}
#pragma unroll
		for (int u = 0; u < U; u++) {
			const int d = d0 + u;
			if (d < d_end) {
				uint32_t w[COLS / 2];
				memcpy(w, &r[u], sizeof(LoadT));
				float o[COLS];
#pragma unroll
				for (int k = 0; k < COLS; k++) {
					const int cst = (w[k >> 1] >> ((k & 1) * 16)) & 0xffff;
					const int s = r1[k] * cst + (j + k + d >= W ? rp[k] : 0) - t[k];
					o[k] = (float)s;
					if (WTA) { if (s <= best[k]) { best[k] = s; bd[k] = d; } }
				}
				if (STORE) {
#pragma unroll
					for (int v = 0; v < COLS / 4; v++)
						__stcs(reinterpret_cast<float4*>(ps + d * sd) + v, make_float4(o[4 * v], o[4 * v + 1], o[4 * v + 2], o[4 * v + 3]));
				}
			}
		}
	}
	if (WTA) {
		if (!keys) {
			// the block saw every disparity: its winners are final (selectedIndexToDisp folded in)
#pragma unroll
			for (int k = 0; k < COLS; k++) dout[i * do0 + (j + k) * do1] = dsign * bd[k] + doffset;
			return;
		}
		// the disparity range is split over blockIdx.z: merge through one 64-bit atomicMin per pixel and chunk on the key
		// (S biased to unsigned) << 32 | (2^32-1 - d): smallest S wins, equal S -> largest d, i.e. the reference's `<=` rule
#pragma unroll
		for (int k = 0; k < COLS; k++) {
			const unsigned long long key = ((unsigned long long)((uint32_t)best[k] ^ 0x80000000u) << 32) | (0xffffffffu - (uint32_t)bd[k]);
			atomicMin(keys + (int64_t)i * W + j + k, key);
		}
	}
}

__global__ void __launch_bounds__(256) k_wta_decode_keys(const unsigned long long* __restrict__ keys, int H, int W, int sign, int offset,
														 int32_t* __restrict__ out, int64_t o0, int64_t o1) {
	const int j = blockIdx.x * 256 + threadIdx.x;
	const int i = blockIdx.y;
	if (j >= W) return;
	const int d = (int)(0xffffffffu - (uint32_t)keys[(int64_t)i * W + j]);
	out[i * o0 + j * o1] = sign * d + offset;
}

// Disparity-only census + SGM in ONE pass over the costs (round 2c). With the collapsed evaluation S(d) = (1+R)*C(d) + R*pout*[j+d >= W] - T,
// 1 + R >= 1, so on each side of the column W - j the order of S is the order of C: the winner under the reference's `<=` scan (last
// minimum wins, correlation_base.h:446) is the largest-d minimum of C on one of the two sides, and a tie between the sides goes
// to the far side (its disparities are the larger ones). The Hamming pass leaves, per pixel, kin / kout = min over the d of a side of
// C*8192 + (8191 - d) (0xffffffff: the side is empty); g = min_d (2*C + pout*[j+d >= W]) follows from the same two numbers.
// (nz: the disparity chunks of the Hamming pass leave one slice of n keys each; their minimum is taken here: no atomics, no pre-fill)
__global__ void __launch_bounds__(256) k_keys_to_gmap(const uint32_t* __restrict__ kin, const uint32_t* __restrict__ kout, int n, int nz, int pout,
													  int32_t* __restrict__ g) {
	const int k = blockIdx.x * 256 + threadIdx.x;
	if (k >= n) return;
	uint32_t a = kin[k], b = kout[k];
	for (int z = 1; z < nz; z++) { a = min(a, kin[(int64_t)z * n + k]); b = min(b, kout[(int64_t)z * n + k]); }
	int v = 0x7fffffff;
	if (a != 0xffffffffu) v = 2 * (int)(a >> 13);
	if (b != 0xffffffffu) v = min(v, 2 * (int)(b >> 13) + pout);
	g[k] = v;
}

__global__ void __launch_bounds__(256) k_keys_to_disp(const uint32_t* __restrict__ kin, const uint32_t* __restrict__ kout,
													  const int32_t* __restrict__ T, const int32_t* __restrict__ R, int H, int W, int nz, int pout,
													  int sign, int offset, int32_t* __restrict__ out, int64_t o0, int64_t o1) {
	const int j = blockIdx.x * 256 + threadIdx.x;
	const int i = blockIdx.y;
	if (j >= W) return;
	const int64_t k = (int64_t)i * W + j, n = (int64_t)H * W;
	uint32_t a = kin[k], b = kout[k];
	for (int z = 1; z < nz; z++) { a = min(a, kin[z * n + k]); b = min(b, kout[z * n + k]); }
	const int r = R[k], t = T[k];
	int d = 0;   // (no disparity at all cannot happen: d = 0 always lies on the near side)
	if (a != 0xffffffffu) d = 8191 - (int)(a & 8191u);
	if (b != 0xffffffffu) {
		const int s_out = (1 + r) * (int)(b >> 13) + r * pout - t;
		if (a == 0xffffffffu || s_out <= (1 + r) * (int)(a >> 13) - t) d = 8191 - (int)(b & 8191u);
	}
	out[i * o0 + j * o1] = sign * d + offset;
}

// any strides: one thread per pixel
__global__ void __launch_bounds__(128) k_sgm_apply_generic(const uint16_t* __restrict__ cv, int64_t c0, int64_t c1, int64_t cd,
														   float* __restrict__ S, int64_t s0, int64_t s1, int64_t sd,
														   const int32_t* __restrict__ T, const int32_t* __restrict__ R,
														   int H, int W, int D, int pout, int32_t* __restrict__ idx, int64_t i0, int64_t i1,
														   int sign, int offset) {
	const int j = blockIdx.x * 128 + threadIdx.x;
	const int i = blockIdx.y;
	if (j >= W) return;
	const int r = R[(int64_t)i * W + j], t = T[(int64_t)i * W + j];
	const uint16_t* pc = cv + i * c0 + j * c1;
	float* ps = S + i * s0 + j * s1;
	int best = 0x7fffffff, bd = 0;
	for (int d = 0; d < D; d++) {
		const int s = (1 + r) * (int)pc[d * cd] + (j + d >= W ? r * pout : 0) - t;
		ps[d * sd] = (float)s;
		if (s <= best) { best = s; bd = d; }
	}
	if (idx) idx[i * i0 + j * i1] = sign * bd + offset;
}

} // namespace stevi

namespace stevi {

static const int32_t STEVI_FALLBACK = -1000; // internal: the collapsed evaluation is not provably exact here

static bool aligned16(const void* p, int64_t a, int64_t b, size_t esz) {
	return ((uintptr_t)p % 16 == 0) && ((a * (int64_t)esz) % 16 == 0) && ((b * (int64_t)esz) % 16 == 0);
}

// the aggregated volume a caller did not ask for (out.p == nullptr) only exists on the paths that cannot do without it
static int32_t ensure_volume(stevi_ctx* c, View& out) {
	if (out.p) return STEVI_OK;
	float* p = nullptr;
	STEVI_TRY(arena_alloc_t(c, (size_t)out.shape[0] * out.shape[1] * out.shape[2], &p));
	out.p = p;
	return STEVI_OK;
}

// Path sums of the collapsed evaluation from the g map: T(p) = sum over passes of m(prev), R(p) = number of passes.
// STEVI_FALLBACK when the float arithmetic this replaces is not provably exact (magnitudes >= 2^24).
static int32_t collapsed_sums(stevi_ctx* c, const int32_t* g, int H, int W, int ndir, int pout, const stevi_margins& mg,
							  int max_cost_hint, int32_t** T_out, int32_t** R_out) {
	const size_t npix = (size_t)H * W;
	int32_t* scratch = nullptr; // [T | R | max_abs_m]
	STEVI_TRY(arena_alloc_t(c, 2 * npix + 64, &scratch));
	int32_t* T = scratch;
	int32_t* R = scratch + npix;
	int32_t* d_max = scratch + 2 * npix;
	STEVI_CUDA(c, cudaMemsetAsync(scratch, 0, (2 * npix + 64) * sizeof(int32_t), c->stream));
	SgmGeom geom{H, W, mg.top, mg.bottom, mg.left, mg.right, 0};
	{
		const int nrows = H - mg.top - mg.bottom;
		dim3 gsweep(cdiv(H + W, 32), ndir == 4 ? 2 : 6);
		static const bool separate = getenv("STEVI_PATHSUM_SEPARATE") != nullptr;   // measurement aid: the two launches of round 2
		if (nrows > 0 && !separate) {
			const int n_rows_x = cdiv(nrows, SWEEP_SEGS), n_rows_blocks = 2 * n_rows_x;
			STEVI_LAUNCH(c, k_sgm_pathsum_both, dim3(n_rows_blocks + gsweep.x * gsweep.y), 32 * SWEEP_SEGS, 0, g, T, R, d_max, geom, n_rows_x,
						 n_rows_blocks, (int)gsweep.x);
		} else {
			STEVI_LAUNCH(c, k_sgm_pathsum_sweep, gsweep, 32 * SWEEP_SEGS, 0, g, T, R, d_max, geom);
			if (nrows > 0) {
				dim3 grows(cdiv(nrows, 4), 2);
				STEVI_LAUNCH(c, k_sgm_pathsum_rows, grows, 128, 0, g, T, R, d_max, geom);
			}
		}
		if (ndir == 16) {
			dim3 gwalk(cdiv(H + W, 128), 8);
			STEVI_LAUNCH(c, k_sgm_pathsum, gwalk, 128, 0, g, T, R, d_max, geom, 8);
		}
	}
	// exactness bound of the float arithmetic this replaces (all magnitudes < 2^24)
	const long max_c = max_cost_hint >= 0 ? max_cost_hint : 65535;
	const long rmax = ndir + 1, apout = labs((long)pout);
	const long lim = 1L << 24;
	const long m_apriori = (2 * max_c + apout) * (long)(H + W);
	if ((1 + rmax) * max_c + rmax * apout + rmax * m_apriori >= lim) {
		int32_t h_max = 0;
		STEVI_CUDA(c, cudaMemcpyAsync(&h_max, d_max, sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
		STEVI_CUDA(c, cudaStreamSynchronize(c->stream));
		if ((1 + rmax) * max_c + rmax * apout + rmax * (long)h_max >= lim) return STEVI_FALLBACK;
	}
	*T_out = T;
	*R_out = R;
	return STEVI_OK;
}

static int32_t sgm_collapsed(stevi_ctx* c, const View& cv, int ndir, int pout, const stevi_margins& mg, View& out,
							 const int32_t* gmap_ready, int max_cost_hint, const WtaSink* wta) {
	const int H = cv.shape[0], W = cv.shape[1], D = cv.shape[2];
	const size_t npix = (size_t)H * W;
	const int32_t* g = gmap_ready;
	const uint16_t* pc = cv.as<uint16_t>();
	if (!g) {
		int32_t* gbuf = nullptr;
		STEVI_TRY(arena_alloc_t(c, npix, &gbuf));
		bool done = false;
		if (cv.st[1] == 1 || cv.st[0] == 1) {
			const int b = cv.st[1] == 1 ? 1 : 0, a = 1 - b;
			if (aligned16(pc, cv.st[a], cv.st[2], 2) && cv.shape[a] <= 65535) {
				dim3 grid(cdiv(cv.shape[b], 1024), cv.shape[a]);
				STEVI_LAUNCH(c, k_sgm_gmap_planes, grid, 128, 0, pc, cv.st[a], cv.st[2], cv.shape[a], cv.shape[b], D, W, b, pout, gbuf);
				done = true;
			}
		}
		if (!done)
			STEVI_LAUNCH(c, k_sgm_gmap_generic, (unsigned)((npix + 7) / 8), 256, 0, pc, cv.st[0], cv.st[1], cv.st[2], H, W, D, pout, gbuf);
		g = gbuf;
	}
	int32_t *T = nullptr, *R = nullptr;
	STEVI_TRY(collapsed_sums(c, g, H, W, ndir, pout, mg, max_cost_hint, &T, &R));
	const bool store = out.p != nullptr || !wta;
	const bool rows = cv.st[1] == 1 && (W % 8) == 0 && aligned16(pc, cv.st[0], cv.st[2], 2) && H <= 65535 &&
					  (!store || (out.st[1] == 1 && aligned16(out.p, out.st[0], out.st[2], 4)));
	if (store || !rows) STEVI_TRY(ensure_volume(c, out));
	float* ps = out.as<float>();
	if (rows) {
		constexpr int cols = 4;   // 8 columns per thread measured 0.82 ms against 0.60 ms at C2 (two 16-byte stores, round 1) and
		                          // 0.586 against 0.557 ms with one 256-bit store per thread and disparity (round 2c)
		const int xb = cdiv(W, 256 * cols);
		// split the disparity range only when a row-per-block grid would leave SMs idle (< 2 waves of 6 resident blocks per
		// SM): every split merges its winners through atomics and re-reads T and R (measured at C2: unsplit 0.567 ms, 2 chunks
		// 0.570, 7 chunks 0.603, 14 chunks 0.655)
		const long want = 2L * 6 * c->sm_count;
		long parts = std::min<long>(std::max(1, D / 32), (want + (long)xb * H - 1) / ((long)xb * H));
		const int dchunk = (int)((D + parts - 1) / parts);
		dim3 grid(xb, H, cdiv(D, dchunk));
		unsigned long long* keys = nullptr;
		const bool split = grid.z > 1;
		if (wta && split) {
			STEVI_TRY(arena_alloc_t(c, npix, &keys));
			STEVI_CUDA(c, cudaMemsetAsync(keys, 0xff, npix * sizeof(unsigned long long), c->stream));
		}
		int32_t* dout = wta ? wta->out.as<int32_t>() : nullptr;
		const int64_t do0 = wta ? wta->out.st[0] : 0, do1 = wta ? wta->out.st[1] : 0;
		const int dsign = wta ? wta->sign : 1, doffset = wta ? wta->offset : 0;
#define STEVI_APPLY(W_, S_, C_) STEVI_LAUNCH(c, (k_sgm_apply_rows<W_, S_, C_>), grid, 256, 0, pc, cv.st[0], cv.st[2], ps, out.st[0], out.st[2], T, R, H, W, D, dchunk, pout, keys, dout, do0, do1, dsign, doffset)
		if (wta && store) STEVI_APPLY(true, true, cols); else if (wta) STEVI_APPLY(true, false, cols); else STEVI_APPLY(false, true, cols);
#undef STEVI_APPLY
		if (wta && split) {
			dim3 gdec(cdiv(W, 256), H);
			STEVI_LAUNCH(c, k_wta_decode_keys, gdec, 256, 0, keys, H, W, wta->sign, wta->offset, wta->out.as<int32_t>(), wta->out.st[0], wta->out.st[1]);
		}
	} else {
		dim3 grid(cdiv(W, 128), H);
		STEVI_LAUNCH(c, k_sgm_apply_generic, grid, 128, 0, pc, cv.st[0], cv.st[1], cv.st[2], ps, out.st[0], out.st[1], out.st[2], T, R, H, W, D, pout,
					 wta ? wta->out.as<int32_t>() : (int32_t*)nullptr, wta ? wta->out.st[0] : 0, wta ? wta->out.st[1] : 0,
					 wta ? wta->sign : 1, wta ? wta->offset : 0);
	}
	return STEVI_OK;
}

int32_t sgm_census_disparity_dev(stevi_ctx* c, const HamSource& h, int ndir, int pout, const stevi_margins& mg, int max_cost,
								 const WtaSink& wta) {
	if (ndir != 4 && ndir != 8 && ndir != 16) return fail(c, STEVI_E_ARG, "sgm: n_directions must be 4, 8 or 16");
	if (mg.left < 0 || mg.right < 0 || mg.top < 0 || mg.bottom < 0) return fail(c, STEVI_E_ARG, "sgm: negative margins");
	if (!hamming_fused_capable(h)) return STEVI_NOT_FUSED;
	const bool r2l = h.direction == STEVI_RIGHT_TO_LEFT;
	const int H = h.fl.shape[0], W = r2l ? h.fr.shape[1] : h.fl.shape[1];
	const size_t npix = (size_t)H * W;
	static const bool two_pass = getenv("STEVI_DISP_TWO_PASS") != nullptr;   // measurement aid: the round-2 form (costs evaluated twice)
	int32_t* g = nullptr;
	STEVI_TRY(arena_alloc_t(c, npix, &g));
	if (!two_pass && npix < (size_t)1 << 31) {
		// one pass over the costs (in registers only): the two side minima per pixel; g and, after the path sums, the winner follow
		uint32_t *kin = nullptr, *kout = nullptr;
		int nz = 1;
		STEVI_TRY(hamming_minkeys_dev(c, h, &kin, &kout, &nz));
		STEVI_LAUNCH(c, k_keys_to_gmap, dim3(cdiv((int)npix, 256)), 256, 0, kin, kout, (int)npix, nz, pout, g);
		int32_t *T = nullptr, *R = nullptr;
		const int32_t s = collapsed_sums(c, g, H, W, ndir, pout, mg, max_cost, &T, &R);
		if (s == STEVI_FALLBACK) return STEVI_NOT_FUSED;
		STEVI_TRY(s);
		dim3 gd(cdiv(W, 256), H);
		STEVI_LAUNCH(c, k_keys_to_disp, gd, 256, 0, kin, kout, T, R, H, W, nz, pout, wta.sign, wta.offset, wta.out.as<int32_t>(), wta.out.st[0], wta.out.st[1]);
		return STEVI_OK;
	}
	// pass 1: g(p) = min_d (2*C + pout*[j+d >= W]); the costs live in registers only
	STEVI_TRY(hamming_gmap_dev(c, h, g, pout));
	int32_t *T = nullptr, *R = nullptr;
	const int32_t s = collapsed_sums(c, g, H, W, ndir, pout, mg, max_cost, &T, &R);
	if (s == STEVI_FALLBACK) return STEVI_NOT_FUSED;
	STEVI_TRY(s);
	// pass 2: the costs again, S = (1+R)*C + R*pout*[j+d >= W] - T and its arg-min
	unsigned long long* keys = nullptr;
	STEVI_TRY(arena_alloc_t(c, npix, &keys));
	STEVI_CUDA(c, cudaMemsetAsync(keys, 0xff, npix * sizeof(unsigned long long), c->stream));
	STEVI_TRY(hamming_sgm_wta_dev(c, h, T, R, pout, 2 * ndir, keys));   // Q4: a line can be walked twice
	dim3 gdec(cdiv(W, 256), H);
	STEVI_LAUNCH(c, k_wta_decode_keys, gdec, 256, 0, keys, H, W, wta.sign, wta.offset, wta.out.as<int32_t>(), wta.out.st[0], wta.out.st[1]);
	return STEVI_OK;
}

int32_t sgm_general(stevi_ctx* c, const View& cv, int ndir, int strategy, float P1, float P2, float Pout,
					const stevi_margins& mg, int semantics, const View& out);

int32_t sgm_dev(stevi_ctx* c, const View& cv, int ndir, int strategy, float P1, float P2, float Pout,
				const stevi_margins& mg, int semantics, const View& out_, const int32_t* gmap_ready, int max_cost_hint,
				const WtaSink* wta) {
	View out = out_;   // out.p == nullptr (pipeline callers that only want the winners): no aggregated volume unless a path needs one
	if (!out.p && !wta) return fail(c, STEVI_E_ARG, "sgm: no output requested");
	if (ndir != 4 && ndir != 8 && ndir != 16) return fail(c, STEVI_E_ARG, "sgm: n_directions must be 4, 8 or 16");
	if (cv.ndim != 3 || (cv.dtype != STEVI_U16 && cv.dtype != STEVI_F32))
		return fail(c, STEVI_E_DTYPE, "sgm: cost volume must be (H,W,D) u16 or f32");
	if (out.ndim != 3 || out.dtype != STEVI_F32 || out.shape[0] != cv.shape[0] || out.shape[1] != cv.shape[1] || out.shape[2] != cv.shape[2])
		return fail(c, STEVI_E_SHAPE, "sgm: out must be an f32 volume of the same shape");
	if (strategy != STEVI_COST && strategy != STEVI_SCORE) return fail(c, STEVI_E_ARG, "sgm: bad strategy");
	if (mg.left < 0 || mg.right < 0 || mg.top < 0 || mg.bottom < 0) return fail(c, STEVI_E_ARG, "sgm: negative margins");
	if (cv.shape[0] == 0 || cv.shape[1] == 0 || cv.shape[2] == 0) return STEVI_OK;
	if (cv.shape[0] > 65535) return fail(c, STEVI_E_UNSUPPORTED, "sgm: more than 65535 rows");
	const bool pout_int = Pout == rintf(Pout) && fabsf(Pout) <= 1.0e6f;
	if (semantics == STEVI_COMPAT && strategy == STEVI_COST && cv.dtype == STEVI_U16 && pout_int) {
		int32_t s = sgm_collapsed(c, cv, ndir, (int)Pout, mg, out, gmap_ready, max_cost_hint, wta);
		if (s != STEVI_FALLBACK) return s;
	}
	STEVI_TRY(ensure_volume(c, out));
	STEVI_TRY(sgm_general(c, cv, ndir, strategy, P1, P2, Pout, mg, semantics, out));
	if (wta) {
		STEVI_TRY(wta_dev(c, out, strategy, wta->out));
		if (wta->sign != 1 || wta->offset != 0)
			STEVI_TRY(index_to_disp_dev(c, wta->out, wta->sign > 0 ? STEVI_RIGHT_TO_LEFT : STEVI_LEFT_TO_RIGHT, wta->offset, wta->out));
	}
	return STEVI_OK;
}

} // namespace stevi
