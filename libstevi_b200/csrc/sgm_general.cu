// General SGM recurrence (correlation/sgm.h:186-404): Score strategy (a genuine P1/P2 recurrence, :218-256), float cost
// volumes in Cost mode (the `c_score` quirk Q3 evaluated per path in the reference's operation order, :257-296) and
// CORRECTED semantics (textbook recurrence, every direction live).
//
// Layout: the volume is brought to a dense f32 [H][W][D] workspace (d fastest) so that a warp reads the D costs of a
// pixel as contiguous 128-byte segments; the aggregated volume is accumulated in a second workspace of that layout and
// written to the caller's strides at the end.
//
// One warp owns one line (sgm.h:186 traverseLine); lane l holds the disparities d = 32*k + l, k < K. Per pixel:
//   max_p  = max_d prev(d)                                  (:220-227)   one warp reduction
//   far(d) = max_{|od-d|>1} prev(od)                        (:236-238)   the global arg-max serves every d outside its
//                                                                        own 3-window; the (at most) three d inside it get
//                                                                        their own masked reductions
//   near   = prev(d-1), prev(d+1) through lane shuffles     (:235)
//   act(d) = c + (max(far-P2, near-P1, prev(d)) [-Pout] - max_p)         (:240-254)
//   S     += act - c                                        (:298-300)
// Subtracting the constant before or after the max is the same in IEEE arithmetic (monotone rounding), non-finite previous
// scores are skipped like the reference's isfinite tests; the float operation order per path and the order of the
// directions and of the two start loops follow the reference, so results are bit-identical to it for the 4- and
// 8-direction sets. (16 directions: lines of one direction overlap, Q8; their contributions are added with float atomics,
// i.e. in a different order than the reference's sequential loop — equal up to float rounding, exact for integer data.)
#include "common.cuh"
#include "kernels.cuh"
#include <cmath>

namespace stevi {

__constant__ int c_gsteps[16][4] = {
	{1, 1, 0, 0}, {-1, -1, 0, 0}, {0, 0, 1, 1}, {0, 0, -1, -1},
	{1, 1, 1, 1}, {-1, -1, -1, -1}, {1, 1, -1, -1}, {-1, -1, 1, 1},
	{0, 1, 1, 1}, {0, -1, -1, -1}, {0, 1, -1, -1}, {0, -1, 1, 1},
	{1, 1, 0, 1}, {-1, -1, 0, -1}, {1, 1, 0, -1}, {-1, -1, 0, 1}};
static const int h_gsteps[16][4] = {
	{1, 1, 0, 0}, {-1, -1, 0, 0}, {0, 0, 1, 1}, {0, 0, -1, -1},
	{1, 1, 1, 1}, {-1, -1, -1, -1}, {1, 1, -1, -1}, {-1, -1, 1, 1},
	{0, 1, 1, 1}, {0, -1, -1, -1}, {0, 1, -1, -1}, {0, -1, 1, 1},
	{1, 1, 0, 1}, {-1, -1, 0, -1}, {1, 1, 0, -1}, {-1, -1, 0, 1}};
static const int h_gorder[16] = {0, 1, 2, 3, 4, 5, 6, 7, 12, 13, 14, 15, 8, 9, 10, 11}; // sgm.h:379-401

// ---- layout conversion: (H,W,D) with any strides <-> dense f32 [H][W][D] -----------------------------------------
// 32x32 tile over (j, d) per row through shared memory: reads coalesced along whichever of j / d the strided side has
// stride 1 on, writes coalesced on the other side.
template<typename T, bool TO_DENSE>
__global__ void __launch_bounds__(256) k_volume_relayout(T* __restrict__ strided, int64_t s0, int64_t s1, int64_t s2,
														 float* __restrict__ dense, int H, int W, int D, int j_fast) {
	__shared__ float tile[32][33];
	const int i = blockIdx.z;
	const int j0 = blockIdx.x * 32, d0 = blockIdx.y * 32;
	const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5; // 32 x 8
	T* srow = strided + i * s0;
	float* drow = dense + (int64_t)i * W * D;
	if (TO_DENSE) {
		for (int r = ty; r < 32; r += 8) {
			// strided side: tx runs along its fast axis
			const int j = j_fast ? j0 + tx : j0 + r, d = j_fast ? d0 + r : d0 + tx;
			if (j < W && d < D) tile[j - j0][d - d0] = (float)srow[j * s1 + d * s2];
		}
		__syncthreads();
		for (int r = ty; r < 32; r += 8) {
			const int j = j0 + r, d = d0 + tx;
			if (j < W && d < D) drow[(int64_t)j * D + d] = tile[r][tx];
		}
	} else {
		for (int r = ty; r < 32; r += 8) {
			const int j = j0 + r, d = d0 + tx;
			if (j < W && d < D) tile[r][tx] = drow[(int64_t)j * D + d];
		}
		__syncthreads();
		for (int r = ty; r < 32; r += 8) {
			const int j = j_fast ? j0 + tx : j0 + r, d = j_fast ? d0 + r : d0 + tx;
			if (j < W && d < D) srow[j * s1 + d * s2] = (T)tile[j - j0][d - d0];
		}
	}
}

template<typename T, bool TO_DENSE>
static int32_t relayout(stevi_ctx* c, const View& v, float* dense) {
	const int H = v.shape[0], W = v.shape[1], D = v.shape[2];
	dim3 grid(cdiv(W, 32), cdiv(D, 32), H);
	const int j_fast = v.st[1] <= v.st[2] ? 1 : 0;
	STEVI_LAUNCH(c, (k_volume_relayout<T, TO_DENSE>), grid, 256, 0, v.as<T>(), v.st[0], v.st[1], v.st[2], dense, H, W, D, j_fast);
	return STEVI_OK;
}

} // namespace stevi

namespace stevi {

enum { GM_SCORE = 0, GM_COST_COMPAT = 1, GM_COST = 2 };

struct GLineParams {
	const float* C; float* S;
	int H, W, D, top, bottom, left, right;
	int dir, kind, n_lines, start_fixed, skip_col;
	float P1, P2, Pout;
};

template<int MODE> __device__ __forceinline__ float g_better(float a, float b) { return MODE == GM_SCORE ? fmaxf(a, b) : fminf(a, b); }
// one instruction on sm_100a (redux.sync on f32 = CREDUX) instead of five dependent shuffle rounds; the operands are never NaN
// (non-finite scores are replaced by the worst value before they get here)
template<int MODE> __device__ __forceinline__ float g_warp_best(float v) {
	float r;
	if (MODE == GM_SCORE) asm volatile("redux.sync.max.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
	else asm volatile("redux.sync.min.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
	return r;
}

template<int K, int MODE>
__global__ void __launch_bounds__(128) k_sgm_lines(const GLineParams p) {
	const int lane = threadIdx.x & 31;
	const int line = blockIdx.x * 4 + (threadIdx.x >> 5);
	if (line >= p.n_lines) return;
	int i = p.kind == 0 ? p.top + line : p.start_fixed;
	int j = p.kind == 0 ? p.start_fixed : p.left + line;
	if (p.kind == 1 && j == p.skip_col) return; // CORRECTED: the corner line was already started by the other loop
	const int sv0 = c_gsteps[p.dir][0], sv1 = c_gsteps[p.dir][1], sh0 = c_gsteps[p.dir][2], sh1 = c_gsteps[p.dir][3];
	const float WORST = MODE == GM_SCORE ? -INFINITY : INFINITY;
	const int D = p.D, W = p.W;
	auto inside = [&](int ii, int jj) { return ii >= p.top && ii < p.H - p.bottom && jj >= p.left && jj < W - p.right; };
	float prev[K], cn[K];
#pragma unroll
	for (int k = 0; k < K; k++) prev[k] = (32 * k + lane < D) ? 0.f : WORST; // previous_cost starts at 0 (sgm.h:206-208)
	if (!inside(i, j)) return;
	{
		const float* pc = p.C + ((int64_t)i * W + j) * D;
#pragma unroll
		for (int k = 0; k < K; k++) cn[k] = (32 * k + lane < D) ? __ldg(pc + 32 * k + lane) : 0.f;
	}
	int step = 0;
	while (true) {
		float c[K];
#pragma unroll
		for (int k = 0; k < K; k++) c[k] = cn[k];
		// next pixel of the line: issue its loads before working on this one
		const int i2 = i + ((step & 1) ? sv1 : sv0), j2 = j + ((step & 1) ? sh1 : sh0);
		const bool more = inside(i2, j2);
		if (more) {
			const float* pc = p.C + ((int64_t)i2 * W + j2) * D;
#pragma unroll
			for (int k = 0; k < K; k++) cn[k] = (32 * k + lane < D) ? __ldg(pc + 32 * k + lane) : 0.f;
		}
		// finite previous scores only (sgm.h:224,243,264,281)
		float pf[K];
		float loc = WORST;
#pragma unroll
		for (int k = 0; k < K; k++) { pf[k] = isfinite(prev[k]) ? prev[k] : WORST; loc = g_better<MODE>(loc, pf[k]); }
		const float best_p = g_warp_best<MODE>(loc); // max_p_cost / min_p_cost
		const bool any = isfinite(best_p);
		float act[K];
		if (MODE == GM_COST_COMPAT) {
			// Q3: min_a_cost ends up being c_score as soon as one previous score is finite
#pragma unroll
			for (int k = 0; k < K; k++) {
				const int d = 32 * k + lane;
				float m = any ? c[k] : INFINITY;
				if (j + d >= W) m += p.Pout;
				act[k] = c[k];
				if (isfinite(m) && any) act[k] += m - best_p;
			}
		} else {
			// position of the extremum (smallest index among equals) and the three masked extrema around it
			int loc_i = 0x7fffffff;
#pragma unroll
			for (int k = 0; k < K; k++) if (pf[k] == best_p && 32 * k + lane < D) loc_i = min(loc_i, 32 * k + lane);
			const int i1 = __reduce_min_sync(0xffffffffu, loc_i);
			float e[3] = {WORST, WORST, WORST}; // best over |od - s| > 1 for s = i1-1, i1, i1+1
#pragma unroll
			for (int k = 0; k < K; k++) {
				const int d = 32 * k + lane;
#pragma unroll
				for (int t = 0; t < 3; t++) if (abs(d - (i1 - 1 + t)) > 1) e[t] = g_better<MODE>(e[t], pf[k]);
			}
#pragma unroll
			for (int t = 0; t < 3; t++) e[t] = g_warp_best<MODE>(e[t]);
#pragma unroll
			for (int k = 0; k < K; k++) {
				const int d = 32 * k + lane;
				float lft = __shfl_up_sync(0xffffffffu, pf[k], 1);
				const float wrap_l = __shfl_sync(0xffffffffu, k > 0 ? pf[k > 0 ? k - 1 : 0] : WORST, 31);
				if (lane == 0) lft = wrap_l;
				float rgt = __shfl_down_sync(0xffffffffu, pf[k], 1);
				const float wrap_r = __shfl_sync(0xffffffffu, k + 1 < K ? pf[k + 1 < K ? k + 1 : K - 1] : WORST, 0);
				if (lane == 31) rgt = wrap_r;
				const int rel = d - i1;
				const float far = (rel < -1 || rel > 1) ? best_p : e[rel + 1];
				float m;
				if (MODE == GM_SCORE) m = fmaxf(fmaxf(far - p.P2, fmaxf(lft, rgt) - p.P1), pf[k]);
				else m = fminf(fminf(far + p.P2, fminf(lft, rgt) + p.P1), pf[k]);
				if (j + d >= W) m = MODE == GM_SCORE ? m - p.Pout : m + p.Pout;
				act[k] = c[k];
				if (isfinite(m) && any) act[k] += m - best_p;
			}
		}
		float* ps = p.S + ((int64_t)i * W + j) * D;
#pragma unroll
		for (int k = 0; k < K; k++) {
			const int d = 32 * k + lane;
			if (d < D) { atomicAdd(ps + d, act[k] - c[k]); prev[k] = act[k]; } // one add per cell and launch for 4/8 directions
		}
		if (!more) break;
		i = i2; j = j2; step++;
	}
}

// Any number of disparities (the reference has no limit): a block per line, the previous scores of the line in shared
// memory, the same steps as k_sgm_lines with block-wide reductions. Slow (a 128-thread block walks one line) but only reached
// by D > 512.
template<int MODE>
__device__ __forceinline__ float g_block_best(float v, float* red) {
	v = g_warp_best<MODE>(v);
	__syncthreads();
	if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
	__syncthreads();
	float r = red[0];
	for (int k = 1; k < (int)(blockDim.x >> 5); k++) r = g_better<MODE>(r, red[k]);
	return r;
}

template<int MODE>
__global__ void __launch_bounds__(1024) k_sgm_lines_any(const GLineParams p) {
	extern __shared__ float sm[];
	float* prev = sm;            // [D] finite previous scores (others replaced by the worst value)
	float* red = sm + p.D;       // [4]
	__shared__ int s_i1;
	const int line = blockIdx.x;
	if (line >= p.n_lines) return;
	int i = p.kind == 0 ? p.top + line : p.start_fixed;
	int j = p.kind == 0 ? p.start_fixed : p.left + line;
	if (p.kind == 1 && j == p.skip_col) return;
	const int sv0 = c_gsteps[p.dir][0], sv1 = c_gsteps[p.dir][1], sh0 = c_gsteps[p.dir][2], sh1 = c_gsteps[p.dir][3];
	const float WORST = MODE == GM_SCORE ? -INFINITY : INFINITY;
	const int D = p.D, W = p.W;
	auto inside = [&](int ii, int jj) { return ii >= p.top && ii < p.H - p.bottom && jj >= p.left && jj < W - p.right; };
	if (!inside(i, j)) return;
	for (int d = threadIdx.x; d < D; d += blockDim.x) prev[d] = 0.f;
	int step = 0;
	while (true) {
		__syncthreads();
		float loc = WORST;
		for (int d = threadIdx.x; d < D; d += blockDim.x) loc = g_better<MODE>(loc, prev[d]);
		const float best_p = g_block_best<MODE>(loc, red);
		const bool any = isfinite(best_p);
		float e[3] = {WORST, WORST, WORST};
		int i1 = 0;
		if (MODE != GM_COST_COMPAT) {
			if (threadIdx.x == 0) s_i1 = 0x7fffffff;
			__syncthreads();
			int li = 0x7fffffff;
			for (int d = threadIdx.x; d < D; d += blockDim.x) if (prev[d] == best_p) { li = d; break; }
			if (li != 0x7fffffff) atomicMin(&s_i1, li);
			__syncthreads();
			i1 = s_i1;
			for (int t = 0; t < 3; t++) {
				float l = WORST;
				for (int d = threadIdx.x; d < D; d += blockDim.x) if (abs(d - (i1 - 1 + t)) > 1) l = g_better<MODE>(l, prev[d]);
				e[t] = g_block_best<MODE>(l, red);
			}
		}
		const float* pc = p.C + ((int64_t)i * W + j) * D;
		float* ps = p.S + ((int64_t)i * W + j) * D;
		// every thread computes its cells from the old scores, then (after a barrier) replaces them
		float newv[8];   // D <= 8 * blockDim.x per pass; larger D loop in passes
		for (int base = 0; base < D; base += 8 * blockDim.x) {
			int n = 0;
			for (int d = base + threadIdx.x; d < D && n < 8; d += blockDim.x, n++) {
				const float c = __ldg(pc + d);
				float act = c;
				if (MODE == GM_COST_COMPAT) {
					float m = any ? c : INFINITY;
					if (j + d >= W) m += p.Pout;
					if (isfinite(m) && any) act += m - best_p;
				} else {
					const float lft = d > 0 ? prev[d - 1] : WORST, rgt = d + 1 < D ? prev[d + 1] : WORST;
					const int rel = d - i1;
					const float far = (rel < -1 || rel > 1) ? best_p : e[rel + 1];
					float m;
					if (MODE == GM_SCORE) m = fmaxf(fmaxf(far - p.P2, fmaxf(lft, rgt) - p.P1), prev[d]);
					else m = fminf(fminf(far + p.P2, fminf(lft, rgt) + p.P1), prev[d]);
					if (j + d >= W) m = MODE == GM_SCORE ? m - p.Pout : m + p.Pout;
					if (isfinite(m) && any) act += m - best_p;
				}
				atomicAdd(ps + d, act - c);
				newv[n] = isfinite(act) ? act : WORST;
			}
			if (base + 8 * (int)blockDim.x < D) {
				// more passes follow: they still need the old neighbours of this pass's last cells; keep it simple: one pass only
			}
			__syncthreads();
			n = 0;
			for (int d = base + threadIdx.x; d < D && n < 8; d += blockDim.x, n++) prev[d] = newv[n];
		}
		const int i2 = i + ((step & 1) ? sv1 : sv0), j2 = j + ((step & 1) ? sh1 : sh0);
		if (!inside(i2, j2)) break;
		i = i2; j = j2; step++;
	}
}

} // namespace stevi

namespace stevi {

static int g_start_pos(int s0, int s1) { return (s0 == 0 && s1 == 0) ? 0 : ((s0 >= 0 && s1 >= 0) ? 1 : 2); }

template<int MODE>
static int32_t launch_lines(stevi_ctx* c, const GLineParams& p) {
	if (p.n_lines <= 0) return STEVI_OK;
	if (p.D > 512) {   // one pass of 8 cells per thread covers D <= 8 * 1024
		const int threads = p.D <= 1024 ? 128 : p.D <= 2048 ? 256 : p.D <= 4096 ? 512 : 1024;
		STEVI_LAUNCH(c, (k_sgm_lines_any<MODE>), dim3(p.n_lines), threads, (size_t)(p.D + 32) * sizeof(float), p);
		return STEVI_OK;
	}
	const int K = p.D <= 32 ? 1 : p.D <= 64 ? 2 : p.D <= 128 ? 4 : p.D <= 256 ? 8 : 16;
	dim3 grid(cdiv(p.n_lines, 4));
	switch (K) {
	case 1: STEVI_LAUNCH(c, (k_sgm_lines<1, MODE>), grid, 128, 0, p); break;
	case 2: STEVI_LAUNCH(c, (k_sgm_lines<2, MODE>), grid, 128, 0, p); break;
	case 4: STEVI_LAUNCH(c, (k_sgm_lines<4, MODE>), grid, 128, 0, p); break;
	case 8: STEVI_LAUNCH(c, (k_sgm_lines<8, MODE>), grid, 128, 0, p); break;
	default: STEVI_LAUNCH(c, (k_sgm_lines<16, MODE>), grid, 128, 0, p); break;
	}
	return STEVI_OK;
}

bool sgm_sweep_capable(const View& cv, int ndir, int strategy, float P1, float P2, int semantics, const View& out);
int32_t sgm_sweep(stevi_ctx* c, const View& cv, int ndir, int strategy, float P1, float P2, float Pout, const stevi_margins& mg,
				  int semantics, const View& out);

int32_t sgm_general(stevi_ctx* c, const View& cv, int ndir, int strategy, float P1, float P2, float Pout,
					const stevi_margins& mg, int semantics, const View& out) {
	const int H = cv.shape[0], W = cv.shape[1], D = cv.shape[2];
	// volumes with contiguous columns (the reference's layout): consumed in place by the sweep kernels (sgm_sweep.cu)
	if (sgm_sweep_capable(cv, ndir, strategy, P1, P2, semantics, out)) return sgm_sweep(c, cv, ndir, strategy, P1, P2, Pout, mg, semantics, out);
	if (D > 8192) return fail(c, STEVI_E_UNSUPPORTED, "sgm: the general recurrence supports up to 8192 disparities");
	const size_t n = (size_t)H * W * D;
	float *Cw = nullptr, *Sw = nullptr;
	STEVI_TRY(arena_alloc_t(c, n, &Cw));
	STEVI_TRY(arena_alloc_t(c, n, &Sw));
	if (cv.dtype == STEVI_U16) STEVI_TRY((relayout<uint16_t, true>(c, cv, Cw))); // static_cast<float>(cv_base.value(..)), sgm.h:233
	else STEVI_TRY((relayout<float, true>(c, cv, Cw)));
	STEVI_CUDA(c, cudaMemcpyAsync(Sw, Cw, n * sizeof(float), cudaMemcpyDeviceToDevice, c->stream)); // sgm.h:371-377
	const bool corrected = semantics == STEVI_CORRECTED;
	const int mode = strategy == STEVI_SCORE ? GM_SCORE : (corrected ? GM_COST : GM_COST_COMPAT);
	GLineParams p;
	p.C = Cw; p.S = Sw; p.H = H; p.W = W; p.D = D;
	p.top = mg.top; p.bottom = mg.bottom; p.left = mg.left; p.right = mg.right;
	p.P1 = P1; p.P2 = P2; p.Pout = Pout;
	const int nrows = H - mg.top - mg.bottom, ncols = W - mg.left - mg.right;
	for (int slot = 0; slot < ndir; slot++) { // directions strictly one after the other, in the reference's order
		const int dir = h_gorder[slot];
		const int* st = h_gsteps[dir];
		const int colStart = g_start_pos(st[0], st[1]), rowStart = g_start_pos(st[2], st[3]);
		p.dir = dir;
		int sj_row = -1;
		for (int kind = 0; kind < 2; kind++) { // lines starting on a column first, then lines starting on a row (sgm.h:329-354)
			const int sp = kind == 0 ? rowStart : colStart;
			if (sp == 0) continue;
			p.kind = kind;
			p.n_lines = kind == 0 ? nrows : ncols;
			// Q4: EndPos is one past the last valid index as written; CORRECTED starts on the last valid index
			if (kind == 0) { p.start_fixed = sp == 1 ? mg.left : W - mg.right - (corrected ? 1 : 0); sj_row = p.start_fixed; }
			else p.start_fixed = sp == 1 ? mg.top : H - mg.bottom - (corrected ? 1 : 0);
			p.skip_col = (corrected && kind == 1 && rowStart != 0) ? sj_row : -1;
			if (nrows <= 0 || ncols <= 0) continue;
			if (mode == GM_SCORE) STEVI_TRY(launch_lines<GM_SCORE>(c, p));
			else if (mode == GM_COST) STEVI_TRY(launch_lines<GM_COST>(c, p));
			else STEVI_TRY(launch_lines<GM_COST_COMPAT>(c, p));
		}
	}
	return relayout<float, false>(c, out, Sw);
}

} // namespace stevi
