// General SGM recurrence (correlation/sgm.h:186-404) on volumes whose columns are contiguous — the reference's own cost-volume
// layout {W*D, 1, W} (cross_correlations.h:220) and any other with a unit column stride — consumed in place: no relayout, no
// copy, the aggregated volume written by plain stores.
//
// Three launches for the 8-direction set (two for 4 directions):
//   k_sgm_sweep_h   the horizontal paths (Left2Right; Right2Left in CORRECTED mode): a warp owns an image row and walks along it,
//                   lane l holds DH consecutive disparities; their contribution `actual_cost - cv_base` (sgm.h:298-300) goes to a
//                   workspace volume A;
//   k_sgm_sweep_v   (down) Up2Down, UpLeft2DownRight and UpRight2DownLeft carried together by one sweep over the rows; reads C and A
//                   once, writes S = (((C + a_U2D) + a_L2R) + a_UL2DR) + a_UR2DL: the reference's own order of additions
//                   (sgm.h:379-388), so the float result is bit-identical to it;
//   k_sgm_sweep_v   (up) the paths that climb (COMPAT: only DownLeft2UpRight lines that start on the left column, Q4), S += a.
//
// The vertical sweep is tiled in SHEARED STRIPS: strip s owns, at sweep step t, the CW columns s*CW + t .. s*CW + t + CW-1 (it
// slides one column to the right per row). A path that runs down-right therefore stays on its lane for its whole life; the
// straight-down path of a column is found one lane to the right in the previous step and the down-left path two lanes to the
// right — so every dependency of a strip points to the strip on its RIGHT, never to the left. Persistent blocks claim strips
// from the rightmost to the leftmost; a strip reads what its right neighbour published for the previous step (two lanes of
// state per step, through global memory = L2, one release/acquire counter per warp) and publishes its own two leftmost lanes.
// One-sided dependencies mean the neighbours settle into a stagger instead of waiting on each other every row, the claim
// order means whatever a strip waits for is already running (no co-residency requirement, no deadlock), and the spin is
// bounded: a wait that does not end raises the ctx fault flag instead of hanging the device.
//
// Inside a strip a thread owns one column and DC consecutive disparities (lanes = 16 columns x 2 disparity chunks, warps = more
// chunks): loads and stores are 64-byte runs along the row, the previous scores of the three paths live in registers, the
// per-column extremum over d is reduced through two shuffles and one shared-memory exchange per step (one __syncthreads per row).
//
// Per cell and path (Score; Cost mirrors it with min/+):  act = c + (max(M - P2, max(L(d-1), L(d+1)) - P1, L(d)) [- Pout] - M),
// M = max_d L(d) over finite previous scores. The reference's loop subtracts P2 only from |od - d| > 1 (sgm.h:236-239); with
// P2 >= P1 >= 0 the three extra candidates L(d) - P2, L(d+-1) - P2 never win, and rounding is monotone, so both give the same
// bits. Other penalties, 16 directions, D > 512 and volumes with strided columns take sgm_general.cu's warp-per-line kernel.
#include "common.cuh"
#include "kernels.cuh"
#include <cmath>
#include <climits>

namespace stevi {

enum { SW_SCORE = 0, SW_COST_COMPAT = 1, SW_COST = 2 };

template<int MODE> __device__ __forceinline__ float sw_worst() { return MODE == SW_SCORE ? -INFINITY : INFINITY; }
template<int MODE> __device__ __forceinline__ float sw_better(float a, float b) { return MODE == SW_SCORE ? fmaxf(a, b) : fminf(a, b); }

// one cell of one path: `in` = previous score at d (non-finite ones already replaced by the worst value), l / r = at d-1 / d+1,
// M = extremum of the previous scores, farv = M -/+ P2. Returns actual_cost (sgm.h:251-254, :291-294).
template<int MODE>
__device__ __forceinline__ float sw_cell(float c, float in, float l, float r, float M, float farv, float P1, float Pout, bool oor) {
	float m;
	if (MODE == SW_SCORE) {
		m = fmaxf(fmaxf(farv, fmaxf(l, r) - P1), in);
		if (oor) m -= Pout;
	} else if (MODE == SW_COST) {
		m = fminf(fminf(farv, fminf(l, r) + P1), in);
		if (oor) m += Pout;
	} else {   // Q3 (sgm.h:281-283): min_a_cost ends up as c_score as soon as one previous score is finite
		m = isfinite(M) ? c : INFINITY;
		if (oor) m += Pout;
	}
	float a = c;
	if (isfinite(m) && isfinite(M)) a += m - M;
	return a;
}

__device__ __forceinline__ int ld_acquire(const int* p) {
	int v;
	asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
	return v;
}
__device__ __forceinline__ void st_release(int* p, int v) {
	asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// ---- horizontal paths ------------------------------------------------------------------------------------------------------

struct HSweep {
	const void* C; int64_t c0, c2; int c_u16;
	float* A; int64_t a0, a2;
	int H, W, D, top, bottom, left, right;
	float P1, P2, Pout;
};

template<int CT> struct VecLoad;
template<> struct VecLoad<1> {
	static __device__ __forceinline__ void f32(const float* p, float* v) { v[0] = __ldg(p); }
	static __device__ __forceinline__ void u16(const uint16_t* p, float* v) { v[0] = (float)__ldg(p); }
	static __device__ __forceinline__ void st(float* p, const float* v) { p[0] = v[0]; }
};
template<> struct VecLoad<4> {
	static __device__ __forceinline__ void f32(const float* p, float* v) {
		const float4 x = __ldg(reinterpret_cast<const float4*>(p));
		v[0] = x.x; v[1] = x.y; v[2] = x.z; v[3] = x.w;
	}
	static __device__ __forceinline__ void u16(const uint16_t* p, float* v) {
		const ushort4 x = __ldg(reinterpret_cast<const ushort4*>(p));
		v[0] = (float)x.x; v[1] = (float)x.y; v[2] = (float)x.z; v[3] = (float)x.w;
	}
	static __device__ __forceinline__ void st(float* p, const float* v) { *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]); }
};

// warp = one image row; lane l = disparities l*DH .. l*DH+DH-1; columns in aligned tiles of CT (CT = 4 needs 16-byte aligned
// rows of C; A is ours and always is). REVERSE walks right to left and adds to what A already holds.
template<int DH, int CT, int MODE, bool REVERSE>
__global__ void __launch_bounds__(128) k_sgm_sweep_h(const HSweep p) {
	const int lane = threadIdx.x & 31;
	const int i = p.top + blockIdx.x * 4 + (threadIdx.x >> 5);
	if (i >= p.H - p.bottom) return;
	const float WORST = sw_worst<MODE>();
	const int d0 = lane * DH;
	const int jb = p.left, je = p.W - p.right;   // columns [jb, je)
	float pf[DH];
#pragma unroll
	for (int k = 0; k < DH; k++) pf[k] = (d0 + k < p.D) ? 0.f : WORST;   // previous_cost starts at 0 (sgm.h:206-208)
	const int64_t crow = (int64_t)i * p.c0 + (int64_t)d0 * p.c2;
	float* arow = p.A + (int64_t)i * p.a0 + (int64_t)d0 * p.a2;
	const int t_first = REVERSE ? (je - 1) / CT : jb / CT, t_last = REVERSE ? jb / CT : (je - 1) / CT;
	auto load_tile = [&](int tile, float (&v)[DH][CT]) {
#pragma unroll
		for (int k = 0; k < DH; k++) {
			if (d0 + k < p.D) {
				const int64_t o = crow + (int64_t)k * p.c2 + (int64_t)tile * CT;
				if (p.c_u16) VecLoad<CT>::u16(static_cast<const uint16_t*>(p.C) + o, v[k]);
				else VecLoad<CT>::f32(static_cast<const float*>(p.C) + o, v[k]);
			} else {
#pragma unroll
				for (int q = 0; q < CT; q++) v[k][q] = 0.f;
			}
		}
	};
	float cn[DH][CT];
	load_tile(t_first, cn);
	for (int tile = t_first; REVERSE ? tile >= t_last : tile <= t_last; tile += REVERSE ? -1 : 1) {
		float c[DH][CT], o[DH][CT];
#pragma unroll
		for (int k = 0; k < DH; k++)
#pragma unroll
			for (int q = 0; q < CT; q++) c[k][q] = cn[k][q];
		if (tile != t_last) load_tile(tile + (REVERSE ? -1 : 1), cn);   // the next tile's loads fly while this one is worked on
		const bool full = tile * CT >= jb && tile * CT + CT <= je;
		if (REVERSE || !full) {   // REVERSE adds to the forward pass's contributions; partial tiles keep what they do not own
#pragma unroll
			for (int k = 0; k < DH; k++) {
				if (d0 + k < p.D) VecLoad<CT>::f32(arow + (int64_t)k * p.a2 + (int64_t)tile * CT, o[k]);
			}
		}
#pragma unroll
		for (int qq = 0; qq < CT; qq++) {
			const int q = REVERSE ? CT - 1 - qq : qq;
			const int j = tile * CT + q;
			if (j < jb || j >= je) continue;
			float mx = WORST;
#pragma unroll
			for (int k = 0; k < DH; k++) mx = sw_better<MODE>(mx, pf[k]);
#pragma unroll
			for (int s = 16; s > 0; s >>= 1) mx = sw_better<MODE>(mx, __shfl_xor_sync(0xffffffffu, mx, s));
			const float M = mx;
			float nlo = __shfl_up_sync(0xffffffffu, pf[DH - 1], 1), nhi = __shfl_down_sync(0xffffffffu, pf[0], 1);
			if (lane == 0) nlo = WORST;
			if (lane == 31) nhi = WORST;
			const float farv = MODE == SW_SCORE ? M - p.P2 : M + p.P2;
			const int kout = p.W - j - d0;   // j + d >= W  <=>  k >= kout
			float nf[DH];
#pragma unroll
			for (int k = 0; k < DH; k++) {
				const float l = k > 0 ? pf[k > 0 ? k - 1 : 0] : nlo, r = k < DH - 1 ? pf[k < DH - 1 ? k + 1 : 0] : nhi;
				const float a = sw_cell<MODE>(c[k][q], pf[k], l, r, M, farv, p.P1, p.Pout, k >= kout);
				const float contrib = a - c[k][q];
				o[k][q] = REVERSE ? o[k][q] + contrib : contrib;
				nf[k] = (isfinite(a) && d0 + k < p.D) ? a : WORST;
			}
#pragma unroll
			for (int k = 0; k < DH; k++) pf[k] = nf[k];
		}
#pragma unroll
		for (int k = 0; k < DH; k++)
			if (d0 + k < p.D) VecLoad<CT>::st(arow + (int64_t)k * p.a2 + (int64_t)tile * CT, o[k]);
	}
}

// ---- vertical and diagonal paths: sheared strips -----------------------------------------------------------------------------

constexpr int SW_CW = 16;          // columns per strip
constexpr int SW_QL = 32 / SW_CW;  // disparity chunks per warp
constexpr int SW_MAXW = 16;        // warps per block
constexpr int SW_SPIN_LIMIT = 1 << 22;

struct VSweep {
	const void* C; int64_t c0, c2; int c_u16;
	float* S; int64_t s0, s2;
	const float* A; int64_t a0, a2;        // horizontal contributions (nullptr: none), added after the straight path's
	float* xchg; int* progress; int* claim; int* fault;
	int H, W, D, Dpad, top, left, nrows, ncols;
	int sweep;                             // +1: rows top -> bottom, -1: bottom -> top
	int en[3];                             // path with shift k (0: along the shear, 1: straight, 2: against the shear) is live
	int rowlive[3], collive[3];            // its lines started by the row-start / the column-start loop are live (Q4)
	int corner_once;                       // CORRECTED: the corner line is started once (COMPAT: twice, Q4)
	int accumulate;                        // 0: S = C + contributions, 1: S += contributions
	int s_lo, s_hi, n_strips;
	float P1, P2, Pout;
};

template<int DC, int MAXW, int MODE>
__global__ void __launch_bounds__(MAXW * 32) k_sgm_sweep_v(const VSweep p) {
	constexpr int CW = SW_CW, QL = SW_QL;
	__shared__ float red[2][3][SW_MAXW][CW];        // per-warp extremum of a column
	__shared__ float wlo[2][3][SW_MAXW][CW];        // previous score at the warp's lowest disparity
	__shared__ float whi[2][3][SW_MAXW][CW];        // ... and at its highest
	__shared__ int s_claim;
	const float WORST = sw_worst<MODE>();
	const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, NW = blockDim.x >> 5;
	const int x = lane % CW, q = lane / CW;
	const int d0 = (w * QL + q) * DC;
	for (;;) {
		if (threadIdx.x == 0) s_claim = atomicAdd(p.claim, 1);
		__syncthreads();
		const int claim = s_claim;
		__syncthreads();
		if (claim >= p.n_strips) return;
		const int s = p.s_hi - claim;   // right to left: whatever this strip waits for has been claimed before
		const int t_begin = max(0, -s * CW - (CW - 1)), t_end = min(p.nrows, p.ncols - s * CW);
		const bool has_right = s + 1 <= p.s_hi;
		const int rt_begin = max(0, -(s + 1) * CW - (CW - 1)), rt_end = min(p.nrows, p.ncols - (s + 1) * CW);
		const float* xin = p.xchg + (int64_t)(s + 1 - p.s_lo) * p.nrows * 3 * p.Dpad + d0;
		float* xout = p.xchg + (int64_t)(s - p.s_lo) * p.nrows * 3 * p.Dpad + d0;
		const int* prog_in = p.progress + (int64_t)(s + 1 - p.s_lo) * SW_MAXW + w;
		int* prog_out = p.progress + (int64_t)(s - p.s_lo) * SW_MAXW + w;
		int seen = 0;
		float st[3][DC];   // what this lane computed in the previous step, per path
#pragma unroll
		for (int k = 0; k < 3; k++)
#pragma unroll
			for (int m = 0; m < DC; m++) st[k][m] = WORST;
		auto row_of = [&](int t) { return p.sweep > 0 ? p.top + t : p.top + p.nrows - 1 - t; };
		auto load_c = [&](int t, int cc, float (&v)[DC]) {
			const bool ok = cc >= 0 && cc < p.ncols && t < t_end;
			const int64_t o = (int64_t)row_of(t) * p.c0 + (int64_t)d0 * p.c2 + (p.left + cc);
#pragma unroll
			for (int m = 0; m < DC; m++) {
				v[m] = 0.f;
				if (ok && d0 + m < p.D) v[m] = p.c_u16 ? (float)__ldg(static_cast<const uint16_t*>(p.C) + o + m * p.c2)
												   : __ldg(static_cast<const float*>(p.C) + o + m * p.c2);
			}
		};
		float cn[DC];
		load_c(t_begin, s * CW + t_begin + x, cn);
		for (int t = t_begin; t < t_end; t++) {
			const int buf = t & 1;
			const int cc = s * CW + t + x;   // column inside the margin rectangle
			const bool valid = cc >= 0 && cc < p.ncols;
			const int i = row_of(t), j = p.left + cc;
			float c[DC];
#pragma unroll
			for (int m = 0; m < DC; m++) c[m] = cn[m];
			load_c(t + 1, cc + 1, cn);
			// this step's other operands: issued now, used after the recurrences
			float av[DC], sacc[DC];
			{
				const int64_t oa = (int64_t)i * p.a0 + (int64_t)d0 * p.a2 + j;
				const int64_t os = (int64_t)i * p.s0 + (int64_t)d0 * p.s2 + j;
#pragma unroll
				for (int m = 0; m < DC; m++) {
					av[m] = 0.f; sacc[m] = 0.f;
					if (valid && d0 + m < p.D) {
						if (p.A) av[m] = __ldg(p.A + oa + m * p.a2);
						if (p.accumulate) sacc[m] = p.S[os + m * p.s2];
					}
				}
			}
			// ---- inputs of the three paths
			// shift 1 / 2: the state of the lane one / two to the right; beyond the strip: what the right neighbour published
			const bool need_in = has_right && t >= 1 && t - 1 >= rt_begin && t - 1 < rt_end && x >= CW - 2 && (p.en[1] || p.en[2]);
			if (need_in && seen < t) {
				int spins = 0;
				while ((seen = ld_acquire(prog_in)) < t) {
					if (++spins > SW_SPIN_LIMIT) { *p.fault = 1; seen = INT_MAX; break; }   // raise the fault flag instead of hanging
					__nanosleep(20);
				}
			}
			float in[3][DC];
#pragma unroll
			for (int m = 0; m < DC; m++) in[0][m] = st[0][m];
#pragma unroll
			for (int k = 1; k < 3; k++) {
				if (!p.en[k]) continue;
#pragma unroll
				for (int m = 0; m < DC; m++) in[k][m] = __shfl_down_sync(0xffffffffu, st[k][m], k, CW);
				if (x + k >= CW) {
					if (need_in) {
						const int slot = k == 1 ? 0 : 1 + (x + 2 - CW);
						const float4* src = reinterpret_cast<const float4*>(xin + ((int64_t)(t - 1) * 3 + slot) * p.Dpad);
#pragma unroll
						for (int m4 = 0; m4 < DC / 4; m4++) {
							const float4 v = __ldcg(src + m4);
							in[k][4 * m4] = v.x; in[k][4 * m4 + 1] = v.y; in[k][4 * m4 + 2] = v.z; in[k][4 * m4 + 3] = v.w;
						}
					} else {
#pragma unroll
						for (int m = 0; m < DC; m++) in[k][m] = WORST;
					}
				}
			}
			// a path whose previous pixel lies outside the margin rectangle starts here: previous_cost = 0 (sgm.h:206-208)
			const bool fresh[3] = {t == 0 || cc == 0, t == 0, t == 0 || cc == p.ncols - 1};
#pragma unroll
			for (int k = 0; k < 3; k++)
				if (fresh[k]) {
#pragma unroll
					for (int m = 0; m < DC; m++) in[k][m] = (d0 + m < p.D) ? 0.f : WORST;
				}
			// ---- per-column extremum over d and the neighbours across chunk borders
			float nlo[3], nhi[3];
#pragma unroll
			for (int k = 0; k < 3; k++) {
				if (!p.en[k]) continue;
				float mx = WORST;
#pragma unroll
				for (int m = 0; m < DC; m++) mx = sw_better<MODE>(mx, in[k][m]);
#pragma unroll
				for (int o = CW; o < 32; o <<= 1) mx = sw_better<MODE>(mx, __shfl_xor_sync(0xffffffffu, mx, o));
				nlo[k] = __shfl_up_sync(0xffffffffu, in[k][DC - 1], CW);     // chunk q-1 of this warp
				nhi[k] = __shfl_down_sync(0xffffffffu, in[k][0], CW);        // chunk q+1
				if (q == 0) { red[buf][k][w][x] = mx; wlo[buf][k][w][x] = in[k][0]; }
				if (q == QL - 1) whi[buf][k][w][x] = in[k][DC - 1];
			}
			__syncthreads();
			// line weights (Q4): how many times the reference walks the line through this pixel for each path
			int wt[3];
#pragma unroll
			for (int k = 0; k < 3; k++) {
				const int sh = 1 - k;
				const int kr = sh > 0 ? cc : (sh < 0 ? p.ncols - 1 - cc : INT_MAX), kc = t;
				int n = ((kr <= kc && p.rowlive[k]) ? 1 : 0) + ((kc <= kr && p.collive[k]) ? 1 : 0);
				if (n == 2 && p.corner_once) n = 1;
				wt[k] = (p.en[k] && valid) ? n : 0;
			}
			if (!p.accumulate) {
#pragma unroll
				for (int m = 0; m < DC; m++) sacc[m] = c[m];   // sgm.h:371-377
			}
			const int kout = p.W - j - d0;
			// the reference's order of additions: Up2Down, Left2Right, UpLeft2DownRight, UpRight2DownLeft (sgm.h:379-388)
			const int order[3] = {1, 0, 2};
#pragma unroll
			for (int oi = 0; oi < 3; oi++) {
				const int k = order[oi];
				if (p.en[k]) {
					float M = WORST;
					for (int ww = 0; ww < NW; ww++) M = sw_better<MODE>(M, red[buf][k][ww][x]);
					float lo = nlo[k], hi = nhi[k];
					if (q == 0) lo = w > 0 ? whi[buf][k][w > 0 ? w - 1 : 0][x] : WORST;
					if (q == QL - 1) hi = w + 1 < NW ? wlo[buf][k][w + 1 < NW ? w + 1 : 0][x] : WORST;
					const float farv = MODE == SW_SCORE ? M - p.P2 : M + p.P2;
#pragma unroll
					for (int m = 0; m < DC; m++) {
						const float l = m > 0 ? in[k][m > 0 ? m - 1 : 0] : lo, r = m < DC - 1 ? in[k][m < DC - 1 ? m + 1 : 0] : hi;
						const float a = sw_cell<MODE>(c[m], in[k][m], l, r, M, farv, p.P1, p.Pout, m >= kout);
						const float contrib = a - c[m];
						if (wt[k] >= 1) sacc[m] += contrib;
						if (wt[k] >= 2) sacc[m] += contrib;
						st[k][m] = (isfinite(a) && d0 + m < p.D) ? a : WORST;
					}
				}
				if (oi == 0 && p.A && valid) {
#pragma unroll
					for (int m = 0; m < DC; m++) sacc[m] += av[m];
				}
			}
			if (valid) {
				const int64_t os = (int64_t)i * p.s0 + (int64_t)d0 * p.s2 + j;
#pragma unroll
				for (int m = 0; m < DC; m++)
					if (d0 + m < p.D) p.S[os + m * p.s2] = sacc[m];
			}
			// ---- publish the two leftmost lanes for the strip on the left
			if (s > p.s_lo && (p.en[1] || p.en[2])) {
				if (x < 2) {
					float* dst = xout + (int64_t)t * 3 * p.Dpad;
					if (x == 0 && p.en[1]) {
#pragma unroll
						for (int m4 = 0; m4 < DC / 4; m4++)
							__stcg(reinterpret_cast<float4*>(dst) + m4, make_float4(st[1][4 * m4], st[1][4 * m4 + 1], st[1][4 * m4 + 2], st[1][4 * m4 + 3]));
					}
					if (p.en[2]) {
						float4* d2 = reinterpret_cast<float4*>(dst + (1 + x) * p.Dpad);
#pragma unroll
						for (int m4 = 0; m4 < DC / 4; m4++)
							__stcg(d2 + m4, make_float4(st[2][4 * m4], st[2][4 * m4 + 1], st[2][4 * m4 + 2], st[2][4 * m4 + 3]));
					}
					__threadfence();
				}
				__syncwarp();
				if (lane == 0) st_release(prog_out, t + 1);
			}
		}
	}
}

// S = (float) C over the whole volume (only needed when margins leave a frame the sweeps never touch)
template<typename T>
__global__ void __launch_bounds__(256) k_sgm_copy_cv(const T* __restrict__ C, int64_t c0, int64_t c2, float* __restrict__ S, int64_t s0, int64_t s2,
													  int H, int W, int D) {
	const int j = blockIdx.x * 256 + threadIdx.x, d = blockIdx.y, i = blockIdx.z;
	if (j < W) S[i * s0 + d * s2 + j] = (float)C[i * c0 + d * c2 + j];
}

template<int MODE, bool REVERSE>
static int32_t launch_h(stevi_ctx* c, const HSweep& p, bool vec) {
	const int rows = p.H - p.top - p.bottom;
	dim3 grid(cdiv(rows, 4));
	const int dh = cdiv(p.D, 32);
#define STEVI_H(DH) do { if (vec) STEVI_LAUNCH(c, (k_sgm_sweep_h<DH, 4, MODE, REVERSE>), grid, 128, 0, p); \
	else STEVI_LAUNCH(c, (k_sgm_sweep_h<DH, 1, MODE, REVERSE>), grid, 128, 0, p); } while (0)
	if (dh <= 1) STEVI_H(1); else if (dh <= 2) STEVI_H(2); else if (dh <= 4) STEVI_H(4); else if (dh <= 8) STEVI_H(8); else STEVI_H(16);
#undef STEVI_H
	return STEVI_OK;
}

template<int MODE>
static int32_t launch_v(stevi_ctx* c, const VSweep& p) {
	const int dc = p.D <= 64 ? 4 : p.D <= 128 ? 8 : 16;
	const int nw = std::max(1, cdiv(cdiv(p.D, dc), SW_QL));
	dim3 grid(std::min(p.n_strips, c->sm_count * (nw <= 8 ? 2 : 1)));
	if (dc == 4) STEVI_LAUNCH(c, (k_sgm_sweep_v<4, 8, MODE>), grid, nw * 32, 0, p);
	else if (dc == 8) STEVI_LAUNCH(c, (k_sgm_sweep_v<8, 8, MODE>), grid, nw * 32, 0, p);
	else if (nw <= 8) STEVI_LAUNCH(c, (k_sgm_sweep_v<16, 8, MODE>), grid, nw * 32, 0, p);
	else STEVI_LAUNCH(c, (k_sgm_sweep_v<16, 16, MODE>), grid, nw * 32, 0, p);
	return STEVI_OK;
}

bool sgm_sweep_capable(const View& cv, int ndir, int strategy, float P1, float P2, int semantics, const View& out) {
	static const bool off = getenv("STEVI_SGM_NO_SWEEP") != nullptr;   // measurement aid: the warp-per-line kernel only
	if (off) return false;
	if (ndir != 4 && ndir != 8) return false;
	if (cv.st[1] != 1 || out.st[1] != 1 || cv.shape[2] > 512) return false;
	const bool compat_cost = strategy == STEVI_COST && semantics != STEVI_CORRECTED;
	if (!compat_cost && !(P2 >= P1 && P1 >= 0.f && std::isfinite(P2))) return false;
	return true;
}

int32_t sgm_sweep(stevi_ctx* c, const View& cv, int ndir, int strategy, float P1, float P2, float Pout, const stevi_margins& mg,
				  int semantics, const View& out) {
	const int H = cv.shape[0], W = cv.shape[1], D = cv.shape[2];
	const bool corrected = semantics == STEVI_CORRECTED;
	const int mode = strategy == STEVI_SCORE ? SW_SCORE : (corrected ? SW_COST : SW_COST_COMPAT);
	const int nrows = H - mg.top - mg.bottom, ncols = W - mg.left - mg.right;
	const bool framed = mg.top || mg.bottom || mg.left || mg.right;
	if (framed || nrows <= 0 || ncols <= 0) {
		dim3 g(cdiv(W, 256), D, H);
		if (cv.dtype == STEVI_U16) STEVI_LAUNCH(c, k_sgm_copy_cv<uint16_t>, g, 256, 0, cv.as<uint16_t>(), cv.st[0], cv.st[2], out.as<float>(), out.st[0], out.st[2], H, W, D);
		else STEVI_LAUNCH(c, k_sgm_copy_cv<float>, g, 256, 0, cv.as<float>(), cv.st[0], cv.st[2], out.as<float>(), out.st[0], out.st[2], H, W, D);
		if (nrows <= 0 || ncols <= 0) return STEVI_OK;
	}
	// ---- horizontal paths -> A (dense [H][D][W4], rows 16-byte aligned)
	const int W4 = (W + 3) / 4 * 4;
	float* A = nullptr;
	STEVI_TRY(arena_alloc_t(c, (size_t)H * D * W4, &A));
	HSweep h;
	h.C = cv.p; h.c0 = cv.st[0]; h.c2 = cv.st[2]; h.c_u16 = cv.dtype == STEVI_U16;
	h.A = A; h.a0 = (int64_t)D * W4; h.a2 = W4;
	h.H = H; h.W = W; h.D = D; h.top = mg.top; h.bottom = mg.bottom; h.left = mg.left; h.right = mg.right;
	h.P1 = P1; h.P2 = P2; h.Pout = Pout;
	const size_t esz = dtype_size(cv.dtype);
	const bool vec = (uintptr_t)cv.p % (4 * esz) == 0 && cv.st[0] % 4 == 0 && cv.st[2] % 4 == 0 && W % 4 == 0;
	if (mode == SW_SCORE) STEVI_TRY((launch_h<SW_SCORE, false>(c, h, vec)));
	else if (mode == SW_COST) STEVI_TRY((launch_h<SW_COST, false>(c, h, vec)));
	else STEVI_TRY((launch_h<SW_COST_COMPAT, false>(c, h, vec)));
	if (corrected) {   // Right2Left is live only with the corrected start rule (Q4)
		if (mode == SW_SCORE) STEVI_TRY((launch_h<SW_SCORE, true>(c, h, vec)));
		else STEVI_TRY((launch_h<SW_COST, true>(c, h, vec)));
	}
	// ---- vertical sweeps
	VSweep v;
	v.C = cv.p; v.c0 = cv.st[0]; v.c2 = cv.st[2]; v.c_u16 = h.c_u16;
	v.S = out.as<float>(); v.s0 = out.st[0]; v.s2 = out.st[2];
	v.a0 = h.a0; v.a2 = h.a2;
	v.H = H; v.W = W; v.D = D; v.top = mg.top; v.left = mg.left; v.nrows = nrows; v.ncols = ncols;
	const int dc = D <= 64 ? 4 : D <= 128 ? 8 : 16;
	v.Dpad = cdiv(D, dc * SW_QL) * dc * SW_QL;
	v.P1 = P1; v.P2 = P2; v.Pout = Pout;
	v.corner_once = corrected ? 1 : 0;
	v.fault = c->fault_dev;
	v.s_lo = -((nrows - 1 + SW_CW - 1) / SW_CW);
	const int s_all = (ncols - 1) / SW_CW;
	const size_t xn = (size_t)(s_all - v.s_lo + 1) * nrows * 3 * v.Dpad;
	const size_t pn = (size_t)(s_all - v.s_lo + 1) * SW_MAXW + 1;
	int* flags = nullptr;
	STEVI_TRY(arena_alloc_t(c, xn, &v.xchg));
	STEVI_TRY(arena_alloc_t(c, 2 * pn, &flags));
	STEVI_CUDA(c, cudaMemsetAsync(flags, 0, 2 * pn * sizeof(int), c->stream));
	for (int pass = 0; pass < 2; pass++) {
		v.sweep = pass == 0 ? 1 : -1;
		v.A = pass == 0 ? A : nullptr;
		v.accumulate = (pass == 1 || framed) ? 1 : 0;
		v.progress = flags + pass * pn;
		v.claim = v.progress + pn - 1;
		bool any = false, any_col = false;
		for (int k = 0; k < 3; k++) {
			const int sh = 1 - k;
			const bool in_set = ndir == 8 || sh == 0;
			// sgm.h:162-184 with :331,:345: ZeroPos starts are live; EndPos starts (negative steps) only with the corrected rule
			v.rowlive[k] = in_set && sh != 0 && (sh > 0 || corrected);
			v.collive[k] = in_set && (v.sweep > 0 || corrected);
			v.en[k] = v.rowlive[k] || v.collive[k];
			any |= v.en[k] != 0;
			any_col |= v.collive[k] != 0;
		}
		if (!any && !(pass == 0)) continue;
		// strips that hold live pixels: lines started by the column loop reach everywhere, lines started on the left column
		// (the only row starts of a sweep without live column starts) stay at or left of the diagonal through the corner
		v.s_hi = (any_col || pass == 0) ? s_all : 0;
		v.n_strips = v.s_hi - v.s_lo + 1;
		if (mode == SW_SCORE) STEVI_TRY(launch_v<SW_SCORE>(c, v));
		else if (mode == SW_COST) STEVI_TRY(launch_v<SW_COST>(c, v));
		else STEVI_TRY(launch_v<SW_COST_COMPAT>(c, v));
	}
	return STEVI_OK;
}

} // namespace stevi
