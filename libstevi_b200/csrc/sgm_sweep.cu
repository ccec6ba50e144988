// General SGM recurrence (correlation/sgm.h:186-404) on volumes whose columns are contiguous — the reference's own cost-volume
// layout {W*D, 1, W} (cross_correlations.h:220) and any other with a unit column stride — consumed in place: no relayout, no
// copy, the aggregated volume written by plain stores.
//
// Three launches for the 8-direction set (two for 4 directions):
//   k_sgm_sweep_h   the horizontal paths (Left2Right; Right2Left in CORRECTED mode): a warp owns an image row and walks along it,
//                   lane l holds DH consecutive disparities; their contribution `actual_cost - cv_base` (sgm.h:298-300) goes to a
//                   workspace volume A;
//   k_sgm_sweep_t   (down) Up2Down, UpLeft2DownRight and UpRight2DownLeft carried together by one sweep over the rows; reads C and A
//                   once, writes S = (((C + a_U2D) + a_L2R) + a_UL2DR) + a_UR2DL: the reference's own order of additions
//                   (sgm.h:379-388), so the float result is bit-identical to it;
//   k_sgm_sweep_t   (up) the paths that climb (COMPAT: only DownLeft2UpRight lines that start on the left column, Q4), S += a.
//
// The vertical sweep is tiled in SHEARED STRIPS: strip s owns, at sweep step t, the 16 columns s*16 + t .. s*16 + t + 15 (it
// slides one column to the right per row). A path that runs down-right therefore stays with its warp for its whole life; the
// straight-down path of a column is found one column to the right in the previous step and the down-left path two columns to
// the right — so every dependency of a strip points to the strip on its RIGHT, never to the left. Persistent blocks claim
// strips from the rightmost to the leftmost; a strip reads what its right neighbour published for the previous row (three
// vectors of D scores per row, through global memory = L2) and publishes its own. One-sided dependencies mean the neighbours
// settle into a stagger instead of waiting on each other every row, the claim order means whatever a strip waits for is
// already running (no co-residency requirement, no deadlock), and the wait is bounded: one that does not end raises the ctx
// fault flag instead of hanging the device. The kernel comment of k_sgm_sweep_t has the inside of a strip.
//
// Per cell and path (Score; Cost mirrors it with min/+):  act = c + (max(M - P2, max(L(d-1), L(d+1)) - P1, L(d)) [- Pout] - M),
// M = max_d L(d) over finite previous scores. The reference's loop subtracts P2 only from |od - d| > 1 (sgm.h:236-239); with
// P2 >= P1 >= 0 the three extra candidates L(d) - P2, L(d+-1) - P2 never win, and rounding is monotone, so both give the same
// bits. Other penalties, 16 directions, D > 512 and volumes with strided columns take sgm_general.cu's line kernels.
#include "common.cuh"
#include "kernels.cuh"
#include <cuda.h>   // CUtensorMap: the type only; cuTensorMapEncodeTiled is fetched through cudaGetDriverEntryPoint (no libcuda link)
#include <cmath>
#include <climits>

namespace stevi {

enum { SW_SCORE = 0, SW_COST_COMPAT = 1, SW_COST = 2 };

template<int MODE> __device__ __forceinline__ float sw_worst() { return MODE == SW_SCORE ? -INFINITY : INFINITY; }
template<int MODE> __device__ __forceinline__ float sw_better(float a, float b) { return MODE == SW_SCORE ? fmaxf(a, b) : fminf(a, b); }
// extremum over the warp in ONE instruction (sm_100a: redux.sync on f32, CREDUX in SASS) instead of five dependent shuffle
// rounds. The operands are never NaN here (non-finite scores have been replaced by the worst value), so this is the value the
// butterfly of fmaxf / fminf gave. (Measured: no change in kernel time — see the note at k_sgm_sweep_h on what bounds it.)
template<int MODE> __device__ __forceinline__ float sw_warp_best(float v) {
	float r;
	if (MODE == SW_SCORE) asm volatile("redux.sync.max.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
	else asm volatile("redux.sync.min.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
	return r;
}

// one cell of one path: `in` = previous score at d (non-finite ones already replaced by the worst value), l / r = at d-1 / d+1,
// farv = M -/+ P2 with M = extremum of the previous scores, FINITE (the caller handles the other case: actual_cost = c).
// Returns actual_cost (sgm.h:251-254, :291-294). The reference's test isfinite(max_a_cost) always holds here in Score and
// corrected Cost mode (M - P2 <= m <= M); in COMPAT Cost mode m = c_score [+ Pout], and for a non-finite c_score the reference's
// "leave actual_cost = c" and c + (m - M) are the same value (NaN stays NaN, +-inf stays +-inf).
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
		m = c;
		if (oor) m += Pout;
	}
	return c + (m - M);
}

// ---- horizontal paths ------------------------------------------------------------------------------------------------------

struct HSweep {
	const void* C; int64_t c0, c2; int c_u16;
	float* A; int64_t a0, a2;
	int H, W, D, top, bottom, left, right;
	float P1, P2, Pout;
	int pf;   // L2 prefetch distance in tiles (0: none)
	int* fault;   // ctx fault flag: raised instead of waiting for ever on a tile that does not arrive
};

__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

// a tile of CT columns of one disparity row: CT = 8 moves whole 32-byte sectors (two 16-byte accesses; one for u16 costs), so
// no sector of C is fetched twice and no sector of A is written in halves; CT = 1 is the any-alignment path
template<int CT> struct Tile;
template<> struct Tile<1> {
	static __device__ __forceinline__ void f32(const float* p, float* v) { v[0] = __ldg(p); }
	static __device__ __forceinline__ void u16(const uint16_t* p, float* v) { v[0] = (float)__ldg(p); }
	static __device__ __forceinline__ void st(float* p, const float* v) { p[0] = v[0]; }
};
template<> struct Tile<8> {
	// one 256-bit access per sector (sm_100a: LDG.E.256 / STG.E.256): the lanes of a warp address 32 different rows of the volume, so
	// every instruction costs 32 tag look-ups whatever its width
	static __device__ __forceinline__ void f32(const float* p, float* v) {
		asm volatile("ld.global.nc.v8.f32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
					 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]) : "l"(p));
	}
	static __device__ __forceinline__ void u16(const uint16_t* p, float* v) {
		const uint4 x = __ldg(reinterpret_cast<const uint4*>(p));
		v[0] = (float)(x.x & 0xffffu); v[1] = (float)(x.x >> 16); v[2] = (float)(x.y & 0xffffu); v[3] = (float)(x.y >> 16);
		v[4] = (float)(x.z & 0xffffu); v[5] = (float)(x.z >> 16); v[6] = (float)(x.w & 0xffffu); v[7] = (float)(x.w >> 16);
	}
	static __device__ __forceinline__ void st(float* p, const float* v) {
		asm volatile("st.global.v8.f32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
					 :: "l"(p), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]), "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7]) : "memory");
	}
};

// one column step of one path for the DN consecutive disparities a thread owns. st: previous scores in, new scores out (non-finite
// ones replaced by the worst value, so that the extremum and the neighbour terms skip them like the reference's isfinite tests);
// lo / hi: previous scores just below / above the thread's range; acc += actual_cost - c (sgm.h:298-300), twice for a line the
// reference walks twice (Q4). c is NaN for disparities past D: those cells stay at the worst value by themselves.
template<int DN, int MODE, bool POUT, bool TWICE>
__device__ __forceinline__ void sw_path(const float (&c)[DN], float (&st)[DN], float lo, float hi, float M, float P1, float P2, float Pout,
										int kout, float (&acc)[DN], bool twice) {
	const float WORST = sw_worst<MODE>();
	if (!isfinite(M)) {   // no finite previous score at all (warp-uniform): actual_cost = c, nothing is added (sgm.h:252, :292)
#pragma unroll
		for (int m = 0; m < DN; m++) {
			const float contrib = c[m] - c[m];   // 0, or NaN for a non-finite cost: what `actual_cost - cv_base` gives (sgm.h:299)
			acc[m] += contrib;
			if (TWICE && twice) acc[m] += contrib;
			st[m] = isfinite(c[m]) ? c[m] : WORST;
		}
		return;
	}
	const float farv = MODE == SW_SCORE ? M - P2 : M + P2;
	float left = lo;
#pragma unroll
	for (int m = 0; m < DN; m++) {
		const float cur = st[m];
		const float right = m < DN - 1 ? st[m < DN - 1 ? m + 1 : 0] : hi;
		const float a = sw_cell<MODE>(c[m], cur, left, right, M, farv, P1, Pout, POUT && m >= kout);
		const float contrib = a - c[m];
		acc[m] += contrib;
		if (TWICE && twice) acc[m] += contrib;
		left = cur;
		st[m] = isfinite(a) ? a : WORST;
	}
}

// warp = one image row; lane l = disparities l*DH .. l*DH+DH-1; columns in aligned tiles of CT. The contributions replace the
// costs in the tile registers and leave as whole tiles. REVERSE walks right to left and adds to what A already holds.
// What bounds it (ncu source page, round 2, profiles/ncu_source_r2c_k_sgm_sweep.txt): 65 % of the stall samples are waits for the
// next tile's loads and for the stores at the tile boundaries, and neither a longer prefetch distance (2, 3, 6 tiles: 1.80, 1.86,
// 2.36 ms), nor whole-line L2 fetches (ld.global.L2::128B), nor the register budget (168 -> 197, spill gone) moved it: every
// access of a warp goes to 32 different rows of the volume, i.e. 32 tag look-ups per instruction whatever its width (l1tex 58 %,
// lts 56 % busy at 0.7 IPC). What did move it is fewer instructions through that stage: 256-bit loads and stores (1.80 -> 1.57
// ms) and no prefetch instructions at all (-> 1.43 ms; the loads of the next tile are in flight for a whole tile anyway). What
// is left is a load path that does not go through the L1 tag stage lane by lane (a tensor-map TMA box [DH*32 rows][8 columns]
// per warp and tile); not built this round.
template<int DH, int CT, int MODE, bool REVERSE>
__global__ void __launch_bounds__(128, 2) k_sgm_sweep_h(const HSweep p) {
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
	const int dt = REVERSE ? -1 : 1;
	const float nanv = __int_as_float(0x7fc00000);
	auto load_tile = [&](int tile, float (&v)[DH][CT]) {
#pragma unroll
		for (int k = 0; k < DH; k++) {
			if (d0 + k < p.D) {
				const int64_t o = crow + (int64_t)k * p.c2 + (int64_t)tile * CT;
				if (p.c_u16) Tile<CT>::u16(static_cast<const uint16_t*>(p.C) + o, v[k]);
				else Tile<CT>::f32(static_cast<const float*>(p.C) + o, v[k]);
			} else {
#pragma unroll
				for (int q = 0; q < CT; q++) v[k][q] = nanv;
			}
		}
	};
	// the sectors two tiles ahead are pulled into L2 while the next tile's loads are in flight
	auto prefetch_tile = [&](int tile) {
		if (CT == 1 || p.pf == 0 || (REVERSE ? tile < t_last : tile > t_last)) return;
#pragma unroll
		for (int k = 0; k < DH; k++) {
			if (d0 + k < p.D) {
				const int64_t o = crow + (int64_t)k * p.c2 + (int64_t)tile * CT;
				prefetch_l2(p.c_u16 ? (const void*)(static_cast<const uint16_t*>(p.C) + o) : (const void*)(static_cast<const float*>(p.C) + o));
			}
		}
	};
	float cn[DH][CT];
	load_tile(t_first, cn);
	for (int k = 1; k < p.pf; k++) prefetch_tile(t_first + k * dt);
	for (int tile = t_first; REVERSE ? tile >= t_last : tile <= t_last; tile += dt) {
		float c[DH][CT];
#pragma unroll
		for (int k = 0; k < DH; k++)
#pragma unroll
			for (int q = 0; q < CT; q++) c[k][q] = cn[k][q];
		if (tile != t_last) load_tile(tile + dt, cn);   // the next tile's loads fly while this one is worked on
		prefetch_tile(tile + p.pf * dt);
		const bool full = tile * CT >= jb && tile * CT + CT <= je;
		float old[REVERSE ? DH : 1][CT];
		if (REVERSE) {
#pragma unroll
			for (int k = 0; k < DH; k++)
				if (d0 + k < p.D) Tile<CT>::f32(arow + (int64_t)k * p.a2 + (int64_t)tile * CT, old[k]);
		}
#pragma unroll
		for (int qq = 0; qq < CT; qq++) {
			const int q = REVERSE ? CT - 1 - qq : qq;
			const int j = tile * CT + q;
			if (j < jb || j >= je) continue;
			float mx = WORST;
#pragma unroll
			for (int k = 0; k < DH; k++) mx = sw_better<MODE>(mx, pf[k]);
			mx = sw_warp_best<MODE>(mx);
			float nlo = __shfl_up_sync(0xffffffffu, pf[DH - 1], 1), nhi = __shfl_down_sync(0xffffffffu, pf[0], 1);
			if (lane == 0) nlo = WORST;
			if (lane == 31) nhi = WORST;
			float cc[DH], acc[DH];
#pragma unroll
			for (int k = 0; k < DH; k++) { cc[k] = c[k][q]; acc[k] = REVERSE ? old[REVERSE ? k : 0][q] : 0.f; }
			// j + d >= W (the Pout term) only concerns the last D columns
			if (j + p.D <= p.W) sw_path<DH, MODE, false, false>(cc, pf, nlo, nhi, mx, p.P1, p.P2, p.Pout, 0, acc, false);
			else sw_path<DH, MODE, true, false>(cc, pf, nlo, nhi, mx, p.P1, p.P2, p.Pout, p.W - j - d0, acc, false);
#pragma unroll
			for (int k = 0; k < DH; k++) c[k][q] = acc[k];
		}
		if (full) {
#pragma unroll
			for (int k = 0; k < DH; k++)
				if (d0 + k < p.D) Tile<CT>::st(arow + (int64_t)k * p.a2 + (int64_t)tile * CT, c[k]);
		} else {   // a tile the margins cut: only its own columns
#pragma unroll
			for (int k = 0; k < DH; k++)
#pragma unroll
				for (int q = 0; q < CT; q++)
					if (d0 + k < p.D && tile * CT + q >= jb && tile * CT + q < je) arow[(int64_t)k * p.a2 + (int64_t)tile * CT + q] = c[k][q];
		}
	}
}

// ---- horizontal paths, tiles moved by the TMA engine ---------------------------------------------------------------------------
// The same walk as k_sgm_sweep_h<DH, 8, MODE, false>, but the tile of a warp — DH*32 disparity rows x 8 columns, 32 bytes out of
// every one of DH*32 rows of the volume — is fetched and written back by ONE tensor-map bulk copy each (UTMALDG / UTMASTG in
// SASS) instead of DH 256-bit accesses per lane that each cost 32 look-ups in the L1 tag stage (what bounded the kernel above).
// The volume is described to the TMA unit as a 4-D tensor (column, lane, k, image row) with d = lane*DH + k, so that the box
// lands in shared memory as [k][lane][8 columns]: lane l reads and writes its own 32 bytes per k with two 16-byte accesses.
// Per warp a ring of three tiles: one being worked on (contributions replace the costs in place), one arriving, one leaving;
// a warp and its three mbarriers are self-contained, no block-wide barrier. Requires D == 32*DH, float costs, no left / right
// margin, 16-byte aligned strides; everything else stays with the kernel above. Measured at C2 size (Score, f32): 1.40 ms with
// the 256-bit accesses -> 1.19 ms; loads by TMA but stores by STG.256 (the stage then frees at once, three tiles of lookahead):
// 1.31 ms. What is left is the memory system serving 32-byte pieces of 2 x 2.1 GB at 3.6 TB/s.
constexpr int SW_HT_STAGES = 3;
constexpr int SW_SPIN_LIMIT = 1 << 22;

__device__ __forceinline__ uint32_t hs_smem(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ bool hs_mbar_try(uint64_t* bar, uint32_t parity) {
	uint32_t ok;
	asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
				 : "=r"(ok) : "r"(hs_smem(bar)), "r"(parity) : "memory");
	return ok != 0;
}

template<int DH, int MODE>
__global__ void __launch_bounds__(128, 2) k_sgm_sweep_h_tma(const HSweep p, const __grid_constant__ CUtensorMap tmC, const __grid_constant__ CUtensorMap tmA) {
	constexpr int TILE = 32 * DH * 8;   // floats per tile
	extern __shared__ __align__(1024) float hsm[];
	__shared__ uint64_t bars[4][SW_HT_STAGES];
	const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
	const int i = p.top + blockIdx.x * 4 + w;
	if (i >= p.H - p.bottom) return;   // (the warps of a block never meet at a barrier)
	float* ring = hsm + w * SW_HT_STAGES * TILE;
	uint64_t* bar = bars[w];
	if (lane == 0) {
#pragma unroll
		for (int s = 0; s < SW_HT_STAGES; s++) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(hs_smem(&bar[s])) : "memory");
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	}
	__syncwarp();
	const float WORST = sw_worst<MODE>();
	const int d0 = lane * DH;
	const int n_tiles = p.W / 8;
	auto issue_load = [&](int tile) {   // lane 0
		const int s = tile % SW_HT_STAGES;
		asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(hs_smem(&bar[s])), "r"(TILE * 4) : "memory");
		asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];"
					 ::"r"(hs_smem(ring + s * TILE)), "l"(&tmC), "r"(tile * 8), "r"(0), "r"(0), "r"(i), "r"(hs_smem(&bar[s])) : "memory");
	};
	if (lane == 0) { issue_load(0); if (n_tiles > 1) issue_load(1); }
	float pf[DH];
#pragma unroll
	for (int k = 0; k < DH; k++) pf[k] = 0.f;   // previous_cost starts at 0 (sgm.h:206-208)
	for (int tile = 0; tile < n_tiles; tile++) {
		const int s = tile % SW_HT_STAGES;
		float* T = ring + s * TILE + lane * 8;
		for (int spins = 0; !hs_mbar_try(&bar[s], (tile / SW_HT_STAGES) & 1); spins++)
			if (spins > SW_SPIN_LIMIT) { *p.fault = 1; break; }   // (bounded like the strip hand-over: a flagged wrong result, not a hung device)
		float c[DH][8];
#pragma unroll
		for (int k = 0; k < DH; k++) {
			const float4 x = *reinterpret_cast<const float4*>(T + k * 256), y = *reinterpret_cast<const float4*>(T + k * 256 + 4);
			c[k][0] = x.x; c[k][1] = x.y; c[k][2] = x.z; c[k][3] = x.w; c[k][4] = y.x; c[k][5] = y.y; c[k][6] = y.z; c[k][7] = y.w;
		}
#pragma unroll
		for (int q = 0; q < 8; q++) {
			const int j = tile * 8 + q;
			float mx = WORST;
#pragma unroll
			for (int k = 0; k < DH; k++) mx = sw_better<MODE>(mx, pf[k]);
			mx = sw_warp_best<MODE>(mx);
			float nlo = __shfl_up_sync(0xffffffffu, pf[DH - 1], 1), nhi = __shfl_down_sync(0xffffffffu, pf[0], 1);
			if (lane == 0) nlo = WORST;
			if (lane == 31) nhi = WORST;
			float cc[DH], acc[DH];
#pragma unroll
			for (int k = 0; k < DH; k++) { cc[k] = c[k][q]; acc[k] = 0.f; }
			// j + d >= W (the Pout term) only concerns the last D columns
			if (j + p.D <= p.W) sw_path<DH, MODE, false, false>(cc, pf, nlo, nhi, mx, p.P1, p.P2, p.Pout, 0, acc, false);
			else sw_path<DH, MODE, true, false>(cc, pf, nlo, nhi, mx, p.P1, p.P2, p.Pout, p.W - j - d0, acc, false);
#pragma unroll
			for (int k = 0; k < DH; k++) c[k][q] = acc[k];
		}
#pragma unroll
		for (int k = 0; k < DH; k++) {
			*reinterpret_cast<float4*>(T + k * 256) = make_float4(c[k][0], c[k][1], c[k][2], c[k][3]);
			*reinterpret_cast<float4*>(T + k * 256 + 4) = make_float4(c[k][4], c[k][5], c[k][6], c[k][7]);
		}
		asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // the lanes' writes, before the TMA unit reads the tile
		__syncwarp();
		if (lane == 0) {
			asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.tile.bulk_group [%0, {%1, %2, %3, %4}], [%5];"
						 ::"l"(&tmA), "r"(tile * 8), "r"(0), "r"(0), "r"(i), "r"(hs_smem(ring + s * TILE)) : "memory");
			asm volatile("cp.async.bulk.commit_group;" ::: "memory");
			// tile + 2 goes where tile - 1 was: its store (committed one iteration ago) must have read the tile
			asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
			if (tile + 2 < n_tiles) issue_load(tile + 2);
		}
	}
	if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

// ---- vertical and diagonal paths: sheared strips -----------------------------------------------------------------------------

// columns per strip = computing warps per block (a warp per column): a template parameter, 16 or 8. Eight-column strips (round 2c) make
// twice as many strips, two blocks per SM and half the bytes per block and row: the sweeps are bound by the per-row latency of a
// block (tile movement, barriers), not by arithmetic. The tile pitch is TW + 1 (odd: a warp reading one column over 32 rows touches
// 32 banks); TW / 2 warps move tiles.

struct VSweep {
	const void* C; int64_t c0, c2; int c_u16;
	float* S; int64_t s0, s2;
	const float* A; int64_t a0, a2;        // horizontal contributions (nullptr: none), added after the straight path's
	uint2* xchg; unsigned tag_base; int* claim; int* fault;   // xchg: {value bits, tag} pairs, see k_sgm_sweep_t
	int H, W, D, top, left, nrows, ncols;
	int sweep;                             // +1: rows top -> bottom, -1: bottom -> top
	int rowlive[3], collive[3];            // path k (0: along the shear, 1: straight, 2: against the shear): its lines started by
	                                       // the row-start / the column-start loop are live (Q4)
	int corner_once;                       // CORRECTED: the corner line is started once (COMPAT: twice, Q4)
	int accumulate;                        // 0: S = C + contributions, 1: S += contributions
	int s_lo, s_hi, n_strips;
	int dbg;                               // measurement aid (STEVI_SGM_DBG): 2 = no wait (results invalid)
	float P1, P2, Pout;
};

__device__ __forceinline__ uint4 ld_relaxed_v4(const uint2* p) {
	uint4 v;
	asm volatile("ld.relaxed.gpu.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
	return v;
}
__device__ __forceinline__ uint2 ld_relaxed_v2(const uint2* p) {
	uint2 v;
	asm volatile("ld.relaxed.gpu.global.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p) : "memory");
	return v;
}
__device__ __forceinline__ void cp_async4s(unsigned dst_smem, const void* src) {
	asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst_smem), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
__device__ __forceinline__ void bar_sync(int id, int nthreads) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory"); }

// One path through one pixel, branch-free so that the three paths of a pixel form one block of independent instruction
// streams: st = previous scores in / new scores out, wsel = 1.f for a line that is walked (0.f otherwise: nothing is added and the
// state is of no consequence), twice = the corner line of the path along the shear (Q4).
template<int DN, int MODE, bool POUT, bool TWICE>
__device__ __forceinline__ void sw_path_bf(const float (&c)[DN], float (&st)[DN], float lo, float hi, float M, float P1, float P2, float Pout,
										   int kout, float (&acc)[DN], bool walked, bool twice) {
	const float WORST = sw_worst<MODE>();
	const bool mfin = isfinite(M);   // no finite previous score at all: actual_cost = c (sgm.h:252, :292)
	const float farv = MODE == SW_SCORE ? M - P2 : M + P2;
	float left = lo;
#pragma unroll
	for (int m = 0; m < DN; m++) {
		const float cur = st[m];
		const float right = m < DN - 1 ? st[m < DN - 1 ? m + 1 : 0] : hi;
		float a = sw_cell<MODE>(c[m], cur, left, right, M, farv, P1, Pout, POUT && m >= kout);
		a = mfin ? a : c[m];
		const float contrib = walked ? a - c[m] : 0.f;
		acc[m] += contrib;
		if (TWICE) acc[m] += twice ? contrib : 0.f;
		left = cur;
		st[m] = isfinite(a) ? a : WORST;
	}
}

// The same for the common pixel: every path walked once, a finite extremum, no Pout term in reach (j + D <= W)
template<int DN, int MODE>
__device__ __forceinline__ void sw_path_fast(const float (&c)[DN], float (&st)[DN], float lo, float hi, float M, float P1, float P2, float (&acc)[DN]) {
	const float WORST = sw_worst<MODE>();
	const float farv = MODE == SW_SCORE ? M - P2 : M + P2;
	float left = lo;
#pragma unroll
	for (int m = 0; m < DN; m++) {
		const float cur = st[m];
		const float right = m < DN - 1 ? st[m < DN - 1 ? m + 1 : 0] : hi;
		const float a = sw_cell<MODE>(c[m], cur, left, right, M, farv, P1, 0.f, false);
		acc[m] += a - c[m];
		left = cur;
		st[m] = isfinite(a) ? a : WORST;
	}
}

// One block walks one strip down (or up) the rows. Per row the strip's slab of the volume, [D][16 columns], is brought into
// shared memory with coalesced 64-byte runs (asynchronous copies, one or two rows ahead) and read
// back TRANSPOSED: warp w works on column w of the tile, lane l holding the DH consecutive disparities l*DH .. l*DH+DH-1 of
// each of the three paths through the pixel (previous scores in registers; extremum over d = one warp reduction; the
// neighbours d-1 / d+1 are in registers except at the lane borders: two shuffles). The tile rows are stored permuted
// ((d % DH)*32 + d / DH) with an odd pitch, so that the transposed reads and writes are free of bank conflicts. The aggregated
// costs are staged in the tile of the horizontal contributions (read a, write S in place) and leave with the same 64-byte runs.
// The tile slides one column to the right per row (the shear), so along the shear a path stays in its registers; the straight
// path is found one column to the right in the previous row, the path against the shear two: after a row each warp leaves those
// two vectors in shared memory for the warps on its left — warps 0 and 1 also for the strip on the left, through global memory:
// every float travels as an 8-byte {bits, tag} pair written by one store (tag = launch epoch << 16 | row + 1), so the reader
// needs no fence and no separate flag: it polls the pairs it needs until their tags are this launch's and this row's. The
// reader is a 17th warp that does nothing else: it collects what the strip on the right left for the NEXT row while the 16
// others work on the current one (its trips to L2 are off their critical path) and joins them at the barrier that opens a row.
// DIRS: bit k = path k is part of the direction set of this sweep.
template<int DH, int MODE, int DIRS, int TWT>
__global__ void __launch_bounds__((TWT + 1 + TWT / 2) * 32, TWT == 8 ? 2 : 1) k_sgm_sweep_t(const VSweep p) {
	constexpr int TW = TWT, TWP = TWT + 1, NWARP = TWT, SW_NIO = TWT / 2, DR = 32 * DH, TILE = DR * TWP;
	constexpr bool EN0 = (DIRS & 1) != 0, EN1 = (DIRS & 2) != 0, EN2 = (DIRS & 4) != 0, XCH = EN1 || EN2;
	constexpr int NCT = NWARP * 32, NIOT = SW_NIO * 32, NALL = NCT + 32 + NIOT;   // computing threads, tile movers, everyone
	extern __shared__ float smem[];
	constexpr int NB = DH <= 8 ? 3 : 2;                 // tile buffers per stream: a slab is asked for NB - 1 rows before it is needed
	float* Ct = smem;                                   // [NB][DR][TWP] costs
	float* Xt = Ct + NB * TILE;                         // [NB][DR][TWP] horizontal contributions (if any)
	float* Yt = Xt + (p.A ? NB * TILE : 0);             // [NB][DR][TWP] the volume being added to (second sweep, framed first sweep)
	float* ex1 = Yt + (p.accumulate ? NB * TILE : 0);    // [NWARP][DR] straight path: previous scores of column w, for column w - 1
	float* ex2 = ex1 + NWARP * DR;                      // [NWARP][DR] path against the shear: of column w, for column w - 2
	float* exr = ex2 + NWARP * DR;                      // [2][3][DR] what the strip on the right left (by row parity): straight path of
	                                                    // its column 0, path against the shear of its columns 0 and 1
	float* Ot = p.A ? Xt : Yt;                          // the tile the results are staged in
	__shared__ int s_claim;
	const float WORST = sw_worst<MODE>();
	const float nanv = __int_as_float(0x7fc00000);
	const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
	const int d0 = lane * DH;
	const int dn = min(DH, p.D - d0);                   // disparities of this lane that exist
	for (;;) {
		if (tid == 0) s_claim = atomicAdd(p.claim, 1);
		__syncthreads();
		const int claim = s_claim;
		__syncthreads();
		if (claim >= p.n_strips) return;
		const int s = p.s_hi - claim;   // right to left: whatever this strip waits for has been claimed before
		const int t_begin = max(0, -s * TW - (TW - 1)), t_end = min(p.nrows, p.ncols - s * TW);
		for (int u = tid; u < (2 * NWARP + 6) * DR; u += NALL) ex1[u] = WORST;
		__syncthreads();   // exchange slots initialised
		auto row_of = [&](int t) { return p.sweep > 0 ? p.top + t : p.top + p.nrows - 1 - t; };

		if (w > NWARP) {
			// ================= tile movers: slab of row t+1 in (asynchronous copies), results of row t out =================
			// thread = tile column fcol, disparity rows frow + 16 u. (Row d sits at tile row (d % DH) * 32 + d / DH.)
			constexpr int FR = NIOT / TW;                   // disparity rows the movers cover at a time (16)
			const int it = tid - NCT - 32, fcol = it % TW, frow = it / TW;   // frow 0..FR-1
			constexpr int NU = DR / FR;
			auto tile_row = [&](int u) { return (frow % DH) * 32 + frow / DH + u * (FR / DH); };   // (FR is a multiple of DH)
			const int nu = min(NU, (p.D - frow + FR - 1) / FR);   // disparity rows of this thread that exist
			const int fcc0 = s * TW + fcol;                 // this thread's column in row t: fcc0 + t
			const int esz = p.c_u16 ? 2 : 4;
			const int64_t rs = p.sweep > 0 ? 1 : -1;
			const int64_t stepC = (rs * p.c0 + 1) * esz, stepA = rs * p.a0 + 1, stepS = rs * p.s0 + 1;
			const int64_t dC = FR * p.c2 * esz, dA = FR * p.a2, dS = FR * p.s2;
			// running pointers of the thread's first disparity row at (row t, its column in row t)
			const int64_t row_b = row_of(t_begin), col_b = p.left + fcc0 + t_begin;
			const char* gC = static_cast<const char*>(p.C) + (row_b * p.c0 + (int64_t)frow * p.c2 + col_b) * esz;
			const float* gA = p.A ? p.A + row_b * p.a0 + (int64_t)frow * p.a2 + col_b : nullptr;
			float* gS = p.S + row_b * p.s0 + (int64_t)frow * p.s2 + col_b;   // (next row to fill from, when accumulating)
			float* gF = gS;                                                  // (next row to flush to)
			const unsigned sC = (unsigned)__cvta_generic_to_shared(Ct) + 4u * fcol, sX = (unsigned)__cvta_generic_to_shared(Xt) + 4u * fcol,
						   sY = (unsigned)__cvta_generic_to_shared(Yt) + 4u * fcol;
			auto fill = [&](int t, int b) {   // the slab of row t (the pointers stand on it) -> buffer b
				const int cc = fcc0 + t;
				if (t < t_end && cc >= 0 && cc < p.ncols) {
					const unsigned bo = b * (4u * TILE);
					const char* pc = gC;
					if (p.c_u16) {   // no 2-byte asynchronous copies: all loads first, then the stores
						uint16_t v[NU];
#pragma unroll
						for (int u = 0; u < NU; u++, pc += dC) v[u] = u < nu ? __ldg(reinterpret_cast<const uint16_t*>(pc)) : (uint16_t)0;
#pragma unroll
						for (int u = 0; u < NU; u++) if (u < nu) Ct[b * TILE + tile_row(u) * TWP + fcol] = (float)v[u];
					} else if (nu == NU) {
#pragma unroll
						for (int u = 0; u < NU; u++, pc += dC) cp_async4s(sC + bo + 4u * (tile_row(u) * TWP), pc);
					} else {
#pragma unroll
						for (int u = 0; u < NU; u++, pc += dC) if (u < nu) cp_async4s(sC + bo + 4u * (tile_row(u) * TWP), pc);
					}
					if (p.A) {
						const float* pa = gA;
						if (nu == NU) {
#pragma unroll
							for (int u = 0; u < NU; u++, pa += dA) cp_async4s(sX + bo + 4u * (tile_row(u) * TWP), pa);
						} else {
#pragma unroll
							for (int u = 0; u < NU; u++, pa += dA) if (u < nu) cp_async4s(sX + bo + 4u * (tile_row(u) * TWP), pa);
						}
					}
					if (p.accumulate) {
						const float* ps = gS;
						if (nu == NU) {
#pragma unroll
							for (int u = 0; u < NU; u++, ps += dS) cp_async4s(sY + bo + 4u * (tile_row(u) * TWP), ps);
						} else {
#pragma unroll
							for (int u = 0; u < NU; u++, ps += dS) if (u < nu) cp_async4s(sY + bo + 4u * (tile_row(u) * TWP), ps);
						}
					}
				}
				asm volatile("cp.async.commit_group;" ::: "memory");
				gC += stepC; gA += stepA; gS += stepS;   // (gA is only dereferenced when there is an A)
			};
			auto flush = [&](int t, int b) {   // results of row t, staged in buffer b -> S
				const int cc = fcc0 + t;
				if (t >= t_begin && cc >= 0 && cc < p.ncols) {
					const float* src = Ot + b * TILE + fcol;
					float* ps = gF;
					if (nu == NU) {
#pragma unroll
						for (int u = 0; u < NU; u++, ps += dS) *ps = src[tile_row(u) * TWP];
					} else {
#pragma unroll
						for (int u = 0; u < NU; u++, ps += dS) if (u < nu) *ps = src[tile_row(u) * TWP];
					}
				}
				gF += stepS;
			};
			constexpr int PD = 6;
			auto prefetch = [&](int t) {   // the tile movers share the disparity rows of the slab PD rows ahead: first and last sector
				if (t >= t_end) return;
				const int64_t r = row_of(t);
				const int j0 = p.left + min(max(s * TW + t, 0), p.ncols - 1), j1 = p.left + min(max(s * TW + t + TW - 1, 0), p.ncols - 1);
				for (int d = it; d < p.D; d += NIOT) {
					if (p.c_u16) {
						const uint16_t* q = static_cast<const uint16_t*>(p.C) + r * p.c0 + (int64_t)d * p.c2;
						prefetch_l2(q + j0); prefetch_l2(q + j1);
					} else {
						const float* q = static_cast<const float*>(p.C) + r * p.c0 + (int64_t)d * p.c2;
						prefetch_l2(q + j0); prefetch_l2(q + j1);
					}
					if (p.A) { const float* q = p.A + r * p.a0 + (int64_t)d * p.a2; prefetch_l2(q + j0); prefetch_l2(q + j1); }
					if (p.accumulate) { const float* q = p.S + r * p.s0 + (int64_t)d * p.s2; prefetch_l2(q + j0); prefetch_l2(q + j1); }
				}
			};
			// the L2 prefetch several rows ahead pays for u16 costs only, whose slabs are fetched by ordinary loads (no 2-byte asynchronous
			// copies); for float costs it costs more than it hides (downward sweep at C2 size 4.12 -> 3.83 ms without it, round 2)
			const bool use_pf = p.c_u16 || (p.dbg & 128);
			if (use_pf) for (int u = NB - 1; u < PD; u++) prefetch(t_begin + u);
#pragma unroll
			for (int u = 0; u < NB - 1; u++) fill(t_begin + u, u);
			int b = 0;   // buffer of row t
			for (int t = t_begin; t < t_end; t++, b = b == NB - 1 ? 0 : b + 1) {
				// the slab of row t has landed (the later ones may still be in flight)
				if (NB == 3) asm volatile("cp.async.wait_group 1;" ::: "memory"); else asm volatile("cp.async.wait_group 0;" ::: "memory");
				bar_sync(1, NALL);                                      // row t opens (and the results of row t - 1 are staged)
				if (p.dbg & 4) continue;   // (measurement aid)
				const int bp = b == 0 ? NB - 1 : b - 1;                 // buffer of row t - 1 = buffer of row t + NB - 1
				if (t > t_begin && !(p.dbg & 32)) flush(t - 1, bp);
				if (!(p.dbg & 64)) fill(t + NB - 1, bp);                // into the buffer just flushed
				if (use_pf) prefetch(t + PD);
			}
			asm volatile("cp.async.wait_all;" ::: "memory");
			__syncthreads();   // the last row's results are staged
			flush(t_end - 1, b == 0 ? NB - 1 : b - 1);
			continue;
		}

		const bool has_right = s + 1 <= p.s_hi;
		const int rt_begin = max(0, -(s + 1) * TW - (TW - 1)), rt_end = min(p.nrows, p.ncols - (s + 1) * TW);
		const uint2* xin = p.xchg + (int64_t)(s + 1 - p.s_lo) * p.nrows * 3 * DR + d0;
		uint2* xout = p.xchg + (int64_t)(s - p.s_lo) * p.nrows * 3 * DR + d0;
		if (w == NWARP) {
			// ================= the collecting warp: what the strip on the right left for row t (its row t - 1) -> exr[t & 1] =========
			// one row ahead of the computing warps; joins them at the barrier that opens the row
			for (int t = t_begin; t < t_end; t++) {
				const bool need = XCH && has_right && t >= 1 && t - 1 >= rt_begin && t - 1 < rt_end && !(p.dbg & 8);
				float* dx = exr + (t & 1) * 3 * DR + lane;
				if (!need) {
#pragma unroll
					for (int k = 0; k < 3; k++)
#pragma unroll
						for (int m = 0; m < DH; m++) dx[k * DR + m * 32] = WORST;
				} else {
					const uint2* src = xin + (int64_t)(t - 1) * 3 * DR;
					const unsigned tag = p.tag_base | (unsigned)t;   // what the neighbour wrote for its row t - 1
					for (int spins = 0;; spins++) {
						bool ok = true;
						float v3[3][DH];
#pragma unroll
						for (int k = 0; k < 3; k++) {
							if (!(k == 0 ? EN1 : EN2)) {   // never published
#pragma unroll
								for (int m = 0; m < DH; m++) v3[k][m] = WORST;
								continue;
							}
							if (DH == 1) {
								const uint2 v = ld_relaxed_v2(src + k * DR);
								ok = ok && v.y == tag;
								v3[k][0] = __uint_as_float(v.x);
							} else {
#pragma unroll
								for (int m = 0; m < DH; m += 2) {
									const uint4 v = ld_relaxed_v4(src + k * DR + m);
									ok = ok && v.y == tag && v.w == tag;
									v3[k][m] = __uint_as_float(v.x);
									v3[k][m + 1 < DH ? m + 1 : 0] = __uint_as_float(v.z);
								}
							}
						}
						const bool all_ok = __all_sync(0xffffffffu, ok) || (p.dbg & 2);
						if (!all_ok && spins > SW_SPIN_LIMIT) *p.fault = 1;   // raise the fault flag instead of hanging
						if (all_ok || spins > SW_SPIN_LIMIT) {
#pragma unroll
							for (int k = 0; k < 3; k++)
#pragma unroll
								for (int m = 0; m < DH; m++) dx[k * DR + m * 32] = v3[k][m];
							break;
						}
						__nanosleep(100);
					}
				}
				bar_sync(1, NALL);
			}
			__syncthreads();
			continue;
		}

		// ================= computing warps: warp w = column w of the tile ============================================================
		auto publish = [&](int t, int slot, const float (&v)[DH]) {   // one vector for the strip on the left
			uint2* dst = xout + ((int64_t)t * 3 + slot) * DR;
			const unsigned tag = p.tag_base | (unsigned)(t + 1);
			if (DH == 1) __stcg(dst, make_uint2(__float_as_uint(v[0]), tag));
			else {
#pragma unroll
				for (int m = 0; m < DH; m += 2)
					__stcg(reinterpret_cast<uint4*>(dst + m), make_uint4(__float_as_uint(v[m]), tag, __float_as_uint(v[m + 1 < DH ? m + 1 : 0]), tag));
			}
		};
		float st[3][DH];   // previous scores of the three paths through the warp's column
#pragma unroll
		for (int k = 0; k < 3; k++)
#pragma unroll
			for (int m = 0; m < DH; m++) st[k][m] = WORST;
		// the line weight of the path along the shear is the same for the whole strip and column (Q4): row starts where the column
		// offset is <= 0, column starts where it is >= 0 (both for the corner line)
		const int sx = s * TW + w;
		int wt0 = ((sx <= 0 && p.rowlive[0]) ? 1 : 0) + ((sx >= 0 && p.collive[0]) ? 1 : 0);
		if (wt0 == 2 && p.corner_once) wt0 = 1;
		// where the vectors from the right are found: the next warps' slots, or (last warps) what the collecting warp brought
		const float* in1 = (w + 1 < NWARP ? ex1 + (w + 1) * DR : exr) + lane;
		const float* in2 = (w + 2 < NWARP ? ex2 + (w + 2) * DR : exr + (1 + (w + 2 - NWARP)) * DR) + lane;
		const int par1 = w + 1 < NWARP ? 0 : 3 * DR, par2 = w + 2 < NWARP ? 0 : 3 * DR;   // offset of the odd rows' buffer
		int bcur = 0;   // tile buffer of row t
		for (int t = t_begin; t < t_end; t++, bcur = bcur == NB - 1 ? 0 : bcur + 1) {
			bar_sync(1, NALL);   // the slab of row t and the vectors from the right are in place; everyone is done with row t - 1
			const int tb = bcur * TILE + lane * TWP + w;
			const int cc = sx + t;   // column inside the margin rectangle
			if (cc >= 0 && cc < p.ncols) {   // (warp-uniform)
				const int j = p.left + cc;
				float c[DH], acc[DH];
#pragma unroll
				for (int m = 0; m < DH; m++) c[m] = m < dn ? Ct[tb + m * 32 * TWP] : nanv;
				if (p.accumulate) {
#pragma unroll
					for (int m = 0; m < DH; m++) acc[m] = Yt[tb + m * 32 * TWP];
				} else {
#pragma unroll
					for (int m = 0; m < DH; m++) acc[m] = c[m];   // sgm.h:371-377
				}
				// previous scores: where each line was one row ago
				if (EN1) {
					const float* q = in1 + (t & 1) * par1;
#pragma unroll
					for (int m = 0; m < DH; m++) st[1][m] = q[m * 32];
				}
				if (EN2) {
					const float* q = in2 + (t & 1) * par2;
#pragma unroll
					for (int m = 0; m < DH; m++) st[2][m] = q[m * 32];
				}
				// a path whose previous pixel lies outside the margin rectangle starts here: previous_cost = 0 (sgm.h:206-208)
				if (t == 0 || cc == 0 || cc == p.ncols - 1) {   // (rare)
					const bool fresh[3] = {t == 0 || cc == 0, t == 0, t == 0 || cc == p.ncols - 1};
#pragma unroll
					for (int k = 0; k < 3; k++) {
						if (!((DIRS >> k) & 1) || !fresh[k]) continue;
#pragma unroll
						for (int m = 0; m < DH; m++) st[k][m] = m < dn ? 0.f : WORST;
					}
				}
				// per path: extremum over d (the three reductions run side by side), neighbours across the lane borders
				float M[3], lo[3], hi[3];
#pragma unroll
				for (int k = 0; k < 3; k++) {
					M[k] = WORST;
					if (!((DIRS >> k) & 1)) continue;
#pragma unroll
					for (int m = 0; m < DH; m++) M[k] = sw_better<MODE>(M[k], st[k][m]);
				}
#pragma unroll
				for (int k = 0; k < 3; k++)
					if ((DIRS >> k) & 1) M[k] = sw_warp_best<MODE>(M[k]);
#pragma unroll
				for (int k = 0; k < 3; k++) {
					if (!((DIRS >> k) & 1)) continue;
					lo[k] = __shfl_up_sync(0xffffffffu, st[k][DH - 1], 1);
					hi[k] = __shfl_down_sync(0xffffffffu, st[k][0], 1);
					if (lane == 0) lo[k] = WORST;
					if (lane == 31) hi[k] = WORST;
				}
				// line weights (Q4): how many times the reference walks the line through this pixel on each path
				const int kr2 = p.ncols - 1 - cc;
				int wt2 = ((kr2 <= t && p.rowlive[2]) ? 1 : 0) + ((t <= kr2 && p.collive[2]) ? 1 : 0);
				if (wt2 == 2 && p.corner_once) wt2 = 1;
				const bool walked[3] = {wt0 > 0, p.collive[1] != 0, wt2 > 0};
				const int kout = p.W - j - d0;   // j + d >= W  <=>  m >= kout
				bool plain = j + p.D <= p.W && wt0 != 2;
#pragma unroll
				for (int k = 0; k < 3; k++)
					if ((DIRS >> k) & 1) plain = plain && walked[k] && isfinite(M[k]);
				// the reference's order of additions: Up2Down, Left2Right, UpLeft2DownRight, UpRight2DownLeft (sgm.h:379-388)
				if (plain) {   // (warp-uniform)
					if (EN1) sw_path_fast<DH, MODE>(c, st[1], lo[1], hi[1], M[1], p.P1, p.P2, acc);
					if (p.A) {
#pragma unroll
						for (int m = 0; m < DH; m++) acc[m] += Xt[tb + m * 32 * TWP];
					}
					if (EN0) sw_path_fast<DH, MODE>(c, st[0], lo[0], hi[0], M[0], p.P1, p.P2, acc);
					if (EN2) sw_path_fast<DH, MODE>(c, st[2], lo[2], hi[2], M[2], p.P1, p.P2, acc);
				} else {
					if (EN1) sw_path_bf<DH, MODE, true, false>(c, st[1], lo[1], hi[1], M[1], p.P1, p.P2, p.Pout, kout, acc, walked[1], false);
					if (p.A) {
#pragma unroll
						for (int m = 0; m < DH; m++) acc[m] += Xt[tb + m * 32 * TWP];
					}
					if (EN0) sw_path_bf<DH, MODE, true, true>(c, st[0], lo[0], hi[0], M[0], p.P1, p.P2, p.Pout, kout, acc, walked[0], wt0 == 2);
					if (EN2) sw_path_bf<DH, MODE, true, false>(c, st[2], lo[2], hi[2], M[2], p.P1, p.P2, p.Pout, kout, acc, walked[2], false);
				}
#pragma unroll
				for (int m = 0; m < DH; m++) if (m < dn) Ot[tb + m * 32 * TWP] = acc[m];
			}
			// hand-over to the strip on the left (global memory)
			if (XCH && s > p.s_lo && !(p.dbg & 16)) {
				if (w == 0) { if (EN1) publish(t, 0, st[1]); if (EN2) publish(t, 1, st[2]); }
				if (w == 1 && EN2) publish(t, 2, st[2]);
			}
			bar_sync(2, NCT);   // everyone has read this row's exchange vectors
			if (XCH) {   // vectors for the warps on the left (next row)
#pragma unroll
				for (int m = 0; m < DH; m++) {
					if (EN1) ex1[w * DR + m * 32 + lane] = st[1][m];
					if (EN2) ex2[w * DR + m * 32 + lane] = st[2][m];
				}
			}
		}
		__syncthreads();
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
#define STEVI_H(DH) do { if (vec) STEVI_LAUNCH(c, (k_sgm_sweep_h<DH, 8, MODE, REVERSE>), grid, 128, 0, p); \
	else STEVI_LAUNCH(c, (k_sgm_sweep_h<DH, 1, MODE, REVERSE>), grid, 128, 0, p); } while (0)
	if (dh <= 1) STEVI_H(1); else if (dh <= 2) STEVI_H(2); else if (dh <= 4) STEVI_H(4); else if (dh <= 8) STEVI_H(8); else STEVI_H(16);
#undef STEVI_H
	return STEVI_OK;
}

// tensor map of a volume with unit column stride for the tiles of k_sgm_sweep_h_tma: dims (column, lane, k, image row), d = lane*dh + k
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
								  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled_fn() {
	static EncodeTiledFn fn = [] {
		void* f = nullptr;
		cudaDriverEntryPointQueryResult qr;
		if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &qr) != cudaSuccess || qr != cudaDriverEntryPointSuccess) f = nullptr;
		(void)cudaGetLastError();
		return reinterpret_cast<EncodeTiledFn>(f);
	}();
	return fn;
}
static bool h_tile_map(CUtensorMap* tm, const float* base, int64_t row_stride, int64_t d_stride, int H, int W, int dh) {
	EncodeTiledFn fn = encode_tiled_fn();
	if (!fn) return false;
	const cuuint64_t dims[4] = {(cuuint64_t)W, 32, (cuuint64_t)dh, (cuuint64_t)H};
	const cuuint64_t strides[3] = {(cuuint64_t)dh * d_stride * 4, (cuuint64_t)d_stride * 4, (cuuint64_t)row_stride * 4};
	const cuuint32_t box[4] = {8, 32, (cuuint32_t)dh, 1}, es[4] = {1, 1, 1, 1};
	// (L2 promotion 64 / 128 / 256 bytes measured: 1.40 / 1.19 / 1.35 ms against 1.21 without, at C2 size: left off)
	return fn(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(base), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
			  CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

template<int MODE>
static int32_t launch_h_tma(stevi_ctx* c, const HSweep& p, const CUtensorMap& tmC, const CUtensorMap& tmA) {
	const int rows = p.H - p.top - p.bottom;
	dim3 grid(cdiv(rows, 4));
	const int dh = p.D / 32;
	const int smem = 4 * SW_HT_STAGES * 32 * dh * 8 * (int)sizeof(float);
#define STEVI_HT(DH) do { \
	STEVI_CUDA(c, cudaFuncSetAttribute((k_sgm_sweep_h_tma<DH, MODE>), cudaFuncAttributeMaxDynamicSharedMemorySize, smem)); \
	STEVI_LAUNCH(c, (k_sgm_sweep_h_tma<DH, MODE>), grid, 128, smem, p, tmC, tmA); } while (0)
	switch (dh) {
	case 1: STEVI_HT(1); break;
	case 2: STEVI_HT(2); break;
	case 4: STEVI_HT(4); break;
	default: STEVI_HT(8); break;
	}
#undef STEVI_HT
	return STEVI_OK;
}

static size_t sweep_smem(int dh, bool has_a, bool accumulate, int tw) {
	const size_t tile = (size_t)32 * dh * (tw + 1);
	return ((dh <= 8 ? 3 : 2) * tile * (1 + (has_a ? 1 : 0) + (accumulate ? 1 : 0)) + (size_t)(2 * tw + 6) * 32 * dh) * sizeof(float);
}

template<int MODE, int DIRS, int TWT>
static int32_t launch_v(stevi_ctx* c, const VSweep& p) {
	const int dh = p.D <= 32 ? 1 : p.D <= 64 ? 2 : p.D <= 128 ? 4 : p.D <= 256 ? 8 : 16;
	const size_t smem = sweep_smem(dh, p.A != nullptr, p.accumulate != 0, TWT);
	if (smem > (227u << 10)) return fail(c, STEVI_E_UNSUPPORTED, "sgm: the sweep tiles do not fit in shared memory");
	// persistent blocks: as many as are resident at once (a strip waits for the strip on its right, which must be running)
	const int want_per_sm = TWT == 8 ? 2 : 1;
	const int per_sm = std::max(1, std::min(want_per_sm, (int)((227u << 10) / (smem + 1024))));
	dim3 grid(std::min(p.n_strips, per_sm * c->sm_count));
#define STEVI_T(DH) do { \
	STEVI_CUDA(c, cudaFuncSetAttribute((k_sgm_sweep_t<DH, MODE, DIRS, TWT>), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
	STEVI_LAUNCH(c, (k_sgm_sweep_t<DH, MODE, DIRS, TWT>), grid, (TWT + 1 + TWT / 2) * 32, smem, p); } while (0)
	switch (dh) {
	case 1: STEVI_T(1); break;
	case 2: STEVI_T(2); break;
	case 4: STEVI_T(4); break;
	case 8: STEVI_T(8); break;
	default: STEVI_T(16); break;
	}
#undef STEVI_T
	return STEVI_OK;
}

template<int MODE>
static int32_t launch_v_dirs(stevi_ctx* c, const VSweep& p, int dirs, int tw) {
	if (tw == 8) {
		switch (dirs) {
		case 7: return launch_v<MODE, 7, 8>(c, p);
		case 2: return launch_v<MODE, 2, 8>(c, p);
		case 1: return launch_v<MODE, 1, 8>(c, p);
		}
	} else {
		switch (dirs) {
		case 7: return launch_v<MODE, 7, 16>(c, p);
		case 2: return launch_v<MODE, 2, 16>(c, p);
		case 1: return launch_v<MODE, 1, 16>(c, p);
		}
	}
	return fail(c, STEVI_E_UNSUPPORTED, "sgm: unexpected direction set of a sweep");
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
	const int W4 = (W + 7) / 8 * 8;   // rows of A start on 32-byte sectors
	float* A = nullptr;
	STEVI_TRY(arena_alloc_t(c, (size_t)H * D * W4, &A));
	HSweep h;
	h.C = cv.p; h.c0 = cv.st[0]; h.c2 = cv.st[2]; h.c_u16 = cv.dtype == STEVI_U16;
	h.A = A; h.a0 = (int64_t)D * W4; h.a2 = W4;
	h.H = H; h.W = W; h.D = D; h.top = mg.top; h.bottom = mg.bottom; h.left = mg.left; h.right = mg.right;
	h.P1 = P1; h.P2 = P2; h.Pout = Pout;
	{ static const char* e = getenv("STEVI_SGM_HPF"); h.pf = e ? atoi(e) : 0; }
	h.fault = c->fault_dev;
	const size_t esz = dtype_size(cv.dtype);
	const bool vec = (uintptr_t)cv.p % (8 * esz) == 0 && cv.st[0] % 8 == 0 && cv.st[2] % 8 == 0 && W % 8 == 0;
	// the forward walk with TMA-moved tiles where the volume allows it (see k_sgm_sweep_h_tma)
	static const bool no_tma = getenv("STEVI_SGM_NO_TMA") != nullptr;   // measurement aid
	CUtensorMap tmC, tmA;
	const bool tma = !no_tma && vec && !h.c_u16 && (D == 32 || D == 64 || D == 128 || D == 256) && mg.left == 0 && mg.right == 0 &&
					 (uintptr_t)cv.p % 16 == 0 && h_tile_map(&tmC, cv.as<float>(), cv.st[0], cv.st[2], H, W, D / 32) &&
					 h_tile_map(&tmA, A, h.a0, h.a2, H, W4, D / 32);
	if (tma) {
		if (mode == SW_SCORE) STEVI_TRY((launch_h_tma<SW_SCORE>(c, h, tmC, tmA)));
		else if (mode == SW_COST) STEVI_TRY((launch_h_tma<SW_COST>(c, h, tmC, tmA)));
		else STEVI_TRY((launch_h_tma<SW_COST_COMPAT>(c, h, tmC, tmA)));
	}
	else if (mode == SW_SCORE) STEVI_TRY((launch_h<SW_SCORE, false>(c, h, vec)));
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
	// strip width per sweep: 8 columns when nothing is exchanged between strips (the along-the-shear path alone: the upward sweep of
	// COMPAT), 16 otherwise (measured at C2 size: upward sweep 1.82 -> 1.29 ms with 8; downward sweep 3.80 -> 5.46 ms: the hand-over
	// chain is as long as the strips are many). The exchange buffer is sized for 16-column strips: only they exchange.
	static const int tw_env = [] { const char* e = getenv("STEVI_SGM_TW"); return e ? atoi(e) : 0; }();   // measurement aid: force 8 or 16
	const int cw = tw_env == 8 ? 8 : 16;   // (the width the exchange buffer is sized for)
	const int dr = 32 * (D <= 32 ? 1 : D <= 64 ? 2 : D <= 128 ? 4 : D <= 256 ? 8 : 16);   // disparity rows of a tile
	v.P1 = P1; v.P2 = P2; v.Pout = Pout;
	v.corner_once = corrected ? 1 : 0;
	v.fault = c->fault_dev;
	{ static const char* e = getenv("STEVI_SGM_DBG"); v.dbg = e ? atoi(e) : 0; }
	// exchange buffer between neighbouring strips: persistent in the ctx (only ever holds this kernel's tagged pairs)
	const size_t xbytes = (size_t)((ncols - 1) / cw + (nrows - 1 + cw - 1) / cw + 1) * nrows * 3 * dr * sizeof(uint2);
	if (c->sgm_xchg_bytes < xbytes) {
		STEVI_CUDA(c, cudaStreamSynchronize(c->stream));
		if (c->sgm_xchg) cudaFree(c->sgm_xchg);
		c->sgm_xchg = nullptr; c->sgm_xchg_bytes = 0;
		STEVI_CUDA(c, cudaMalloc(&c->sgm_xchg, xbytes));
		c->sgm_xchg_bytes = xbytes;
		c->sgm_epoch = 0;
	}
	v.xchg = static_cast<uint2*>(c->sgm_xchg);
	int* claims = nullptr;
	STEVI_TRY(arena_alloc_t(c, 2, &claims));
	STEVI_CUDA(c, cudaMemsetAsync(claims, 0, 2 * sizeof(int), c->stream));
	for (int pass = 0; pass < 2; pass++) {
		v.sweep = pass == 0 ? 1 : -1;
		v.A = pass == 0 ? A : nullptr;
		v.accumulate = (pass == 1 || framed) ? 1 : 0;
		v.claim = claims + pass;
		bool any_col = false;
		int dirs = 0;
		for (int k = 0; k < 3; k++) {
			const int sh = 1 - k;
			const bool in_set = ndir == 8 || sh == 0;
			// sgm.h:162-184 with :331,:345: ZeroPos starts are live; EndPos starts (negative steps) only with the corrected rule
			v.rowlive[k] = in_set && sh != 0 && (sh > 0 || corrected);
			v.collive[k] = in_set && (v.sweep > 0 || corrected);
			if (v.rowlive[k] || v.collive[k]) dirs |= 1 << k;
			any_col |= v.collive[k] != 0;
		}
		if (!dirs) continue;
		const int tw = tw_env == 8 || tw_env == 16 ? tw_env : (dirs == 1 ? 8 : 16);   // (4-column strips: 2.66 ms for the upward sweep, half-sector accesses)
		v.s_lo = -((nrows - 1 + tw - 1) / tw);
		const int s_all = (ncols - 1) / tw;
		// strips that hold live pixels: lines started by the column loop reach everywhere, lines started on the left column
		// (the only row starts of a sweep without live column starts) stay at or left of the diagonal through the corner
		v.s_hi = any_col ? s_all : 0;
		v.n_strips = v.s_hi - v.s_lo + 1;
		// a tag no earlier launch on this buffer has used (zeroed buffer = tag 0 = never valid)
		if (c->sgm_epoch == 0 || c->sgm_epoch >= 0xffffu) {
			STEVI_CUDA(c, cudaMemsetAsync(c->sgm_xchg, 0, c->sgm_xchg_bytes, c->stream));
			c->sgm_epoch = 0;
		}
		v.tag_base = (unsigned)(++c->sgm_epoch) << 16;
		if (mode == SW_SCORE) STEVI_TRY(launch_v_dirs<SW_SCORE>(c, v, dirs, tw));
		else if (mode == SW_COST) STEVI_TRY(launch_v_dirs<SW_COST>(c, v, dirs, tw));
		else STEVI_TRY(launch_v_dirs<SW_COST_COMPAT>(c, v, dirs, tw));
	}
	return STEVI_OK;
}

} // namespace stevi
