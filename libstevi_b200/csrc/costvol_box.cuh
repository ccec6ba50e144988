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
// The block's band of image rows (R + 2vr rows of 136 source and 168 target columns) is brought into shared memory by
// the TMA engine (cp.async.bulk, one mbarrier) from zero-padded, unit-de-interleaved float copies of the images, and
// read back with conflict-free 16-byte LDS; the per-pixel statistics rows follow through a 4-slot shared-memory ring
// filled the same way three rows ahead (as plain LDGs they queued behind the stores: 0.46 ms vs 0.31 ms without them);
// stores are 2 x 16 bytes per (row, d) and thread — a half-warp writes 512 contiguous bytes of the reference's
// [H][D][W] layout.
#pragma once
#include "common.cuh"

namespace stevi {

enum { BOX_PROD = 0, BOX_SQ = 1, BOX_ABS = 2 };
// a block = 16 column groups (128 source columns) x 8 disparity groups (32 disparities); three blocks per SM leave
// 168 registers per thread, enough for the 64 window sums plus the row being consumed without spilling
constexpr int BOX_THREADS = 128, BOX_BLOCKS_PER_SM = 3, BOX_D_PER_BLOCK = 32;
// rows per kernel band: fixed, so that band boundaries (and with them the rounding of the running sums) depend only on the
// global row index; (2vr+1) + 2*31 row updates per 32 output rows = 1.1x the steady-state work at vr = 4
constexpr int BOX_BAND_ROWS = 32;

// Geometry of the padded planes (images and statistics maps share it). A padded row holds `pitch` floats (a multiple of
// 64); padded column X = pad_left + j. The row is stored DE-INTERLEAVED by 16-byte units: unit u = X/4 lives at
// (u&1)*pitch/2 + (u>>1)*4. Threads of a warp own 8 consecutive columns each, i.e. their windows start two units apart:
// with this layout the k-th 16-byte load of all lanes is one contiguous run (one wavefront per 128 bytes) instead of
// 16-byte pieces at a 32-byte stride (two).
__host__ __device__ __forceinline__ int pad_unit_off(int u, int half) { return (u & 1) * half + (u >> 1) * 4; }
__host__ __device__ __forceinline__ int pad_col_off(int X, int half) { return pad_unit_off(X >> 2, half) + (X & 3); }

struct BoxParams {
	const float* ps; const float* pt;   // padded image planes, pointing at padded column 0 of image row 0
	const float* As; const float* Bs;   // source window statistics, same geometry
	const float* At; const float* Bt;   // target window statistics
	int64_t pitch;
	int half, pad_left;
	int H, Ws, D, vr, R, row_shift;
	int d_first = 0;                    // disparities below this one are evaluated but not stored (2-D search: the target planes can only be shifted by whole pairs of units)
	float Ff;                           // number of features F as float
	float* out; int64_t o0, o1, o2;
	int vec_store;                      // o1 == 1 and rows aligned to 16 bytes (1) or to 32 bytes (2)
	// optional fused winner-take-all (extractSelectedIndex, correlation_base.h:427-464): dense [H][Ws] keys, cleared to 0
	// beforehand; key = (order-preserving bits of the score, negated for costs) << 32 | d, merged with atomicMax, so equal
	// values resolve to the larger d = the reference's `>=` / `<=` (last extremum wins, Q5); a NaN at d = 0 "sticks"
	// (:440-452: no later comparison is true) and is encoded as the largest key
	unsigned long long* keys;
};

__device__ __forceinline__ unsigned wta_order_bits(float v) {
	const unsigned u = __float_as_uint(v);
	return (u >> 31) ? ~u : (u | 0x80000000u);
}

// NU consecutive units starting at a unit whose offset is offA (the next one is at offB, then offA+4, offB+4, ...)
template<int NU> __device__ __forceinline__ void load_units(float (&dst)[4 * NU], const float* __restrict__ row, int offA, int offB) {
#pragma unroll
	for (int k = 0; k < NU; k++) {
		const float4 v = __ldg(reinterpret_cast<const float4*>(row + ((k & 1) ? offB : offA) + 4 * (k >> 1)));
		dst[4 * k] = v.x; dst[4 * k + 1] = v.y; dst[4 * k + 2] = v.z; dst[4 * k + 3] = v.w;
	}
}

__device__ __forceinline__ float4 load_unit(const float* __restrict__ row, int off) {
	return __ldg(reinterpret_cast<const float4*>(row + off));
}

// ---- shared-memory band tile, filled by the TMA engine (1-D bulk copies completing on an mbarrier) -------------------
// per tile row: [s: 17 units of the first parity | 17 of the other | t: 21 units of the first parity | 21 of the other]
constexpr int BOX_S_UNITS = 17, BOX_T_UNITS = 21;
constexpr int BOX_ROW_FLOATS = 4 * (2 * BOX_S_UNITS + 2 * BOX_T_UNITS);   // 304 floats = 1216 bytes
// ring of per-pixel statistics rows: [As: 16 units | 16][Bs: 16 | 16][At: 20 | 20][Bt: 20 | 20] (parity halves)
constexpr int BOX_SS_UNITS = 16, BOX_ST_UNITS = 20, BOX_STAT_SLOTS = 4;
constexpr int BOX_STAT_FLOATS = 4 * (4 * BOX_SS_UNITS + 4 * BOX_ST_UNITS);  // 576 floats = 2304 bytes
constexpr int BOX_MAX_TILE_ROWS = 54;   // 3 blocks/SM x (54 x 1216 + 4 x 2304 + static) B = 219 KB of the 227 KB
constexpr int BOX_SMEM_MAX = BOX_MAX_TILE_ROWS * BOX_ROW_FLOATS * 4 + BOX_STAT_SLOTS * BOX_STAT_FLOATS * 4;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
	asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
	asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
	asm volatile("{\n\t.reg .pred p;\n\tBOX_WAIT:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra BOX_DONE;\n\tbra BOX_WAIT;\n\tBOX_DONE:\n\t}"
				 :: "r"(smem_u32(bar)), "r"(parity) : "memory");
}
// global -> shared bulk copy (UBLKCP): 16-byte aligned on both sides, size a multiple of 16
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
	asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
				 :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ float4 lds_unit(const float* __restrict__ row, int off) {
	return *reinterpret_cast<const float4*>(row + off);
}

// V[q][c] += / -= op(s(c), t(c + SIGN*(d0+q))) for one tile row. The row is consumed unit by unit (4 source columns
// against a sliding pair of target units) so that few loaded values are live at a time.
template<int OP, int NV, int SIGN, bool ADD>
__device__ __forceinline__ void box_accum_row(float (&V)[4][NV], const float* __restrict__ row, int sA, int sB, int tA, int tB) {
	float4 t0 = lds_unit(row, tA);
#pragma unroll
	for (int k = 0; k < NV / 4; k++) {
		const float4 sv = lds_unit(row, ((k & 1) ? sB : sA) + 4 * (k >> 1));
		const float4 t1 = lds_unit(row, (((k + 1) & 1) ? tB : tA) + 4 * ((k + 1) >> 1));
		const float s[4] = {sv.x, sv.y, sv.z, sv.w};
		const float t[8] = {t0.x, t0.y, t0.z, t0.w, t1.x, t1.y, t1.z, t1.w};
#pragma unroll
		for (int cc = 0; cc < 4; cc++) {
#pragma unroll
			for (int q = 0; q < 4; q++) {
				const float tv = SIGN > 0 ? t[cc + q] : t[cc + 4 - q];
				float& v = V[q][4 * k + cc];
				if (OP == BOX_PROD) {
					v = fmaf(ADD ? s[cc] : -s[cc], tv, v);
				} else if (OP == BOX_SQ) {
					const float e = s[cc] - tv;
					v = fmaf(ADD ? e : -e, e, v);
				} else {
					const float e = fabsf(s[cc] - tv);
					v = ADD ? v + e : v - e;
				}
			}
		}
		t0 = t1;
	}
}

// winner keys of one row, [d group][column slot]: slot = k * BOX_KEY_PITCH + cg for column jb + 8*cg + k, so that both the
// per-thread writes (fixed k, consecutive cg) and the per-column reads (consecutive columns) are bank-conflict free
constexpr int BOX_KEY_PITCH = 18;
constexpr int BOX_KEY_ROW = 8 * BOX_KEY_PITCH;

template<int OP, int HR, int SIGN, bool WTA>
__global__ void __launch_bounds__(BOX_THREADS, BOX_BLOCKS_PER_SM) k_cv_box(const BoxParams p) {
	static_assert(HR >= 1 && HR <= 4, "window half-width");
	constexpr int NV = 16;                  // V columns j0-4 .. j0+11
	constexpr int LH = 4;                   // left halo
	extern __shared__ __align__(128) float smem[];
	__shared__ uint64_t bar_img, bar_st[BOX_STAT_SLOTS];
	__shared__ unsigned long long kb[WTA ? 2 * 8 * BOX_KEY_ROW : 1];   // double-buffered: row i is merged while row i+1 is computed
	float* stat_ring = smem;                                  // BOX_STAT_SLOTS rows of per-pixel statistics
	float* tile = smem + BOX_STAT_SLOTS * BOX_STAT_FLOATS;    // the band of image rows
	const int cg = threadIdx.x & 15, dg = threadIdx.x >> 4;
	const int jb = blockIdx.x * 128, db = blockIdx.y * BOX_D_PER_BLOCK;
	const int j0 = jb + cg * 8;
	const int d0 = db + dg * 4;
	// kernel row bands are anchored at GLOBAL multiples of R (row_shift = global row of image row 0, mod R): the running
	// sums of a row then round the same way whether the caller passes the whole frame or a row band of it
	const int i0 = max((int)blockIdx.z * p.R - p.row_shift, 0);
	const int i_end = min((int)(blockIdx.z + 1) * p.R - p.row_shift, p.H);
	if (i0 >= i_end) return;
	const int n_rows = p.R + 2 * p.vr;      // tile row r <-> image row i0 - vr + r
	// first unit of the block's source window, and of its target window (lowest column any of its threads reads)
	const int us_blk = (p.pad_left + jb - LH) >> 2;
	const int ut_blk = (p.pad_left + (SIGN > 0 ? (jb - LH + db) : (jb - LH - db - BOX_D_PER_BLOCK))) >> 2;
	if (threadIdx.x == 0) {
		mbar_init(&bar_img, 1);
#pragma unroll
		for (int k = 0; k < BOX_STAT_SLOTS; k++) mbar_init(&bar_st[k], 1);
		asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	}
	__syncthreads();
	// one row of statistics -> ring slot: 8 bulk copies (two parity halves of As, Bs, At, Bt), one per lane of warp 0.
	// source statistics: 32 units from column jb; target statistics: 40 units from the unit after the target window's first
	auto issue_stats = [&](int row, int slot) {
		const int lane = threadIdx.x;
		if (lane == 0) mbar_expect_tx(&bar_st[slot], BOX_STAT_FLOATS * 4);
		__syncwarp();
		if (lane < 8) {
			const int plane = lane >> 1, hf = lane & 1;
			const float* src = plane == 0 ? p.As : plane == 1 ? p.Bs : plane == 2 ? p.At : p.Bt;
			const int units = plane < 2 ? BOX_SS_UNITS : BOX_ST_UNITS;
			const int u = (plane < 2 ? us_blk : ut_blk) + 1 + hf;
			const int dst = (plane == 0 ? 0 : plane == 1 ? 8 * BOX_SS_UNITS : plane == 2 ? 16 * BOX_SS_UNITS : 16 * BOX_SS_UNITS + 8 * BOX_ST_UNITS)
							+ hf * 4 * units;
			bulk_g2s(stat_ring + slot * BOX_STAT_FLOATS + dst, src + (int64_t)row * p.pitch + pad_unit_off(u, p.half), units * 16, &bar_st[slot]);
		}
	};
	if (threadIdx.x < 32) {
		if (threadIdx.x == 0) mbar_expect_tx(&bar_img, (uint32_t)n_rows * BOX_ROW_FLOATS * 4);
		__syncwarp();
		const int gsA = pad_unit_off(us_blk, p.half), gsB = pad_unit_off(us_blk + 1, p.half);
		const int gtA = pad_unit_off(ut_blk, p.half), gtB = pad_unit_off(ut_blk + 1, p.half);
		for (int r = threadIdx.x; r < n_rows; r += 32) {
			const int64_t go = (int64_t)(i0 - p.vr + r) * p.pitch;
			float* dst = tile + r * BOX_ROW_FLOATS;
			bulk_g2s(dst, p.ps + go + gsA, BOX_S_UNITS * 16, &bar_img);
			bulk_g2s(dst + 4 * BOX_S_UNITS, p.ps + go + gsB, BOX_S_UNITS * 16, &bar_img);
			bulk_g2s(dst + 8 * BOX_S_UNITS, p.pt + go + gtA, BOX_T_UNITS * 16, &bar_img);
			bulk_g2s(dst + 8 * BOX_S_UNITS + 4 * BOX_T_UNITS, p.pt + go + gtB, BOX_T_UNITS * 16, &bar_img);
		}
		for (int k = 0; k < BOX_STAT_SLOTS - 1; k++)
			if (i0 + k < i_end) issue_stats(i0 + k, k);
	}
	// V column c <-> source column j0-LH+c; target column = source column + SIGN*(d0+q):
	//   SIGN>0: first target column j0-LH+d0, register c+q ; SIGN<0: first target column j0-LH-d0-4, register c+4-q
	// offsets inside a tile row: unit x of the block window sits in parity half (x&1) at index x>>1
	const int sA = 4 * cg, sB = 4 * (BOX_S_UNITS + cg);
	const int xt = 2 * cg + (SIGN > 0 ? dg : (BOX_D_PER_BLOCK / 4 - 1 - dg));
	const int tbase = 8 * BOX_S_UNITS;
	const int tA = tbase + ((xt & 1) ? 4 * (BOX_T_UNITS + (xt >> 1)) : 4 * (xt >> 1));
	const int tB = tbase + ((xt & 1) ? 4 * ((xt + 1) >> 1) : 4 * (BOX_T_UNITS + (xt >> 1)));
	// offsets inside a ring slot: source units of columns j0, j0+4; target units xt, xt+1, xt+2 of the 40
	const int qs0 = 4 * cg, qs1 = 4 * (BOX_SS_UNITS + cg);
	const int qt0 = ((xt & 1) ? 4 * (BOX_ST_UNITS + (xt >> 1)) : 4 * (xt >> 1));
	const int qt1 = ((xt & 1) ? 4 * ((xt + 1) >> 1) : 4 * (BOX_ST_UNITS + (xt >> 1)));
	const bool live = d0 < p.D && j0 < p.Ws;
	float V[4][NV];
#pragma unroll
	for (int q = 0; q < 4; q++)
#pragma unroll
		for (int c = 0; c < NV; c++) V[q][c] = 0.f;
	// merge the 8 disparity groups' winners of one column and publish them: one 64-bit atomicMax per column, row and block
	const int kslot = (threadIdx.x & 7) * BOX_KEY_PITCH + (threadIdx.x >> 3);
	// The threads leave {best value (raw float bits) << 32 | index q of its disparity among the thread's four, -1: none, -2: the
	// NaN at d = 0 that sticks}; the merging thread of a column walks the 8 groups in ascending d with the reference's >= / <=
	// (a later equal value wins, -0 == +0) and converts the one winner into the ordered key.
	constexpr unsigned long long KEY_NONE = 0x00000000FFFFFFFFull;
	auto merge_keys = [&](int row, int buf) {
		if (jb + (int)threadIdx.x >= p.Ws) return;
		const unsigned long long* src = kb + buf * 8 * BOX_KEY_ROW + kslot;
		float bv = OP == BOX_PROD ? __int_as_float(0xff800000) : __int_as_float(0x7f800000);
		int bd = -1;
		bool stick = false;
#pragma unroll
		for (int g = 0; g < 8; g++) {
			const unsigned long long e = src[g * BOX_KEY_ROW];
			const float v = __uint_as_float((unsigned)(e >> 32));
			const int q = (int)(unsigned)e;
			if (g == 0) stick = q == -2;
			if (q >= 0 && (OP == BOX_PROD ? (v >= bv) : (v <= bv))) { bv = v; bd = db + 4 * g + q; }
		}
		unsigned long long key = 0;
		// + 0.f: -0 -> +0, the reference's comparisons see them as equal
		if (bd >= 0) key = ((unsigned long long)wta_order_bits((OP == BOX_PROD ? bv : -bv) + 0.f) << 32) | (unsigned)bd;
		if (stick) key = 0xFFFFFFFF00000000ull;
		if (key) atomicMax(p.keys + (int64_t)row * p.Ws + jb + threadIdx.x, key);
	};
	if (WTA && !live) {
#pragma unroll
		for (int k = 0; k < 8; k++) { kb[dg * BOX_KEY_ROW + k * BOX_KEY_PITCH + cg] = KEY_NONE; kb[(8 + dg) * BOX_KEY_ROW + k * BOX_KEY_PITCH + cg] = KEY_NONE; }
	}
	mbar_wait(&bar_img, 0);
	if (live)
		for (int r = 0; r <= 2 * p.vr; r++)
			box_accum_row<OP, NV, SIGN, true>(V, tile + r * BOX_ROW_FLOATS, sA, sB, tA, tB);
	// loop-invariant: which of the thread's four disparities exist and are stored, how its 8 columns leave, where row i0 starts
	bool qok[4];
#pragma unroll
	for (int q = 0; q < 4; q++) qok[q] = d0 + q < p.D && d0 + q >= p.d_first;
	const int smode = (j0 + 8 <= p.Ws) ? p.vec_store : 0;
	float* orow = p.out + (int64_t)i0 * p.o0 + (int64_t)d0 * p.o2 + (smode ? j0 : 0);
	const bool first_d = d0 == 0;   // only these threads can see the NaN-at-d-0 rule
	for (int i = i0; i < i_end; i++, orow += p.o0) {
		const int n = i - i0, slot = n % BOX_STAT_SLOTS;
		// every warp has finished row i-1, so its ring slot can take row i + BOX_STAT_SLOTS - 1
		__syncthreads();
		if (threadIdx.x < 32 && i + BOX_STAT_SLOTS - 1 < i_end) issue_stats(i + BOX_STAT_SLOTS - 1, (n + BOX_STAT_SLOTS - 1) % BOX_STAT_SLOTS);
		if (WTA && n > 0) merge_keys(i - 1, (n - 1) & 1);
		if (!live) continue;
		if (n > 0) {
			box_accum_row<OP, NV, SIGN, true>(V, tile + (n + 2 * p.vr) * BOX_ROW_FLOATS, sA, sB, tA, tB);
			box_accum_row<OP, NV, SIGN, false>(V, tile + (n - 1) * BOX_ROW_FLOATS, sA, sB, tA, tB);
		}
		mbar_wait(&bar_st[slot], (n / BOX_STAT_SLOTS) & 1);
		const float* st = stat_ring + slot * BOX_STAT_FLOATS;
		const float4 as0 = lds_unit(st, qs0), as1 = lds_unit(st, qs1);
		const float4 bs0 = lds_unit(st + 8 * BOX_SS_UNITS, qs0), bs1 = lds_unit(st + 8 * BOX_SS_UNITS, qs1);
		const float* stt = st + 16 * BOX_SS_UNITS;
		float at[12];
		if (OP != BOX_ABS) {
			const float4 at0 = lds_unit(stt, qt0), at1 = lds_unit(stt, qt1), at2 = lds_unit(stt, qt0 + 4);
			at[0] = at0.x; at[1] = at0.y; at[2] = at0.z; at[3] = at0.w; at[4] = at1.x; at[5] = at1.y; at[6] = at1.z; at[7] = at1.w;
			at[8] = at2.x; at[9] = at2.y; at[10] = at2.z; at[11] = at2.w;
		}
		const float4 bt0 = lds_unit(stt + 8 * BOX_ST_UNITS, qt0), bt1 = lds_unit(stt + 8 * BOX_ST_UNITS, qt1), bt2 = lds_unit(stt + 8 * BOX_ST_UNITS, qt0 + 4);
		const float as[8] = {as0.x, as0.y, as0.z, as0.w, as1.x, as1.y, as1.z, as1.w};
		const float bs[8] = {bs0.x, bs0.y, bs0.z, bs0.w, bs1.x, bs1.y, bs1.z, bs1.w};
		const float bt[12] = {bt0.x, bt0.y, bt0.z, bt0.w, bt1.x, bt1.y, bt1.z, bt1.w, bt2.x, bt2.y, bt2.z, bt2.w};
		float bestv[8];
		int bestq[8];
		unsigned nan0 = 0;
#pragma unroll
		for (int k = 0; k < 8; k++) { bestv[k] = OP == BOX_PROD ? __int_as_float(0xff800000) : __int_as_float(0x7f800000); bestq[k] = -1; }
		float* dst = orow;
#pragma unroll
		for (int q = 0; q < 4; q++, dst += p.o2) {
			if (!qok[q]) continue;
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
				if (WTA) {
					if (OP == BOX_PROD ? (v >= bestv[k]) : (v <= bestv[k])) { bestv[k] = v; bestq[k] = q; }
				}
			}
			if (WTA && q == 0 && first_d) {
#pragma unroll
				for (int k = 0; k < 8; k++) if (o[k] != o[k]) nan0 |= 1u << k;
			}
			if (smode == 2) {   // 32-byte aligned rows: one 256-bit store per sector (sm_100a, STG.E.256)
				asm volatile("st.global.cs.v8.f32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
							 :: "l"(dst), "f"(o[0]), "f"(o[1]), "f"(o[2]), "f"(o[3]), "f"(o[4]), "f"(o[5]), "f"(o[6]), "f"(o[7]) : "memory");
			} else if (smode) {
				__stcs(reinterpret_cast<float4*>(dst), make_float4(o[0], o[1], o[2], o[3]));
				__stcs(reinterpret_cast<float4*>(dst + 4), make_float4(o[4], o[5], o[6], o[7]));
			} else {
#pragma unroll
				for (int k = 0; k < 8; k++)
					if (j0 + k < p.Ws) dst[(int64_t)(j0 + k) * p.o1] = o[k];
			}
		}
		if (WTA) {
			if (first_d && nan0) {
#pragma unroll
				for (int k = 0; k < 8; k++) if ((nan0 >> k) & 1u) bestq[k] = -2;
			}
			uint2* dstk = reinterpret_cast<uint2*>(kb + ((n & 1) * 8 + dg) * BOX_KEY_ROW + cg);
#pragma unroll
			for (int k = 0; k < 8; k++) dstk[k * BOX_KEY_PITCH] = make_uint2((unsigned)bestq[k], __float_as_uint(bestv[k]));
		}
	}
	if (WTA) {
		__syncthreads();
		merge_keys(i_end - 1, (i_end - 1 - i0) & 1);
	}
}

template<int OP, int HR>
static int32_t launch_cv_box_hr(stevi_ctx* c, const BoxParams& p, int sign, dim3 grid) {
	const int smem = (p.R + 2 * p.vr) * BOX_ROW_FLOATS * 4 + BOX_STAT_SLOTS * BOX_STAT_FLOATS * 4;
#define STEVI_BOX_LAUNCH(S, W)                                                                                                     \
	do {                                                                                                                           \
		STEVI_CUDA(c, cudaFuncSetAttribute(k_cv_box<OP, HR, S, W>, cudaFuncAttributeMaxDynamicSharedMemorySize, BOX_SMEM_MAX));    \
		STEVI_LAUNCH(c, (k_cv_box<OP, HR, S, W>), grid, BOX_THREADS, smem, p);                                                     \
	} while (0)
	if (sign > 0) {
		if (p.keys) STEVI_BOX_LAUNCH(1, true);
		else STEVI_BOX_LAUNCH(1, false);
	} else {
		if (p.keys) STEVI_BOX_LAUNCH(-1, true);
		else STEVI_BOX_LAUNCH(-1, false);
	}
#undef STEVI_BOX_LAUNCH
	return STEVI_OK;
}

template<int OP>
int32_t launch_cv_box(stevi_ctx* c, const BoxParams& p, int hr, int sign) {
	dim3 grid(cdiv(p.Ws, 128), cdiv(p.D, BOX_D_PER_BLOCK), cdiv(p.H + p.row_shift, p.R));
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
