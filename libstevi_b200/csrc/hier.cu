// Average-pooling pyramid, guided cost volume and the coarse-to-fine driver.
//
//   interpolation/downsampling.h:67-118  averagePoolingDownsample: out ceil(H/wv) x ceil(W/wh); mean over the in-image part
//                                        of each window, accumulated in row-major window order, divided by the count.
//                                        Quirks kept: the inner loops both run to windows.horizontal() (:97,:101) and the
//                                        row offset uses the WIDTH remainder, the column offset the HEIGHT remainder (:80-84)
//   hierarchical.h:73-233                computeGuidedCV: bilinear guide upsampling in GLOBAL float coordinates
//                                        i*(h_g-1)/(h-1) (Q7: int product, float division), d0 = sign*round(2*interp),
//                                        best of the 2r+1 candidates with STRICT compare (first wins, unlike the dense WTA),
//                                        output window re-centred on the winner (:194-226)
//   hierarchical.h:235-314               hiearchicalTruncatedCostVolume (sic): pyramid by 2, full cost volume + WTA at the
//                                        coarsest level with disp_width halved per level, guided volumes on the way up
#include "common.cuh"
#include <type_traits>
#include "kernels.cuh"
#include "costvol_box.cuh"

namespace stevi {

namespace {

template<typename TI>
__global__ void __launch_bounds__(128) k_avgpool(const TI* __restrict__ img, int64_t s0, int64_t s1, int H, int W, int wh, int wv,
												 int off_v, int off_h, float* __restrict__ out, int64_t o0, int64_t o1, int nH, int nW) {
	const int j = blockIdx.x * 128 + threadIdx.x, i = blockIdx.y;
	if (j >= nW) return;
	float val = 0.f;
	int count = 0;
	for (int dv = 0; dv < wh; dv++) {          // sic: horizontal() bounds both loops
		const int pi = i * wv - off_v + dv;
		for (int dh = 0; dh < wh; dh++) {
			const int pj = j * wh - off_h + dh;
			if (pi >= 0 && pi < H && pj >= 0 && pj < W) { val = __fadd_rn(val, (float)img[pi * s0 + pj * s1]); count++; }
		}
	}
	out[i * o0 + j * o1] = __fdiv_rn(val, (float)count);
}

// 2x2 windows over an even-sized dense f32 image (every level of the coarse-to-fine pyramid of an even frame): a thread makes
// two output pixels from two 16-byte loads; same accumulation order ((a00 + a01) + a10) + a11, same division
__global__ void __launch_bounds__(128) k_avgpool_2x2(const float* __restrict__ img, int64_t s0, float* __restrict__ out, int64_t o0, int nW) {
	const int j = (blockIdx.x * 128 + threadIdx.x) * 2, i = blockIdx.y;
	if (j >= nW) return;
	const float4 a = __ldg(reinterpret_cast<const float4*>(img + (int64_t)(2 * i) * s0 + 2 * j));
	const float4 b = __ldg(reinterpret_cast<const float4*>(img + (int64_t)(2 * i + 1) * s0 + 2 * j));
	float2 v;
	v.x = __fdiv_rn(__fadd_rn(__fadd_rn(__fadd_rn(__fadd_rn(0.f, a.x), a.y), b.x), b.y), 4.f);
	v.y = __fdiv_rn(__fadd_rn(__fadd_rn(__fadd_rn(__fadd_rn(0.f, a.z), a.w), b.z), b.w), 4.f);
	*reinterpret_cast<float2*>(out + (int64_t)i * o0 + j) = v;
}

struct Guided {
	const float* ps; const float* pt; int64_t pitch; int half, pad_left;
	const float* Ms; const float* Ns; const float* Mt; const float* Nt;   // STATS_RAW maps (mean, norm)
	const int32_t* guide; int64_t g0, g1; int hg, wg;   // hg, wg: GLOBAL guide size (interpolation coordinates)
	int H, Ws, Wt, hr, vr, r, sign, op, zero_mean, normalized, cost;   // H: GLOBAL image height
	int n_rows, out_row0, img_row0, guide_row0;       // row band: output rows out_row0.., image/guide buffers start at these global rows
	float* tcv; int64_t t0, t1, t2;
	int32_t* disp; int64_t d0s, d1s;
};

constexpr int GUIDED_MAX_R = 4; // candidates at offsets -2r..2r around the guide: 4r+1 <= 17 accumulators

// One thread per pixel. The 2r+1 candidates around the upsampled guide and the 2r+1 cells around the winner are all
// comparisons against target columns j + d0 + e, e in [-2r, 2r]: each is evaluated once, over the raw window pixels
//   correlations: (sum s*t - F*m_s*m_t) / (n_s*n_t);  costs: sum op((s - m_s) - (t - m_t))
// instead of over materialised (v-mean)/norm feature volumes (49 x 4 bytes per pixel and image at C4).
__global__ void __launch_bounds__(128) k_guided_cv(const Guided p) {
	const int j = blockIdx.x * 128 + threadIdx.x, io = blockIdx.y;
	if (j >= p.Ws) return;
	const int gi = p.out_row0 + io;        // global row: the guide is upsampled in global coordinates (Q7)
	const int i = gi - p.img_row0;         // row inside the image band
	// hierarchical.h:116-148, same operation order in fp32
	const float v_pos = __fdiv_rn((float)(gi * (p.hg - 1)), (float)(p.H - 1));
	int v0 = (int)floorf(v_pos), v1 = (int)ceilf(v_pos);
	if (v0 == v1) v1 += 1;
	if (v1 == p.hg) { v0 -= 1; v1 -= 1; }
	const float h_pos = __fdiv_rn((float)(j * (p.wg - 1)), (float)(p.Ws - 1));
	int h0 = (int)floorf(h_pos), h1 = (int)ceilf(h_pos);
	if (h0 == h1) h1 += 1;
	if (h1 == p.wg) { h0 -= 1; h1 -= 1; }
	const float av = __fsub_rn(v_pos, (float)v0), bv = __fsub_rn((float)v1, v_pos);
	const float ah = __fsub_rn(h_pos, (float)h0), bh = __fsub_rn((float)h1, h_pos);
	float id = 0.f;
	const int64_t gr0 = (int64_t)(v0 - p.guide_row0) * p.g0, gr1 = (int64_t)(v1 - p.guide_row0) * p.g0;
	id = __fadd_rn(id, __fmul_rn(__fmul_rn(av, ah), (float)p.guide[gr1 + h1 * p.g1]));
	id = __fadd_rn(id, __fmul_rn(__fmul_rn(bv, ah), (float)p.guide[gr0 + h1 * p.g1]));
	id = __fadd_rn(id, __fmul_rn(__fmul_rn(av, bh), (float)p.guide[gr1 + h0 * p.g1]));
	id = __fadd_rn(id, __fmul_rn(__fmul_rn(bv, bh), (float)p.guide[gr0 + h0 * p.g1]));
	id = __fmul_rn(id, 2.f);
	const int d0 = p.sign * (int)roundf(id);

	const int hs = 2 * p.hr + 1, vs = 2 * p.vr + 1, nc = 4 * p.r + 1;
	const float F = (float)(hs * vs);
	const int64_t os = (int64_t)i * p.pitch + pad_col_off(p.pad_left + j, p.half);
	const float ms = p.Ms[os], ns = p.Ns[os];
	float acc[4 * GUIDED_MAX_R + 1];
#pragma unroll
	for (int e = 0; e < 4 * GUIDED_MAX_R + 1; e++) acc[e] = 0.f;
	const int jt0 = j + d0 - 2 * p.r;   // target column of candidate e = 0
	float mtv[4 * GUIDED_MAX_R + 1];    // target window means (costs of zero-mean functions subtract them per pixel)
#pragma unroll
	for (int e = 0; e < 4 * GUIDED_MAX_R + 1; e++) {
		const int jt = jt0 + e;
		mtv[e] = 0.f;
		if (e < nc && p.op != BOX_PROD && p.zero_mean && jt >= 0 && jt < p.Wt)
			mtv[e] = p.Mt[(int64_t)i * p.pitch + pad_col_off(p.pad_left + jt, p.half)];
	}
	for (int k = 0; k < vs; k++) {
		const float* srow = p.ps + (int64_t)(i - p.vr + k) * p.pitch;
		const float* trow = p.pt + (int64_t)(i - p.vr + k) * p.pitch;
		for (int l = 0; l < hs; l++) {
			float s = __ldg(srow + pad_col_off(p.pad_left + j - p.hr + l, p.half));
			if (p.op != BOX_PROD && p.zero_mean) s -= ms;
#pragma unroll
			for (int e = 0; e < 4 * GUIDED_MAX_R + 1; e++) {
				if (e >= nc) break;
				const int jt = jt0 + e;
				float t = 0.f;
				if (jt >= 0 && jt < p.Wt) t = __ldg(trow + pad_col_off(p.pad_left + jt - p.hr + l, p.half));
				if (p.op == BOX_PROD) acc[e] = fmaf(s, t, acc[e]);
				else if (p.op == BOX_SQ) { const float q = s - (t - mtv[e]); acc[e] = fmaf(q, q, acc[e]); }
				else acc[e] += fabsf(s - (t - mtv[e]));
			}
		}
	}
	if (p.op == BOX_PROD) {
#pragma unroll
		for (int e = 0; e < 4 * GUIDED_MAX_R + 1; e++) {
			if (e >= nc) break;
			const int jt = jt0 + e;
			float v = 0.f;
			if (jt >= 0 && jt < p.Wt) {
				const int64_t ot = (int64_t)i * p.pitch + pad_col_off(p.pad_left + jt, p.half);
				v = acc[e];
				if (p.zero_mean) v = fmaf(-F * ms, p.Mt[ot], v);
				if (p.normalized) {                          // flat windows: NaN features in the reference (Q6)
					const float nt = p.Nt[ot];
					v = (ns == 0.f || nt == 0.f) ? __int_as_float(0x7fc00000) : v / (ns * nt);
				}
			} else if (p.normalized && ns == 0.f) {
				v = __int_as_float(0x7fc00000);              // NaN source features times a zero target vector
			}
			acc[e] = v;
		}
	}
	// best of the candidates d0-r .. d0+r, strict compare: the first extremum wins (hierarchical.h:179,184)
	float score = p.cost ? __int_as_float(0x7f800000) : __int_as_float(0xff800000);
	int d_r = d0;
	for (int dd = -p.r; dd <= p.r; dd++) {
		float cmp = 0.f;
#pragma unroll
		for (int e = 0; e < 4 * GUIDED_MAX_R + 1; e++) if (e == dd + 2 * p.r) cmp = acc[e];
		if (p.cost ? (cmp < score) : (cmp > score)) { score = cmp; d_r = d0 + dd; }
	}
	p.disp[io * p.d0s + j * p.d1s] = p.sign * d_r;
	// cell dd holds the comparison against target column j + d_r + sign*(dd - r) (after the reference's shift-and-refill)
	float* o = p.tcv + io * p.t0 + j * p.t1;
	for (int dd = 0; dd <= 2 * p.r; dd++) {
		const int want = (d_r - d0) + p.sign * (dd - p.r) + 2 * p.r;
		float cmp = 0.f;
#pragma unroll
		for (int e = 0; e < 4 * GUIDED_MAX_R + 1; e++) if (e == want) cmp = acc[e];
		o[dd * p.t2] = cmp;
	}
}


// hierarchical.h:116-148: the guide upsampled at global row gi, column j, in the reference's fp32 operation order
__device__ __forceinline__ int guided_d0(const Guided& p, int gi, int j) {
	const float v_pos = __fdiv_rn((float)(gi * (p.hg - 1)), (float)(p.H - 1));
	int v0 = (int)floorf(v_pos), v1 = (int)ceilf(v_pos);
	if (v0 == v1) v1 += 1;
	if (v1 == p.hg) { v0 -= 1; v1 -= 1; }
	const float h_pos = __fdiv_rn((float)(j * (p.wg - 1)), (float)(p.Ws - 1));
	int h0 = (int)floorf(h_pos), h1 = (int)ceilf(h_pos);
	if (h0 == h1) h1 += 1;
	if (h1 == p.wg) { h0 -= 1; h1 -= 1; }
	const float av = __fsub_rn(v_pos, (float)v0), bv = __fsub_rn((float)v1, v_pos);
	const float ah = __fsub_rn(h_pos, (float)h0), bh = __fsub_rn((float)h1, h_pos);
	const int64_t gr0 = (int64_t)(v0 - p.guide_row0) * p.g0, gr1 = (int64_t)(v1 - p.guide_row0) * p.g0;
	float id = 0.f;
	id = __fadd_rn(id, __fmul_rn(__fmul_rn(av, ah), (float)p.guide[gr1 + h1 * p.g1]));
	id = __fadd_rn(id, __fmul_rn(__fmul_rn(bv, ah), (float)p.guide[gr0 + h1 * p.g1]));
	id = __fadd_rn(id, __fmul_rn(__fmul_rn(av, bh), (float)p.guide[gr1 + h0 * p.g1]));
	id = __fadd_rn(id, __fmul_rn(__fmul_rn(bv, bh), (float)p.guide[gr0 + h0 * p.g1]));
	id = __fmul_rn(id, 2.f);
	return p.sign * (int)roundf(id);
}

// the same split in its column part (once per thread) and its row part (once per pixel): identical operations, identical result
struct GuideCol { int h0, h1; float ah, bh; };
__device__ __forceinline__ GuideCol guided_col(const Guided& p, int j) {
	const float h_pos = __fdiv_rn((float)(j * (p.wg - 1)), (float)(p.Ws - 1));
	GuideCol c;
	c.h0 = (int)floorf(h_pos); c.h1 = (int)ceilf(h_pos);
	if (c.h0 == c.h1) c.h1 += 1;
	if (c.h1 == p.wg) { c.h0 -= 1; c.h1 -= 1; }
	c.ah = __fsub_rn(h_pos, (float)c.h0); c.bh = __fsub_rn((float)c.h1, h_pos);
	return c;
}
__device__ __forceinline__ int guided_d0_row(const Guided& p, int gi, const GuideCol& c) {
	const float v_pos = __fdiv_rn((float)(gi * (p.hg - 1)), (float)(p.H - 1));
	int v0 = (int)floorf(v_pos), v1 = (int)ceilf(v_pos);
	if (v0 == v1) v1 += 1;
	if (v1 == p.hg) { v0 -= 1; v1 -= 1; }
	const float av = __fsub_rn(v_pos, (float)v0), bv = __fsub_rn((float)v1, v_pos);
	const int64_t gr0 = (int64_t)(v0 - p.guide_row0) * p.g0, gr1 = (int64_t)(v1 - p.guide_row0) * p.g0;
	float id = 0.f;
	id = __fadd_rn(id, __fmul_rn(__fmul_rn(av, c.ah), (float)p.guide[gr1 + c.h1 * p.g1]));
	id = __fadd_rn(id, __fmul_rn(__fmul_rn(bv, c.ah), (float)p.guide[gr0 + c.h1 * p.g1]));
	id = __fadd_rn(id, __fmul_rn(__fmul_rn(av, c.bh), (float)p.guide[gr1 + c.h0 * p.g1]));
	id = __fadd_rn(id, __fmul_rn(__fmul_rn(bv, c.bh), (float)p.guide[gr0 + c.h0 * p.g1]));
	id = __fmul_rn(id, 2.f);
	return p.sign * (int)roundf(id);
}

// computeGuidedCV on caller-supplied feature volumes — the reference's own signature (hierarchical.h:73-233): the candidates and
// the re-centred window are comparisons of the source vector with target vectors (zeros outside the target image, :170),
// accumulated over the channels in the reference's order.
struct GuidedFeat {
	Guided g;
	const float* fs; int64_t s0, s1, s2;
	const float* ft; int64_t t0, t1, t2;
	int F;
};
__device__ __forceinline__ float guided_compare(const GuidedFeat& q, const float* s, const float* t, bool inside) {
	float acc = 0.f;
	for (int c = 0; c < q.F; c++) {
		const float sv = s[c * q.s2], tv = inside ? t[c * q.t2] : 0.f;
		if (q.g.op == BOX_PROD) acc = __fadd_rn(acc, __fmul_rn(sv, tv));
		else if (q.g.op == BOX_SQ) { const float e = __fsub_rn(sv, tv); acc = __fadd_rn(acc, __fmul_rn(e, e)); }
		else acc = __fadd_rn(acc, fabsf(__fsub_rn(sv, tv)));
	}
	return acc;
}
__global__ void __launch_bounds__(128) k_guided_cv_features(const GuidedFeat q) {
	const Guided& p = q.g;
	const int j = blockIdx.x * 128 + threadIdx.x, i = blockIdx.y;
	if (j >= p.Ws) return;
	const int d0 = guided_d0(p, i, j);
	const float* s = q.fs + i * q.s0 + j * q.s1;
	const float* trow = q.ft + i * q.t0;
	float score = p.cost ? __int_as_float(0x7f800000) : __int_as_float(0xff800000);
	int d_r = d0;
	for (int dd = -p.r; dd <= p.r; dd++) {   // strict compare: the first extremum wins (:179,:184)
		const int jt = j + d0 + dd;
		const bool inside = jt >= 0 && jt < p.Wt;
		const float cmp = guided_compare(q, s, trow + (int64_t)(inside ? jt : 0) * q.t1, inside);
		if (p.cost ? (cmp < score) : (cmp > score)) { score = cmp; d_r = d0 + dd; }
	}
	p.disp[i * p.d0s + j * p.d1s] = p.sign * d_r;
	// after the reference's shift-and-refill (:194-226) cell dd holds the comparison against target column j + d_r + sign*(dd - r)
	float* o = p.tcv + i * p.t0 + j * p.t1;
	for (int dd = 0; dd <= 2 * p.r; dd++) {
		const int jt = j + d_r + p.sign * (dd - p.r);
		const bool inside = jt >= 0 && jt < p.Wt;
		o[dd * p.t2] = guided_compare(q, s, trow + (int64_t)(inside ? jt : 0) * q.t1, inside);
	}
}

// Tiled form for windows up to 9 wide and upscale radii 1..2: a block owns 128 consecutive pixels of GT_ROWS consecutive
// rows. The image rows under those windows (GT_ROWS + 2vr of them) are staged once in shared memory, 16 bytes at a time:
// 136 source columns and the span of all candidate columns of the block (the guide being smooth). A thread then walks its
// GT_ROWS pixels; per window row it reads the target row once into registers and feeds all 4r+1 candidates from it
// (2hr+1 + 4r loads for (2hr+1)(4r+1) FMAs). Blocks whose candidates spread over more than GT_SPREAD columns (disparity
// edges) read the target through global memory instead.
constexpr int GT_SPREAD = 96;
constexpr int GT_ROWS = 8;
constexpr int GT_SRC_UNITS = (128 + 8) / 4;   // source columns jb-4 .. jb+131

template<int R> struct GtGeom {
	static constexpr int NC = 4 * R + 1;
	static constexpr int TGT_UNITS = (128 + 8 + NC + GT_SPREAD + 3 + 3) / 4;   // from the unit holding t_first-4
};

template<int HR, int R, int OP>
__global__ void __launch_bounds__(128) k_guided_cv_tile(const Guided p) {
	constexpr int HS = 2 * HR + 1, NC = 4 * R + 1, TN = HS + NC - 1;
	constexpr int SW = 4 * GT_SRC_UNITS, TU = GtGeom<R>::TGT_UNITS, TW = 4 * TU;
	extern __shared__ __align__(16) float gsm[];
	__shared__ int s_min[4], s_max[4];
	const int vs = 2 * p.vr + 1;
	const int tx = threadIdx.x, jb = blockIdx.x * 128, j = jb + tx, io0 = blockIdx.y * GT_ROWS;
	const int n_out = min(GT_ROWS, p.n_rows - io0);      // rows of this block
	const int n_stage = n_out + 2 * p.vr;
	float* ssrc = gsm;                     // n_stage x SW, tile row q <-> image row i0 - vr + q
	float* stgt = gsm + n_stage * SW;      // n_stage x TW
	float* smt = stgt + n_stage * TW;      // n_out x TW: target means under the staged columns
	float* snt = smt + n_out * TW;         // n_out x TW: target norms
	const int gi0 = p.out_row0 + io0, i0 = gi0 - p.img_row0;
	const bool live = j < p.Ws;
	// guide disparities of this thread's pixels, and the span of the block's candidate columns
	int d0r[GT_ROWS];
	int lo = 0x7fffffff, hi = -0x7fffffff;
	const GuideCol gcol = guided_col(p, live ? j : 0);
#pragma unroll
	for (int rr = 0; rr < GT_ROWS; rr++) {
		d0r[rr] = (live && rr < n_out) ? guided_d0_row(p, gi0 + rr, gcol) : 0;
		if (live && rr < n_out) { lo = min(lo, j + d0r[rr] - 2 * R); hi = max(hi, j + d0r[rr] - 2 * R); }
	}
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) { lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o)); hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o)); }
	if ((tx & 31) == 0) { s_min[tx >> 5] = lo; s_max[tx >> 5] = hi; }
	__syncthreads();
	const int tmin = min(min(s_min[0], s_min[1]), min(s_min[2], s_min[3])), tmax = max(max(s_max[0], s_max[1]), max(s_max[2], s_max[3]));
	// columns left of -(pad_left-8) are not in the padded planes: clamp the staged range to what exists; candidates
	// outside the target image are masked below and never read the tile
	const int t_first = max(tmin, -HR - 8);
	const int tu0 = (p.pad_left + t_first - 4) >> 2;     // first staged target unit; staged column c <-> image column 4*tu0 - pad_left + c
	const int t_col0 = 4 * tu0 - p.pad_left;
	const bool tiled = (tmax + NC + HR) - t_col0 <= TW && 4 * (tu0 + TU) <= (int)p.pitch - 8;   // last column read: tmax + NC-1 + HR
	{
		const int su0 = (p.pad_left + jb - 4) >> 2;
		for (int n = tx; n < n_stage * GT_SRC_UNITS; n += 128) {
			const int q = n / GT_SRC_UNITS, u = n - q * GT_SRC_UNITS;
			const float4 v = __ldg(reinterpret_cast<const float4*>(p.ps + (int64_t)(i0 - p.vr + q) * p.pitch + pad_unit_off(su0 + u, p.half)));
			*reinterpret_cast<float4*>(ssrc + q * SW + 4 * u) = v;
		}
		if (tiled)
			for (int n = tx; n < n_stage * TU; n += 128) {
				const int q = n / TU, u = n - q * TU;
				const float4 v = __ldg(reinterpret_cast<const float4*>(p.pt + (int64_t)(i0 - p.vr + q) * p.pitch + pad_unit_off(tu0 + u, p.half)));
				*reinterpret_cast<float4*>(stgt + q * TW + 4 * u) = v;
			}
		if (tiled)
			for (int n = tx; n < n_out * TU; n += 128) {
				const int q = n / TU, u = n - q * TU;
				const int64_t o = (int64_t)(i0 + q) * p.pitch + pad_unit_off(tu0 + u, p.half);
				*reinterpret_cast<float4*>(smt + q * TW + 4 * u) = __ldg(reinterpret_cast<const float4*>(p.Mt + o));
				*reinterpret_cast<float4*>(snt + q * TW + 4 * u) = __ldg(reinterpret_cast<const float4*>(p.Nt + o));
			}
	}
	__syncthreads();
	if (!live) return;
	const float F = (float)(HS * vs);
	const bool sub_mean = OP != BOX_PROD && p.zero_mean;
	// two copies of the pixel loop, so that the address arithmetic of the global-memory fallback is not hoisted into the
	// tiled one
	auto run = [&](auto tiled_tag) {
	constexpr bool TILED = decltype(tiled_tag)::value;
#pragma unroll 1
	for (int rr = 0; rr < n_out; rr++) {
		const int io = io0 + rr, i = i0 + rr;
		int d0 = 0;
#pragma unroll
		for (int q = 0; q < GT_ROWS; q++) if (q == rr) d0 = d0r[q];
		const int jt0 = j + d0 - 2 * R;
		const int64_t os = (int64_t)i * p.pitch + pad_col_off(p.pad_left + j, p.half);
		const float ms = p.Ms[os], ns = p.Ns[os];
		float acc[NC], mtv[NC], zacc = 0.f;   // zacc: the cost against the all-zero target vector
		// statistics of the candidate columns: from the staged rows, else through global memory. Threads whose candidates
		// all lie left of the staged range (outside the image, masked below) read column 0.
		const int tstat0 = rr * TW + max(jt0 - t_col0, 0);
		auto target_mean = [&](int e) { return TILED ? smt[tstat0 + e] : p.Mt[(int64_t)i * p.pitch + pad_col_off(p.pad_left + jt0 + e, p.half)]; };
		auto target_norm = [&](int e) { return TILED ? snt[tstat0 + e] : p.Nt[(int64_t)i * p.pitch + pad_col_off(p.pad_left + jt0 + e, p.half)]; };
#pragma unroll
		for (int e = 0; e < NC; e++) {
			acc[e] = 0.f; mtv[e] = 0.f;
			const int jt = jt0 + e;
			if (sub_mean && jt >= 0 && jt < p.Wt) mtv[e] = target_mean(e);
		}
		for (int k = 0; k < vs; k++) {
			float t[TN];
			if (TILED) {
				// threads whose candidates all lie left of the staged range (hence outside the image, masked below) read column 0
				const float* T = stgt + (rr + k) * TW + max(jt0 - HR - t_col0, 0);
#pragma unroll
				for (int n = 0; n < TN; n++) t[n] = T[n];
			} else {
				const float* trow = p.pt + (int64_t)(i - p.vr + k) * p.pitch;
#pragma unroll
				for (int n = 0; n < TN; n++) {
					const int X = jt0 - HR + n;   // may lie far outside the padded row when the guide points off the image
					t[n] = (X >= -p.pad_left + 8 && X < (int)p.pitch - p.pad_left - 8) ? __ldg(trow + pad_col_off(p.pad_left + X, p.half)) : 0.f;
				}
			}
			const float* S = ssrc + (rr + k) * SW + tx + 4 - HR;
#pragma unroll
			for (int l = 0; l < HS; l++) {
				float sv = S[l];
				if (sub_mean) sv -= ms;
				if (OP == BOX_SQ) zacc = fmaf(sv, sv, zacc);
				else if (OP == BOX_ABS) zacc += fabsf(sv);
#pragma unroll
				for (int e = 0; e < NC; e++) {
					if (OP == BOX_PROD) acc[e] = fmaf(sv, t[l + e], acc[e]);
					else if (OP == BOX_SQ) { const float q = sv - (t[l + e] - mtv[e]); acc[e] = fmaf(q, q, acc[e]); }
					else acc[e] += fabsf(sv - (t[l + e] - mtv[e]));
				}
			}
		}
		if (OP == BOX_PROD && TILED) {
			// branch-free: the 4R+1 candidates of a pixel are finished as straight-line code (ncu, round 2: this part was 575 of the
			// 1420 instructions per pixel, a third of it branches around the IEEE division and the inside / flat-window cases). The
			// quotient is num * rcp(den) (2 ulp; the volume is tolerance-matched at 1e-4); the cases that decide NaN
			// and 0 are the same exact tests as before.
			const float nan = __int_as_float(0x7fc00000);
			const float Fms = p.zero_mean ? F * ms : 0.f;
			const bool norm = p.normalized != 0, ns0 = norm && ns == 0.f;
			const float ns_eff = norm ? ns : 1.f;
			const float outside_v = ns0 ? nan : 0.f;   // NaN source features times a zero target vector (Q6)
#pragma unroll
			for (int e = 0; e < NC; e++) {
				const float mt = smt[tstat0 + e], nt = snt[tstat0 + e];
				const float num = fmaf(-Fms, mt, acc[e]);
				const float den = ns_eff * (norm ? nt : 1.f);
				float v = __fdividef(num, den);
				v = (ns0 || (norm && nt == 0.f)) ? nan : v;
				acc[e] = ((unsigned)(jt0 + e) < (unsigned)p.Wt) ? v : outside_v;
			}
		} else {
#pragma unroll
		for (int e = 0; e < NC; e++) {
			const int jt = jt0 + e;
			const bool inside = jt >= 0 && jt < p.Wt;
			float v;
			if (OP == BOX_PROD) {
				v = 0.f;
				if (inside) {
					v = acc[e];
					if (p.zero_mean) v = fmaf(-F * ms, target_mean(e), v);
					if (p.normalized) { const float nt = target_norm(e); v = (ns == 0.f || nt == 0.f) ? __int_as_float(0x7fc00000) : v / (ns * nt); }
				} else if (p.normalized && ns == 0.f) v = __int_as_float(0x7fc00000);
			} else v = inside ? acc[e] : zacc;
			acc[e] = v;
		}
		}
		float score = p.cost ? __int_as_float(0x7f800000) : __int_as_float(0xff800000);
		int best = R;   // index of the candidate at d0 (hierarchical.h: d_r starts at d0 and moves only on a strict improvement)
		bool moved = false;
#pragma unroll
		for (int dd = 0; dd <= 2 * R; dd++) {
			const float cmp = acc[R + dd];
			if (p.cost ? (cmp < score) : (cmp > score)) { score = cmp; best = dd; moved = true; }
		}
		const int d_r = moved ? d0 + best - R : d0;
		p.disp[io * p.d0s + j * p.d1s] = p.sign * d_r;
		float* o = p.tcv + io * p.t0 + j * p.t1;
		// re-centred window (hierarchical.h:194-226): cell dd is candidate (d_r - d0) + sign*(dd - R) + 2R, i.e. the 2R+1
		// consecutive candidates from `first` on, in the direction of the sign
		const int first = moved ? best : R;
#pragma unroll
		for (int b = 0; b <= 2 * R; b++) {
			if (first == b) {
#pragma unroll
				for (int dd = 0; dd <= 2 * R; dd++) o[dd * p.t2] = acc[b + (p.sign > 0 ? dd : 2 * R - dd)];
			}
		}
	}
	};
	if (tiled) run(std::true_type{}); else run(std::false_type{});
}

template<int HR, int R>
static int32_t launch_guided_tile(stevi_ctx* c, const Guided& p) {
	const int smem = ((GT_ROWS + 2 * p.vr) * (GT_SRC_UNITS + GtGeom<R>::TGT_UNITS) + 2 * GT_ROWS * GtGeom<R>::TGT_UNITS) * 16;
	dim3 grid(cdiv(p.Ws, 128), cdiv(p.n_rows, GT_ROWS));
#define GT_OP(OP_) { STEVI_CUDA(c, cudaFuncSetAttribute(k_guided_cv_tile<HR, R, OP_>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024)); \
		STEVI_LAUNCH(c, (k_guided_cv_tile<HR, R, OP_>), grid, 128, smem, p); return STEVI_OK; }
	if (p.op == BOX_PROD) GT_OP(BOX_PROD)
	if (p.op == BOX_SQ) GT_OP(BOX_SQ)
	GT_OP(BOX_ABS)
#undef GT_OP
}

static int32_t launch_guided(stevi_ctx* c, const Guided& p, dim3 grid) {
	if (p.hr >= 1 && p.hr <= 4 && (p.r == 1 || p.r == 2) &&
		((GT_ROWS + 2 * p.vr) * (GT_SRC_UNITS + GtGeom<2>::TGT_UNITS) + 2 * GT_ROWS * GtGeom<2>::TGT_UNITS) * 16 <= 100 * 1024) {
#define GT_CASE(HR_) case HR_: return p.r == 1 ? launch_guided_tile<HR_, 1>(c, p) : launch_guided_tile<HR_, 2>(c, p);
		switch (p.hr) { GT_CASE(1) GT_CASE(2) GT_CASE(3) GT_CASE(4) }
#undef GT_CASE
	}
	STEVI_LAUNCH(c, k_guided_cv, grid, 128, 0, p);
	return STEVI_OK;
}

} // namespace

namespace {
inline bool h_zero_mean(int mf) { return mf == STEVI_MF_ZCC || mf == STEVI_MF_ZNCC || mf == STEVI_MF_ZSSD || mf == STEVI_MF_ZSAD; }
inline bool h_normalized(int mf) { return mf == STEVI_MF_NCC || mf == STEVI_MF_ZNCC; }
inline bool h_score(int mf) { return mf == STEVI_MF_CC || mf == STEVI_MF_NCC || mf == STEVI_MF_ZCC || mf == STEVI_MF_ZNCC; }
inline int h_op(int mf) { return h_score(mf) ? BOX_PROD : (mf == STEVI_MF_SSD || mf == STEVI_MF_ZSSD) ? BOX_SQ : BOX_ABS; }
inline bool h_known(int mf) {
	switch (mf) {
	case STEVI_MF_CC: case STEVI_MF_NCC: case STEVI_MF_ZCC: case STEVI_MF_ZNCC:
	case STEVI_MF_SSD: case STEVI_MF_ZSSD: case STEVI_MF_SAD: case STEVI_MF_ZSAD: return true;
	default: return false;
	}
}

// dense (H,W) workspace array
template<typename T> int32_t alloc_map(stevi_ctx* c, int H, int W, int dtype, View& v) {
	T* p = nullptr;
	STEVI_TRY(arena_alloc_t(c, (size_t)H * W, &p));
	v.p = p; v.dtype = dtype; v.ndim = 2;
	v.shape[0] = H; v.shape[1] = W; v.st[0] = W; v.st[1] = 1;
	return STEVI_OK;
}
} // namespace

int32_t avgpool_dev(stevi_ctx* c, const View& img, int wh, int wv, const View& out) {
	if (img.ndim != 2 || (img.dtype != STEVI_F32 && img.dtype != STEVI_U8)) return fail(c, STEVI_E_DTYPE, "avgpool: image must be (H,W) u8|f32");
	if (wh < 1 || wv < 1) return fail(c, STEVI_E_ARG, "avgpool: window");
	const int H = img.shape[0], W = img.shape[1];
	const int nH = (H + wv - 1) / wv, nW = (W + wh - 1) / wh;
	if (out.ndim != 2 || out.dtype != STEVI_F32 || out.shape[0] != nH || out.shape[1] != nW)
		return fail(c, STEVI_E_SHAPE, "avgpool: out must be (%d,%d) f32", nH, nW);
	if (nH == 0 || nW == 0) return STEVI_OK;
	if (nH > 65535) return fail(c, STEVI_E_UNSUPPORTED, "avgpool: more than 65535 rows");
	// downsampling.h:80-84 (names as in the reference): hRem from the height, vRem from the width; the ROW index uses
	// initialVOffset (= vRem/2) and the COLUMN index initialHOffset (= hRem/2)
	const int hRem = nH * wv - H, vRem = nW * wh - W;
	dim3 grid(cdiv(nW, 128), nH);
	if (img.dtype == STEVI_F32 && wh == 2 && wv == 2 && H % 2 == 0 && W % 4 == 0 && img.st[1] == 1 && out.st[1] == 1 && img.st[0] % 4 == 0 &&
		out.st[0] % 2 == 0 && (uintptr_t)img.p % 16 == 0 && (uintptr_t)out.p % 8 == 0) {
		dim3 g2(cdiv(nW, 256), nH);
		STEVI_LAUNCH(c, k_avgpool_2x2, g2, 128, 0, img.as<float>(), img.st[0], out.as<float>(), out.st[0], nW);
		return STEVI_OK;
	}
	if (img.dtype == STEVI_F32) {
		STEVI_LAUNCH(c, k_avgpool<float>, grid, 128, 0, img.as<float>(), img.st[0], img.st[1], H, W, wh, wv, vRem / 2, hRem / 2,
					 out.as<float>(), out.st[0], out.st[1], nH, nW);
	} else {
		STEVI_LAUNCH(c, k_avgpool<uint8_t>, grid, 128, 0, img.as<uint8_t>(), img.st[0], img.st[1], H, W, wh, wv, vRem / 2, hRem / 2,
					 out.as<float>(), out.st[0], out.st[1], nH, nW);
	}
	return STEVI_OK;
}

int32_t guided_cv_dev(stevi_ctx* c, int mf, const View& l, const View& r, int hr, int vr, const View& guide, int radius, int direction,
					  const stevi_band* band, const View& tcv, const View& disp) {
	if (!h_known(mf)) return fail(c, STEVI_E_UNSUPPORTED, "guided_cv: matching function %d not available", mf);
	if (radius < 0 || radius > GUIDED_MAX_R) return fail(c, STEVI_E_UNSUPPORTED, "guided_cv: upscale radius must be 0..%d", GUIDED_MAX_R);
	if (hr < 0 || vr < 0 || hr > 127 || vr > 127) return fail(c, STEVI_E_ARG, "guided_cv: radii out of range");
	if (l.ndim == 2 && l.dtype == STEVI_U8 && (h_zero_mean(mf) || h_normalized(mf)))
		return fail(c, STEVI_E_UNSUPPORTED, "guided_cv: pass float images for zero-mean / normalised functions");
	if (guide.ndim != 2 || guide.dtype != STEVI_I32) return fail(c, STEVI_E_DTYPE, "guided_cv: guide must be (h_g,w_g) i32");
	PaddedPair pp;
	STEVI_TRY(pp.build(c, l, r, hr, vr, 0, 1));
	const bool r2l = direction == STEVI_RIGHT_TO_LEFT;
	const int si = r2l ? 1 : 0, ti = 1 - si;
	const int Ws = pp.W[si];
	// whole image: output rows = image rows, everything starts at global row 0
	const int H = band ? band->img_height : pp.H, hg = band ? band->guide_height : guide.shape[0];
	const int n_rows = band ? tcv.shape[0] : pp.H;
	// hierarchical.h:89-91 returns an empty volume on a shape mismatch; 1-pixel axes divide by zero (Q7)
	if (H < 2 || Ws < 2 || hg < 2 || guide.shape[1] < 2) return fail(c, STEVI_E_SHAPE, "guided_cv: every axis needs at least 2 pixels (hierarchical.h:116,132)");
	if (band) {
		const int lo = band->out_row0 - band->img_row0;
		if (band->img_row0 < 0 || lo < 0 || lo + n_rows > pp.H || band->out_row0 + n_rows > H || band->img_row0 + pp.H > H)
			return fail(c, STEVI_E_ARG, "guided_cv: output rows must lie inside the image band and the image");
		// the window of an output row may leave the band only where it also leaves the image
		if ((lo < vr && band->img_row0 > 0) || (lo + n_rows + vr > pp.H && band->img_row0 + pp.H < H))
			return fail(c, STEVI_E_ARG, "guided_cv: the image band needs v_radius halo rows on its inner edges");
		// guide rows touched by the bilinear upsampling of the band's rows (same arithmetic as the kernel, conservatively)
		const int v_lo = std::max(0, (int)((int64_t)band->out_row0 * (hg - 1) / (H - 1)) - 1);
		const int v_hi = std::min(hg - 1, (int)((int64_t)(band->out_row0 + n_rows - 1) * (hg - 1) / (H - 1)) + 2);
		if (band->guide_row0 > v_lo || band->guide_row0 + guide.shape[0] - 1 < v_hi || band->guide_row0 < 0)
			return fail(c, STEVI_E_ARG, "guided_cv: the guide band must cover global guide rows %d..%d", v_lo, v_hi);
	}
	if (tcv.ndim != 3 || tcv.dtype != STEVI_F32 || tcv.shape[0] != n_rows || tcv.shape[1] != Ws || tcv.shape[2] != 2 * radius + 1)
		return fail(c, STEVI_E_SHAPE, "guided_cv: tcv must be (%d,%d,%d) f32", n_rows, Ws, 2 * radius + 1);
	if (disp.ndim != 2 || disp.dtype != STEVI_I32 || disp.shape[0] != n_rows || disp.shape[1] != Ws)
		return fail(c, STEVI_E_SHAPE, "guided_cv: disp must be (%d,%d) i32", n_rows, Ws);
	if (n_rows == 0) return STEVI_OK;
	STEVI_TRY(pp.stats(c, si, hr, vr, mf, 0 /* raw: mean, norm */));
	STEVI_TRY(pp.stats(c, ti, hr, vr, mf, 0));
	Guided p;
	p.ps = pp.img[si]; p.pt = pp.img[ti]; p.pitch = pp.pitch; p.half = pp.half; p.pad_left = pp.pad_left;
	p.Ms = pp.A[si]; p.Ns = pp.B[si]; p.Mt = pp.A[ti]; p.Nt = pp.B[ti];
	p.guide = guide.as<int32_t>(); p.g0 = guide.st[0]; p.g1 = guide.st[1]; p.hg = hg; p.wg = guide.shape[1];
	p.n_rows = n_rows; p.out_row0 = band ? band->out_row0 : 0; p.img_row0 = band ? band->img_row0 : 0; p.guide_row0 = band ? band->guide_row0 : 0;
	p.H = H; p.Ws = Ws; p.Wt = pp.W[ti]; p.hr = hr; p.vr = vr; p.r = radius; p.sign = r2l ? 1 : -1;
	p.op = h_op(mf); p.zero_mean = h_zero_mean(mf); p.normalized = h_normalized(mf); p.cost = h_score(mf) ? 0 : 1;
	p.tcv = tcv.as<float>(); p.t0 = tcv.st[0]; p.t1 = tcv.st[1]; p.t2 = tcv.st[2];
	p.disp = disp.as<int32_t>(); p.d0s = disp.st[0]; p.d1s = disp.st[1];
	dim3 grid(cdiv(Ws, 128), n_rows);
	return launch_guided(c, p, grid);
}

int32_t guided_cv_features_dev(stevi_ctx* c, int mf, const View& fl, const View& fr, const View& guide, int radius, int direction,
							   const View& tcv, const View& disp) {
	if (!h_known(mf)) return fail(c, STEVI_E_UNSUPPORTED, "guided_cv_features: matching function %d not available", mf);
	if (radius < 0 || radius > 64) return fail(c, STEVI_E_ARG, "guided_cv_features: upscale radius");
	if (fl.ndim != 3 || fr.ndim != 3 || fl.dtype != STEVI_F32 || fr.dtype != STEVI_F32) return fail(c, STEVI_E_DTYPE, "guided_cv_features: feature volumes must be (H,W,F) f32");
	if (guide.ndim != 2 || guide.dtype != STEVI_I32) return fail(c, STEVI_E_DTYPE, "guided_cv_features: guide must be (h_g,w_g) i32");
	// hierarchical.h:89-91: an empty volume when the heights differ
	if (fl.shape[0] != fr.shape[0] || fl.shape[2] != fr.shape[2]) return fail(c, STEVI_E_SHAPE, "guided_cv_features: the feature volumes do not match");
	const bool r2l = direction == STEVI_RIGHT_TO_LEFT;
	const View& src = r2l ? fr : fl;
	const View& tgt = r2l ? fl : fr;
	const int H = src.shape[0], Ws = src.shape[1];
	if (H < 2 || Ws < 2 || guide.shape[0] < 2 || guide.shape[1] < 2) return fail(c, STEVI_E_SHAPE, "guided_cv_features: every axis needs at least 2 pixels (hierarchical.h:116,132)");
	if (H > 65535) return fail(c, STEVI_E_UNSUPPORTED, "guided_cv_features: more than 65535 rows");
	if (tcv.ndim != 3 || tcv.dtype != STEVI_F32 || tcv.shape[0] != H || tcv.shape[1] != Ws || tcv.shape[2] != 2 * radius + 1)
		return fail(c, STEVI_E_SHAPE, "guided_cv_features: tcv must be (%d,%d,%d) f32", H, Ws, 2 * radius + 1);
	if (disp.ndim != 2 || disp.dtype != STEVI_I32 || disp.shape[0] != H || disp.shape[1] != Ws)
		return fail(c, STEVI_E_SHAPE, "guided_cv_features: disp must be (%d,%d) i32", H, Ws);
	GuidedFeat q;
	Guided& p = q.g;
	p = Guided{};
	p.guide = guide.as<int32_t>(); p.g0 = guide.st[0]; p.g1 = guide.st[1]; p.hg = guide.shape[0]; p.wg = guide.shape[1];
	p.H = H; p.Ws = Ws; p.Wt = tgt.shape[1]; p.r = radius; p.sign = r2l ? 1 : -1;
	p.op = h_op(mf); p.cost = h_score(mf) ? 0 : 1;
	p.n_rows = H;
	p.tcv = tcv.as<float>(); p.t0 = tcv.st[0]; p.t1 = tcv.st[1]; p.t2 = tcv.st[2];
	p.disp = disp.as<int32_t>(); p.d0s = disp.st[0]; p.d1s = disp.st[1];
	q.fs = src.as<float>(); q.s0 = src.st[0]; q.s1 = src.st[1]; q.s2 = src.st[2];
	q.ft = tgt.as<float>(); q.t0 = tgt.st[0]; q.t1 = tgt.st[1]; q.t2 = tgt.st[2];
	q.F = src.shape[2];
	dim3 grid(cdiv(Ws, 128), H);
	STEVI_LAUNCH(c, k_guided_cv_features, grid, 128, 0, q);
	return STEVI_OK;
}

// hierarchical.h:235-314. h_radii / v_radii hold depth+1 entries, coarsest level first (like the reference's arrays).
int32_t hierarchical_dev(stevi_ctx* c, int mf, int depth, const View& l, const View& r, const int32_t* h_radii, const int32_t* v_radii,
						 int D, int radius, int direction, const View& tcv, const View& disp) {
	if (depth < 1 || depth > 8) return fail(c, STEVI_E_ARG, "hierarchical: depth must be 1..8");
	if (l.ndim != 2 || r.ndim != 2 || l.dtype != STEVI_F32 || r.dtype != STEVI_F32)
		return fail(c, STEVI_E_UNSUPPORTED, "hierarchical: images must be (H,W) f32");
	if (l.shape[0] != r.shape[0]) return fail(c, STEVI_E_SHAPE, "hierarchical: image heights differ");
	// pyramid: level `depth` = the input, level 0 = the coarsest
	std::vector<View> pl(depth + 1), pr(depth + 1);
	std::vector<int> dw(depth + 1);
	pl[depth] = l; pr[depth] = r; dw[depth] = D;
	for (int k = depth - 1; k >= 0; k--) {
		STEVI_TRY(alloc_map<float>(c, (pl[k + 1].shape[0] + 1) / 2, (pl[k + 1].shape[1] + 1) / 2, STEVI_F32, pl[k]));
		STEVI_TRY(alloc_map<float>(c, (pr[k + 1].shape[0] + 1) / 2, (pr[k + 1].shape[1] + 1) / 2, STEVI_F32, pr[k]));
		STEVI_TRY(avgpool_dev(c, pl[k + 1], 2, 2, pl[k]));
		STEVI_TRY(avgpool_dev(c, pr[k + 1], 2, 2, pr[k]));
		dw[k] = (dw[k + 1] + 1) / 2;
	}
	const bool r2l = direction == STEVI_RIGHT_TO_LEFT;
	// coarsest level: full cost volume + winner-take-all (raw indices are the guide, hierarchical.h:262)
	const View& sl = r2l ? pr[0] : pl[0];
	View cv, guide;
	{
		float* p = nullptr;
		STEVI_TRY(arena_alloc_t(c, (size_t)sl.shape[0] * sl.shape[1] * dw[0], &p));
		cv.p = p; cv.dtype = STEVI_F32; cv.ndim = 3;
		cv.shape[0] = sl.shape[0]; cv.shape[1] = sl.shape[1]; cv.shape[2] = dw[0];
		cv.st[0] = (int64_t)sl.shape[1] * dw[0]; cv.st[1] = 1; cv.st[2] = sl.shape[1];
	}
	STEVI_TRY(alloc_map<int32_t>(c, sl.shape[0], sl.shape[1], STEVI_I32, guide));
	// the winner-take-all rides on the cost-volume kernel where the box kernel applies (falls back to the staged WTA inside)
	STEVI_TRY(float_cv_dev(c, mf, pl[0], pr[0], h_radii[0], v_radii[0], dw[0], direction, cv, false, &guide));
	for (int k = 1; k <= depth; k++) {
		const View& s = r2l ? pr[k] : pl[k];
		View t, d;
		if (k == depth) { t = tcv; d = disp; }
		else {
			float* p = nullptr;
			STEVI_TRY(arena_alloc_t(c, (size_t)s.shape[0] * s.shape[1] * (2 * radius + 1), &p));
			t.p = p; t.dtype = STEVI_F32; t.ndim = 3;
			t.shape[0] = s.shape[0]; t.shape[1] = s.shape[1]; t.shape[2] = 2 * radius + 1;
			t.st[0] = (int64_t)s.shape[1] * t.shape[2]; t.st[1] = t.shape[2]; t.st[2] = 1;
			STEVI_TRY(alloc_map<int32_t>(c, s.shape[0], s.shape[1], STEVI_I32, d));
		}
		STEVI_TRY(guided_cv_dev(c, mf, pl[k], pr[k], h_radii[k], v_radii[k], guide, radius, direction, nullptr, t, d));
		guide = d;
	}
	return STEVI_OK;
}

} // namespace stevi
