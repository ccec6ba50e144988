// Float cost volumes: CC / NCC / ZCC / ZNCC / SSD / ZSSD / SAD / ZSAD.
//
//   cross_correlations.h:740-765  unfoldBasedCostVolume = unfold both images -> getFeatureVolumeForMatchFunc -> aggregateCost
//   correlation_base.h:1100-1136  channelsMean: mean = (sum_c v) * (1/F)     (sum in channel order, then one multiply)
//   cross_correlations.h:61-104   channelsZeroMeanNorm: norm = sqrt(sum_c (v-mean)^2);  :149-191 channelsNorm: sqrt(sum_c v^2)
//   cross_correlations.h:452      feature = (v-mean)/norm, no epsilon: a flat window gives 0/0 = NaN (Q6)
//
// Two evaluation paths:
//   * box path (costvol_box.cuh): O(1) per cell, used for window half-widths 1..4 and every function but ZSAD;
//   * direct path (k_cv_direct): the reference's own O(F) sum per cell, in its channel order — any window, ZSAD too.
// Both read zero-padded float copies of the images (k_pad_image) and per-pixel window statistics (k_window_stats*).
#include "common.cuh"
#include "kernels.cuh"
#include "costvol_box.cuh"

namespace stevi {

namespace {

inline bool mf_zero_mean(int mf) { return mf == STEVI_MF_ZCC || mf == STEVI_MF_ZNCC || mf == STEVI_MF_ZSSD || mf == STEVI_MF_ZSAD; }
inline bool mf_normalized(int mf) { return mf == STEVI_MF_NCC || mf == STEVI_MF_ZNCC; }
__host__ __device__ inline int mf_op(int mf) {
	switch (mf) {
	case STEVI_MF_CC: case STEVI_MF_NCC: case STEVI_MF_ZCC: case STEVI_MF_ZNCC: return BOX_PROD;
	case STEVI_MF_SSD: case STEVI_MF_ZSSD: return BOX_SQ;
	default: return BOX_ABS;
	}
}

// zero-padded float copy in the unit-de-interleaved row layout of costvol_box.cuh: pixel (i,j) at
// out[i*pitch + pad_col_off(pad_left + j)] (out points at image row 0); the buffer was cleared beforehand
template<typename TI>
__global__ void __launch_bounds__(256) k_pad_image(const TI* __restrict__ img, int64_t s0, int64_t s1, int H, int W,
												   float* __restrict__ out, int64_t pitch, int half, int pad_left) {
	const int j = blockIdx.x * 256 + threadIdx.x, i = blockIdx.y;
	if (j < W) out[(int64_t)i * pitch + pad_col_off(pad_left + j, half)] = (float)img[i * s0 + j * s1];
}

// the same for rows that can be read 4 pixels at a time (column stride 1, 4-pixel-aligned rows): a thread moves one unit
template<typename TI>
__global__ void __launch_bounds__(256) k_pad_image_units(const TI* __restrict__ img, int64_t s0, int H, int W,
														 float* __restrict__ out, int64_t pitch, int half, int pad_left) {
	const int j = (blockIdx.x * 256 + threadIdx.x) * 4, i = blockIdx.y;
	if (j >= W) return;
	float4 v;
	if (j + 4 <= W) {
		if (sizeof(TI) == 4) v = __ldg(reinterpret_cast<const float4*>(img + i * s0 + j));
		else { const uchar4 b = __ldg(reinterpret_cast<const uchar4*>(img + i * s0 + j)); v = make_float4((float)b.x, (float)b.y, (float)b.z, (float)b.w); }
	} else {
		const TI* q = img + i * s0 + j;
		v = make_float4((float)q[0], j + 1 < W ? (float)q[1] : 0.f, j + 2 < W ? (float)q[2] : 0.f, 0.f);
	}
	*reinterpret_cast<float4*>(out + (int64_t)i * pitch + pad_unit_off((pad_left + j) >> 2, half)) = v;
}

// what the statistics maps hold
enum { STATS_RAW = 0,       // A = mean, B = norm                               (direct path, feature volumes)
	   STATS_PROD_SRC = 1,  // A = F*mean, B = 1/norm (NaN when norm == 0)
	   STATS_PROD_TGT = 2,  // A = mean,   B = 1/norm
	   STATS_DIFF_SRC = 3,  // A = mean,   B = cost of the window against the all-zero vector (target outside the image)
	   STATS_DIFF_TGT = 4 };// A = mean,   B = 1

struct StatsParams {
	const float* pimg;       // padded image plane at (image row 0, padded column 0)
	int64_t pitch;
	int half, pad_left;
	int H, W, hr, vr;
	int zero_mean, normalized, op, kind;
	float* A; float* B;      // same geometry as the padded image
};

__device__ __forceinline__ void stats_finish(const StatsParams& p, float mean, float acc, float& a, float& b) {
	// acc: PROD -> sum of squares (normalised functions only); DIFF -> cost against zero
	const float F = (float)((2 * p.hr + 1) * (2 * p.vr + 1));
	float norm = p.normalized ? sqrtf(acc) : 1.f;
	switch (p.kind) {
	case STATS_RAW: a = mean; b = norm; break;
	case STATS_PROD_SRC: a = F * mean; b = p.normalized ? (norm == 0.f ? __int_as_float(0x7fc00000) : 1.f / norm) : 1.f; break;
	case STATS_PROD_TGT: a = mean; b = p.normalized ? (norm == 0.f ? __int_as_float(0x7fc00000) : 1.f / norm) : 1.f; break;
	case STATS_DIFF_SRC: a = mean; b = acc; break;
	default: a = mean; b = 1.f; break;
	}
}

// acc + term(f). Squares go through one fused multiply-add: the reference's `acc += f*f` is contracted the same way by
// its compiler whenever the host has FMA (GCC's default -ffp-contract=fast), and it is a third fewer instructions here.
__device__ __forceinline__ float stats_accum(int op, bool normalized, float acc, float f) {
	if (op == BOX_PROD) return normalized ? __fmaf_rn(f, f, acc) : acc;
	if (op == BOX_SQ) return __fmaf_rn(f, f, acc);
	return __fadd_rn(acc, fabsf(f));
}

// any window: one thread per pixel, channel order c = (2hr+1)k + l like the reference
__global__ void __launch_bounds__(128) k_window_stats_generic(const StatsParams p) {
	const int j = blockIdx.x * 128 + threadIdx.x, i = blockIdx.y;
	if (j >= p.W) return;
	const float* base = p.pimg + (int64_t)(i - p.vr) * p.pitch;
	const int X0 = p.pad_left + j - p.hr;
	const int hs = 2 * p.hr + 1, vs = 2 * p.vr + 1;
	float mean = 0.f;
	if (p.zero_mean) {
		for (int k = 0; k < vs; k++) for (int l = 0; l < hs; l++) mean = __fadd_rn(mean, __ldg(base + k * p.pitch + pad_col_off(X0 + l, p.half)));
		mean = __fmul_rn(mean, 1.f / (float)(hs * vs));
	}
	float acc = 0.f;
	if (p.op != BOX_PROD || p.normalized) {
		for (int k = 0; k < vs; k++) for (int l = 0; l < hs; l++) {
			float f = __ldg(base + k * p.pitch + pad_col_off(X0 + l, p.half));
			if (p.zero_mean) f = __fsub_rn(f, mean);
			acc = stats_accum(p.op, p.normalized, acc, f);
		}
	}
	float a, b;
	stats_finish(p, mean, acc, a, b);
	const int64_t o = (int64_t)i * p.pitch + pad_col_off(p.pad_left + j, p.half);
	p.A[o] = a;
	p.B[o] = b;
}

// windows up to 9 wide: a thread owns 8 consecutive pixels (two 16-byte units) and reads each window row as four aligned
// 16-byte loads (units u-1 .. u+2 hold every column its windows touch); same summation order as the reference, a
// twentieth of the load instructions and half the L1 traffic per pixel of a 4-pixel thread (the kernel is L1-bound)
template<int HR>
__global__ void __launch_bounds__(128) k_window_stats(const StatsParams p) {
	constexpr int PX = 8;
	const int j0 = (blockIdx.x * 128 + threadIdx.x) * PX, i = blockIdx.y;
	if (j0 >= p.W) return;
	const float* base = p.pimg + (int64_t)(i - p.vr) * p.pitch;
	const int u = (p.pad_left + j0) >> 2;   // first of this thread's two units (even: pad_left and j0 are multiples of 8)
	const int oA = pad_unit_off(u - 1, p.half), oB = pad_unit_off(u, p.half);
	const int vs = 2 * p.vr + 1;
	constexpr int HS = 2 * HR + 1;
	float mean[PX];
#pragma unroll
	for (int q = 0; q < PX; q++) mean[q] = 0.f;
	if (p.zero_mean) {
		for (int k = 0; k < vs; k++) {
			float r[16];
			load_units<4>(r, base + k * p.pitch, oA, oB);
#pragma unroll
			for (int q = 0; q < PX; q++)
#pragma unroll
				for (int l = 0; l < HS; l++) mean[q] = __fadd_rn(mean[q], r[4 - HR + q + l]);
		}
		const float sc = 1.f / (float)(HS * vs);
#pragma unroll
		for (int q = 0; q < PX; q++) mean[q] = __fmul_rn(mean[q], sc);
	}
	float acc[PX];
#pragma unroll
	for (int q = 0; q < PX; q++) acc[q] = 0.f;
	// the accumulation is decided once, outside the unrolled loops (as per-element selects it was 49 of 270 instructions per pixel);
	// without zero mean, mean[] is 0 and f - 0 is f, bit for bit
	const int acc_mode = p.op == BOX_PROD ? (p.normalized ? 1 : 0) : (p.op == BOX_SQ ? 1 : 2);   // 1: squares, 2: absolute values
	if (acc_mode == 1) {
		for (int k = 0; k < vs; k++) {
			float r[16];
			load_units<4>(r, base + k * p.pitch, oA, oB);
#pragma unroll
			for (int q = 0; q < PX; q++)
#pragma unroll
				for (int l = 0; l < HS; l++) {
					const float f = __fsub_rn(r[4 - HR + q + l], mean[q]);
					acc[q] = __fmaf_rn(f, f, acc[q]);
				}
		}
	} else if (acc_mode == 2) {
		for (int k = 0; k < vs; k++) {
			float r[16];
			load_units<4>(r, base + k * p.pitch, oA, oB);
#pragma unroll
			for (int q = 0; q < PX; q++)
#pragma unroll
				for (int l = 0; l < HS; l++) acc[q] = __fadd_rn(acc[q], fabsf(__fsub_rn(r[4 - HR + q + l], mean[q])));
		}
	}
	float a[PX], b[PX];
#pragma unroll
	for (int q = 0; q < PX; q++) stats_finish(p, mean[q], acc[q], a[q], b[q]);
	// columns past W inside the last group must stay zero (they are "outside the image" for the box kernel)
	const int64_t row = (int64_t)i * p.pitch;
#pragma unroll
	for (int q = 0; q < PX; q++) if (j0 + q < p.W) {
		const int64_t o = row + (q < 4 ? oB + q : oA + 4 + (q - 4));
		p.A[o] = a[q]; p.B[o] = b[q];
	}
}

// The reference's own sum per cell: features rebuilt on the fly as (v-mean)/norm, accumulated in channel order.
struct DirectParams {
	const float* ps; const float* pt; int64_t pitch; int half, pad_left;
	const float* Ms; const float* Ns; const float* Mt; const float* Nt;   // STATS_RAW maps
	int H, Ws, Wt, D, hr, vr, sign, mf, zero_mean, normalized;
	int dv, col_off;   // 2-D search volumes: the target pixel of (i, j, d) is (i + dv, j + sign*d + col_off); 0, 0 otherwise
	float* out; int64_t o0, o1, o2;
};

__global__ void __launch_bounds__(128) k_cv_direct(const DirectParams p) {
	const int j = blockIdx.x * 128 + threadIdx.x, d = blockIdx.y, i = blockIdx.z;
	if (j >= p.Ws) return;
	const int jt = j + p.sign * d + p.col_off, it = i + p.dv;
	const bool inside = jt >= 0 && jt < p.Wt && it >= 0 && it < p.H;
	const int hs = 2 * p.hr + 1, vs = 2 * p.vr + 1;
	const float* sb = p.ps + (int64_t)(i - p.vr) * p.pitch;
	const float* tb = p.pt + (int64_t)(it - p.vr) * p.pitch;
	const int Xs = p.pad_left + j - p.hr, Xt = p.pad_left + jt - p.hr;
	const int64_t os = (int64_t)i * p.pitch + pad_col_off(p.pad_left + j, p.half);
	const float ms = p.Ms[os], ns = p.Ns[os];
	float mt = 0.f, nt = 1.f;
	if (inside) {
		const int64_t ot = (int64_t)it * p.pitch + pad_col_off(p.pad_left + jt, p.half);
		mt = p.Mt[ot]; nt = p.Nt[ot];
	}
	const int op = mf_op(p.mf);
	float acc = 0.f;
	for (int k = 0; k < vs; k++) for (int l = 0; l < hs; l++) {
		float s = __ldg(sb + k * p.pitch + pad_col_off(Xs + l, p.half));
		if (p.zero_mean) s = __fsub_rn(s, ms);
		if (p.normalized) s = __fdiv_rn(s, ns);
		float t = 0.f;
		if (inside) {
			t = __ldg(tb + k * p.pitch + pad_col_off(Xt + l, p.half));
			if (p.zero_mean) t = __fsub_rn(t, mt);
			if (p.normalized) t = __fdiv_rn(t, nt);
		}
		if (op == BOX_PROD) acc = __fadd_rn(acc, __fmul_rn(s, t));
		else if (op == BOX_SQ) { float e = __fsub_rn(s, t); acc = __fadd_rn(acc, __fmul_rn(e, e)); }
		else acc = __fadd_rn(acc, fabsf(__fsub_rn(s, t)));
	}
	p.out[(int64_t)i * p.o0 + (int64_t)j * p.o1 + (int64_t)d * p.o2] = acc;
}

// ---- 2-D search volumes: feature planes + a register-tiled comparison -------------------------------------------------
// k_cv_direct rebuilds both feature vectors of every cell ((v - mean) / norm per element: two IEEE divisions per window
// element and cell), which is what makes it exact and what makes it slow once a search window has hundreds of candidates.
// Here each image's features are evaluated once, with the very same operations, into F planes of H x W floats
// (k_feature_planes), and the volume kernel only reads them: cell (i, j, a, b) accumulates op(S[f][i][j], T[f][it][jt]) over
// f in the reference's channel order with separately rounded operations - the same bits as k_cv_direct. A thread owns one
// source pixel and four consecutive column candidates: per feature one source load and four target loads that are
// contiguous across the warp.
struct PlaneParams {
	const float* img; int64_t pitch; int half, pad_left;
	const float* M; const float* N;
	int H, W, hr, vr, zero_mean, normalized;
	float* planes;   // [F][H][W]
};

__global__ void __launch_bounds__(128) k_feature_planes(const PlaneParams p) {
	// thread per pixel, all F features: mean and norm read once, plane writes contiguous across the warp
	const int j = blockIdx.x * 128 + threadIdx.x, i = blockIdx.y;
	if (j >= p.W) return;
	const int hs = 2 * p.hr + 1, vs = 2 * p.vr + 1;
	const int64_t o = (int64_t)i * p.pitch + pad_col_off(p.pad_left + j, p.half);
	const float m = p.zero_mean ? p.M[o] : 0.f, n = p.normalized ? p.N[o] : 1.f;
	const int64_t plane = (int64_t)p.H * p.W;
	float* dst = p.planes + (int64_t)i * p.W + j;
	for (int k = 0; k < vs; k++) {
		const float* row = p.img + (int64_t)(i - p.vr + k) * p.pitch;
		for (int l = 0; l < hs; l++) {
			float v = __ldg(row + pad_col_off(p.pad_left + j - p.hr + l, p.half));
			if (p.zero_mean) v = __fsub_rn(v, m);
			if (p.normalized) v = __fdiv_rn(v, n);
			*dst = v;
			dst += plane;
		}
	}
}

struct Cv2dParams {
	const float* S; const float* T;   // [F][H][W]
	int H, W, F, dH, dW, lo0, lo1;
	float* out; int64_t o0, o1, o2, o3;
};

template<int OP>
__global__ void __launch_bounds__(128) k_cv2d_planes(const Cv2dParams p) {
	const int j = blockIdx.x * 128 + threadIdx.x, i = blockIdx.z;
	if (j >= p.W) return;
	const int groups = (p.dW + 3) >> 2;
	const int a = blockIdx.y / groups, b0 = (blockIdx.y - a * groups) * 4;
	const int it = i + a + p.lo0, jt0 = j + b0 + p.lo1;
	const bool row_in = it >= 0 && it < p.H;
	bool in[4];
#pragma unroll
	for (int q = 0; q < 4; q++) in[q] = row_in && jt0 + q >= 0 && jt0 + q < p.W;
	const int64_t plane = (int64_t)p.H * p.W;
	const float* s = p.S + (int64_t)i * p.W + j;
	const float* t = p.T + (int64_t)(row_in ? it : 0) * p.W + jt0;
	float acc[4] = {0.f, 0.f, 0.f, 0.f};
	for (int f = 0; f < p.F; f++) {
		const float sv = __ldg(s);
#pragma unroll
		for (int q = 0; q < 4; q++) {
			const float tv = in[q] ? __ldg(t + q) : 0.f;   // zero features outside the target image (valueOrAlt)
			if (OP == BOX_PROD) acc[q] = __fadd_rn(acc[q], __fmul_rn(sv, tv));
			else if (OP == BOX_SQ) { const float e = __fsub_rn(sv, tv); acc[q] = __fadd_rn(acc[q], __fmul_rn(e, e)); }
			else acc[q] = __fadd_rn(acc[q], fabsf(__fsub_rn(sv, tv)));
		}
		s += plane; t += plane;
	}
	float* o = p.out + (int64_t)i * p.o0 + (int64_t)j * p.o1 + (int64_t)a * p.o2 + (int64_t)b0 * p.o3;
#pragma unroll
	for (int q = 0; q < 4; q++) if (b0 + q < p.dW) o[q * p.o3] = acc[q];
}

// Rows of the planes that are multiples of 16 bytes (W % 4 == 0): a thread owns four consecutive source pixels and four
// consecutive column offsets whose first one keeps the target columns 16-byte aligned. The sixteen cells read the source
// pixels as one 16-byte load and the seven target columns they touch as two: 3 wide loads per feature for 16 cells
// (the four-candidate kernel: 5 scalar loads for 4). Summation per cell unchanged. off0 = the aligned offset of group 0.
template<int OP>
__global__ void __launch_bounds__(128) k_cv2d_planes_x4(const Cv2dParams p, int groups, int off0, int vec_out) {
	const int j = (blockIdx.x * 128 + threadIdx.x) * 4, i = blockIdx.z;
	if (j >= p.W) return;
	const int a = blockIdx.y / groups, off = off0 + (blockIdx.y - a * groups) * 4;
	const int it = i + a + p.lo0, jt = j + off;
	const bool row_in = it >= 0 && it < p.H;
	const bool in0 = row_in && jt >= 0 && jt + 4 <= p.W, in1 = row_in && jt + 4 >= 0 && jt + 8 <= p.W;
	const int64_t plane4 = (int64_t)p.H * p.W / 4;
	const float4* s = reinterpret_cast<const float4*>(p.S + (int64_t)i * p.W + j);
	const float4* t = reinterpret_cast<const float4*>(p.T + (int64_t)(row_in ? it : 0) * p.W + jt);
	float acc[4][4];
#pragma unroll
	for (int q = 0; q < 4; q++)
#pragma unroll
		for (int r = 0; r < 4; r++) acc[q][r] = 0.f;
	const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
	for (int f = 0; f < p.F; f++) {
		const float4 sv4 = __ldg(s);
		const float4 t0 = in0 ? __ldg(t) : zero, t1 = in1 ? __ldg(t + 1) : zero;   // zero features outside the target image
		const float sv[4] = {sv4.x, sv4.y, sv4.z, sv4.w};
		const float tv[8] = {t0.x, t0.y, t0.z, t0.w, t1.x, t1.y, t1.z, t1.w};
#pragma unroll
		for (int q = 0; q < 4; q++)
#pragma unroll
			for (int r = 0; r < 4; r++) {
				if (OP == BOX_PROD) acc[q][r] = __fadd_rn(acc[q][r], __fmul_rn(sv[q], tv[q + r]));
				else if (OP == BOX_SQ) { const float e = __fsub_rn(sv[q], tv[q + r]); acc[q][r] = __fadd_rn(acc[q][r], __fmul_rn(e, e)); }
				else acc[q][r] = __fadd_rn(acc[q][r], fabsf(__fsub_rn(sv[q], tv[q + r])));
			}
		s += plane4; t += plane4;
	}
	float* o = p.out + (int64_t)i * p.o0 + (int64_t)j * p.o1 + (int64_t)a * p.o2;
#pragma unroll
	for (int r = 0; r < 4; r++) {
		const int b = off + r - p.lo1;
		if (b < 0 || b >= p.dW) continue;
		float* ob = o + (int64_t)b * p.o3;
		if (vec_out) *reinterpret_cast<float4*>(ob) = make_float4(acc[0][r], acc[1][r], acc[2][r], acc[3][r]);
		else { ob[0] = acc[0][r]; ob[p.o1] = acc[1][r]; ob[2 * p.o1] = acc[2][r]; ob[3 * p.o1] = acc[3][r]; }
	}
}

} // namespace

int32_t PaddedPair::build(stevi_ctx* c, const View& l, const View& r, int hr, int vr, int D, int band_rows, int extra_rows) {
	if (l.ndim != 2 || r.ndim != 2) return fail(c, STEVI_E_UNSUPPORTED, "images must be 2-D (H,W)");
	if (l.dtype != r.dtype || (l.dtype != STEVI_U8 && l.dtype != STEVI_F32))
		return fail(c, STEVI_E_DTYPE, "images must both be u8 or both f32");
	if (l.shape[0] != r.shape[0]) return fail(c, STEVI_E_SHAPE, "image heights differ (cross_correlations.h:751)");
	H = l.shape[0];
	if (H > 65535) return fail(c, STEVI_E_UNSUPPORTED, "more than 65535 rows");
	const int Wmax = std::max(l.shape[1], r.shape[1]);
	pad_left = (D + hr + 48 + 7) / 8 * 8;
	pad_top = vr + 1 + extra_rows;
	pitch = (pad_left + (Wmax + 127) / 128 * 128 + pad_left + 64 + 63) / 64 * 64;
	half = (int)(pitch / 2);
	rows = pad_top + (H + band_rows) + vr + 1 + extra_rows;
	plane = (size_t)rows * pitch;
	float* buf = nullptr;
	STEVI_TRY(arena_alloc_t(c, plane * 6, &buf));
	STEVI_CUDA(c, cudaMemsetAsync(buf, 0, plane * 6 * sizeof(float), c->stream));
	const int64_t org = (int64_t)pad_top * pitch;
	for (int k = 0; k < 2; k++) {
		img[k] = buf + plane * (3 * k) + org;
		A[k] = buf + plane * (3 * k + 1) + org;
		B[k] = buf + plane * (3 * k + 2) + org;
		const View& v = k == 0 ? l : r;
		W[k] = v.shape[1];
		if (H == 0 || W[k] == 0) continue;
		dim3 grid(cdiv(W[k], 256), H);
		const size_t esz = v.dtype == STEVI_U8 ? 1 : 4;
		if (v.st[1] == 1 && (v.st[0] * esz) % (4 * esz) == 0 && (uintptr_t)v.p % (4 * esz) == 0) {
			dim3 gu(cdiv(W[k], 1024), H);
			if (v.dtype == STEVI_U8) STEVI_LAUNCH(c, k_pad_image_units<uint8_t>, gu, 256, 0, v.as<uint8_t>(), v.st[0], H, W[k], img[k], pitch, half, pad_left);
			else STEVI_LAUNCH(c, k_pad_image_units<float>, gu, 256, 0, v.as<float>(), v.st[0], H, W[k], img[k], pitch, half, pad_left);
			continue;
		}
		if (v.dtype == STEVI_U8) {
			STEVI_LAUNCH(c, k_pad_image<uint8_t>, grid, 256, 0, v.as<uint8_t>(), v.st[0], v.st[1], H, W[k], img[k], pitch, half, pad_left);
		} else {
			STEVI_LAUNCH(c, k_pad_image<float>, grid, 256, 0, v.as<float>(), v.st[0], v.st[1], H, W[k], img[k], pitch, half, pad_left);
		}
	}
	return STEVI_OK;
}

int32_t PaddedPair::stats(stevi_ctx* c, int which, int hr, int vr, int mf, int kind) {
	if (H == 0 || W[which] == 0) return STEVI_OK;
	StatsParams p;
	p.pimg = img[which]; p.pitch = pitch; p.half = half; p.pad_left = pad_left; p.H = H; p.W = W[which]; p.hr = hr; p.vr = vr;
	p.zero_mean = mf_zero_mean(mf); p.normalized = mf_normalized(mf); p.op = mf_op(mf); p.kind = kind;
	p.A = A[which]; p.B = B[which];
	if (hr >= 1 && hr <= 4) {
		dim3 grid(cdiv(p.W, 1024), H);
		switch (hr) {
		case 1: STEVI_LAUNCH(c, k_window_stats<1>, grid, 128, 0, p); break;
		case 2: STEVI_LAUNCH(c, k_window_stats<2>, grid, 128, 0, p); break;
		case 3: STEVI_LAUNCH(c, k_window_stats<3>, grid, 128, 0, p); break;
		default: STEVI_LAUNCH(c, k_window_stats<4>, grid, 128, 0, p); break;
		}
	} else {
		dim3 grid(cdiv(p.W, 128), H);
		STEVI_LAUNCH(c, k_window_stats_generic, grid, 128, 0, p);
	}
	return STEVI_OK;
}

// unfoldBased2dDisparityCostVolume (cross_correlations.h:794-822 over aggregateCost :308-371): cell (i, j, dh, dw) compares
// the source window at (i, j) with the target window at (i + dh + lower0, j + dw + lower1), zero features outside the
// target image, for both directions (the direction only picks the source image). One launch of the reference-order kernel
// per row offset; search = {lower0, upper0, lower1, upper1}; out (H, W, dH, dW).
int32_t float_cv2d_dev(stevi_ctx* c, int mf, const View& l, const View& r, int hr, int vr, const int32_t* search, int direction,
					   const View& out, bool exact) {
	switch (mf) {
	case STEVI_MF_CC: case STEVI_MF_NCC: case STEVI_MF_ZCC: case STEVI_MF_ZNCC:
	case STEVI_MF_SSD: case STEVI_MF_ZSSD: case STEVI_MF_SAD: case STEVI_MF_ZSAD: break;
	default: return fail(c, STEVI_E_UNSUPPORTED, "cost_volume_2d: matching function %d not available", mf);
	}
	if (!search) return fail(c, STEVI_E_ARG, "cost_volume_2d: null search window");
	if (hr < 0 || vr < 0 || hr > 127 || vr > 127) return fail(c, STEVI_E_ARG, "cost_volume_2d: radii out of range");
	if (l.ndim != 2 || r.ndim != 2 || l.dtype != STEVI_F32 || r.dtype != STEVI_F32)
		return fail(c, STEVI_E_DTYPE, "cost_volume_2d: images must be (H,W) f32");
	if (l.shape[0] != r.shape[0] || l.shape[1] != r.shape[1])
		return fail(c, STEVI_E_SHAPE, "cost_volume_2d: image sizes differ (cross_correlations.h:804-810)");
	const int dH = search[1] - search[0] + 1, dW = search[3] - search[2] + 1;
	if (dH <= 0 || dW <= 0) return fail(c, STEVI_E_SHAPE, "cost_volume_2d: empty search window (cross_correlations.h:336-338)");
	if (dH > 4096 || dW > 65535 || abs(search[0]) > 65535 || abs(search[1]) > 65535 || abs(search[2]) > 65535 || abs(search[3]) > 65535)
		return fail(c, STEVI_E_UNSUPPORTED, "cost_volume_2d: search window too large");
	const int H = l.shape[0], W = l.shape[1];
	if (out.ndim != 4 || out.dtype != STEVI_F32 || out.shape[0] != H || out.shape[1] != W || out.shape[2] != dH || out.shape[3] != dW)
		return fail(c, STEVI_E_SHAPE, "cost_volume_2d: out must be (%d,%d,%d,%d) f32", H, W, dH, dW);
	if (H == 0 || W == 0) return STEVI_OK;
	const bool r2l = direction == STEVI_RIGHT_TO_LEFT;
	const int reach = std::max(abs(search[2]), abs(search[3]));   // the padded rows must hold every target column a cell can name
	PaddedPair pp;
	const int si = r2l ? 1 : 0, ti = 1 - si;
	if (!exact && mf != STEVI_MF_ZSAD && hr >= 1 && hr <= 4 && BOX_BAND_ROWS + 2 * vr <= BOX_MAX_TILE_ROWS && dW <= 65535 - 8) {
		// One launch of the 1-D box kernel (§ costvol_box.cuh: O(1) per cell from running window sums) per row offset: row offset a of the
		// search is the 1-D volume against the target planes read a rows further down, column offsets lower1 .. upper1 are its
		// "disparities". Rows and columns outside the target image are zero image / zero statistics in the padded planes, which is
		// how the kernel recognises the reference's zero target vector. The de-interleaved planes can be shifted by whole pairs of
		// 16-byte units (8 columns) only: the first stored column offset is the d_first-th one the kernel evaluates.
		const int vreach = std::max(abs(search[0]), abs(search[1]));
		STEVI_TRY(pp.build(c, l, r, hr, vr, reach + 9, BOX_BAND_ROWS, vreach));
		const int op = mf_op(mf);
		STEVI_TRY(pp.stats(c, si, hr, vr, mf, op == BOX_PROD ? STATS_PROD_SRC : STATS_DIFF_SRC));
		STEVI_TRY(pp.stats(c, ti, hr, vr, mf, op == BOX_PROD ? STATS_PROD_TGT : STATS_DIFF_TGT));
		const int lo8 = search[2] >= 0 ? search[2] / 8 * 8 : -((-search[2] + 7) / 8 * 8);
		const int d_first = search[2] - lo8;
		const int64_t col_shift = (int64_t)(lo8 / 8) * 4;   // 8 columns = two units = 4 floats in the de-interleaved rows
		for (int a = 0; a < dH; a++) {
			const int64_t shift = (int64_t)(search[0] + a) * pp.pitch + col_shift;
			BoxParams p;
			p.ps = pp.img[si]; p.pt = pp.img[ti] + shift; p.As = pp.A[si]; p.Bs = pp.B[si]; p.At = pp.A[ti] + shift; p.Bt = pp.B[ti] + shift;
			p.pitch = pp.pitch; p.half = pp.half; p.pad_left = pp.pad_left; p.H = H; p.Ws = W; p.D = dW + d_first; p.vr = vr; p.R = BOX_BAND_ROWS;
			p.row_shift = ((c->row_origin % BOX_BAND_ROWS) + BOX_BAND_ROWS) % BOX_BAND_ROWS;
			p.d_first = d_first;
			p.Ff = (float)((2 * hr + 1) * (2 * vr + 1));
			p.out = out.as<float>() + (int64_t)a * out.st[2] - (int64_t)d_first * out.st[3];
			p.o0 = out.st[0]; p.o1 = out.st[1]; p.o2 = out.st[3];
			p.keys = nullptr;
			p.vec_store = out.st[1] == 1 && (uintptr_t)out.p % 16 == 0 && out.st[0] % 4 == 0 && out.st[2] % 4 == 0 && out.st[3] % 4 == 0;
			if (p.vec_store && (uintptr_t)out.p % 32 == 0 && out.st[0] % 8 == 0 && out.st[2] % 8 == 0 && out.st[3] % 8 == 0) p.vec_store = 2;
			switch (op) {
			case BOX_PROD: STEVI_TRY(launch_cv_box_prod(c, p, hr, 1)); break;
			case BOX_SQ: STEVI_TRY(launch_cv_box_sq(c, p, hr, 1)); break;
			default: STEVI_TRY(launch_cv_box_abs(c, p, hr, 1)); break;
			}
		}
		return STEVI_OK;
	}
	STEVI_TRY(pp.build(c, l, r, hr, vr, reach + 1, 1));
	STEVI_TRY(pp.stats(c, si, hr, vr, mf, STATS_RAW));
	STEVI_TRY(pp.stats(c, ti, hr, vr, mf, STATS_RAW));
	const int F = (2 * hr + 1) * (2 * vr + 1);
	const size_t plane_floats = (size_t)F * H * W;
	if ((int64_t)dH * ((dW + 3) / 4) <= 65535 && plane_floats * 8 <= ((size_t)24 << 30)) {
		// feature planes of both images, then the register-tiled comparison (same arithmetic as k_cv_direct, see above)
		float* planes[2] = {nullptr, nullptr};
		for (int k = 0; k < 2; k++) {
			STEVI_TRY(arena_alloc_t(c, plane_floats, &planes[k]));
			PlaneParams q;
			q.img = pp.img[k]; q.pitch = pp.pitch; q.half = pp.half; q.pad_left = pp.pad_left; q.M = pp.A[k]; q.N = pp.B[k];
			q.H = H; q.W = W; q.hr = hr; q.vr = vr; q.zero_mean = mf_zero_mean(mf); q.normalized = mf_normalized(mf); q.planes = planes[k];
			dim3 gp(cdiv(W, 128), H);
			STEVI_LAUNCH(c, k_feature_planes, gp, 128, 0, q);
		}
		Cv2dParams v;
		v.S = planes[si]; v.T = planes[ti]; v.H = H; v.W = W; v.F = F; v.dH = dH; v.dW = dW; v.lo0 = search[0]; v.lo1 = search[2];
		v.out = out.as<float>(); v.o0 = out.st[0]; v.o1 = out.st[1]; v.o2 = out.st[2]; v.o3 = out.st[3];
		if (W % 4 == 0) {
			// aligned groups of four column offsets covering [lower1, upper1]
			const int off0 = search[2] >= 0 ? search[2] / 4 * 4 : -((-search[2] + 3) / 4 * 4);
			const int groups = (search[3] - off0) / 4 + 1;
			if ((int64_t)dH * groups <= 65535) {
				const int vec_out = out.st[1] == 1 && (uintptr_t)out.p % 16 == 0 && out.st[0] % 4 == 0 && out.st[2] % 4 == 0 && out.st[3] % 4 == 0;
				dim3 g4(cdiv(W / 4, 128), dH * groups, H);
				switch (mf_op(mf)) {
				case BOX_PROD: STEVI_LAUNCH(c, k_cv2d_planes_x4<BOX_PROD>, g4, 128, 0, v, groups, off0, vec_out); break;
				case BOX_SQ: STEVI_LAUNCH(c, k_cv2d_planes_x4<BOX_SQ>, g4, 128, 0, v, groups, off0, vec_out); break;
				default: STEVI_LAUNCH(c, k_cv2d_planes_x4<BOX_ABS>, g4, 128, 0, v, groups, off0, vec_out); break;
				}
				return STEVI_OK;
			}
		}
		dim3 gv(cdiv(W, 128), dH * ((dW + 3) / 4), H);
		switch (mf_op(mf)) {
		case BOX_PROD: STEVI_LAUNCH(c, k_cv2d_planes<BOX_PROD>, gv, 128, 0, v); break;
		case BOX_SQ: STEVI_LAUNCH(c, k_cv2d_planes<BOX_SQ>, gv, 128, 0, v); break;
		default: STEVI_LAUNCH(c, k_cv2d_planes<BOX_ABS>, gv, 128, 0, v); break;
		}
		return STEVI_OK;
	}
	// windows or frames too large for the planes: one launch of the per-cell kernel per row offset
	DirectParams p;
	p.ps = pp.img[si]; p.pt = pp.img[ti]; p.pitch = pp.pitch; p.half = pp.half; p.pad_left = pp.pad_left;
	p.Ms = pp.A[si]; p.Ns = pp.B[si]; p.Mt = pp.A[ti]; p.Nt = pp.B[ti];
	p.H = H; p.Ws = W; p.Wt = W; p.D = dW; p.hr = hr; p.vr = vr; p.sign = 1; p.mf = mf;
	p.zero_mean = mf_zero_mean(mf); p.normalized = mf_normalized(mf);
	p.col_off = search[2];
	p.o0 = out.st[0]; p.o1 = out.st[1]; p.o2 = out.st[3];
	dim3 grid(cdiv(W, 128), dW, H);
	for (int a = 0; a < dH; a++) {
		p.dv = a + search[0];
		p.out = out.as<float>() + (int64_t)a * out.st[2];
		STEVI_LAUNCH(c, k_cv_direct, grid, 128, 0, p);
	}
	return STEVI_OK;
}

int32_t float_cv_dev(stevi_ctx* c, int mf, const View& l, const View& r, int hr, int vr, int D, int direction, const View& out,
					 bool force_direct, const View* wta_out) {
	switch (mf) {
	case STEVI_MF_CC: case STEVI_MF_NCC: case STEVI_MF_ZCC: case STEVI_MF_ZNCC:
	case STEVI_MF_SSD: case STEVI_MF_ZSSD: case STEVI_MF_SAD: case STEVI_MF_ZSAD: break;
	default: return fail(c, STEVI_E_UNSUPPORTED, "cost_volume: matching function %d not available", mf);
	}
	if (hr < 0 || vr < 0 || hr > 127 || vr > 127) return fail(c, STEVI_E_ARG, "cost_volume: radii out of range");
	if (D < 0) return fail(c, STEVI_E_ARG, "cost_volume: negative disparity width");
	if (l.ndim == 2 && r.ndim == 2 && l.dtype == STEVI_U8 && r.dtype == STEVI_U8 && (mf_zero_mean(mf) || mf_normalized(mf)))
		return fail(c, STEVI_E_UNSUPPORTED, "cost_volume: the reference's int16 fixed-point features for u8 images (cross_correlations.h:416-462) are not available; pass float images");
	const bool r2l = direction == STEVI_RIGHT_TO_LEFT;
	const bool direct = force_direct || mf == STEVI_MF_ZSAD || hr < 1 || hr > 4 || BOX_BAND_ROWS + 2 * vr > BOX_MAX_TILE_ROWS;
	const int R = direct ? 1 : BOX_BAND_ROWS;
	PaddedPair pp;
	STEVI_TRY(pp.build(c, l, r, hr, vr, D, R));
	const int si = r2l ? 1 : 0, ti = 1 - si;
	const int H = pp.H, Ws = pp.W[si], Wt = pp.W[ti];
	if (out.ndim != 3 || out.shape[0] != H || out.shape[1] != Ws || out.shape[2] != D || out.dtype != STEVI_F32)
		return fail(c, STEVI_E_SHAPE, "cost_volume: out must be (%d,%d,%d) f32", H, Ws, D);
	const int op = mf_op(mf);
	if (H == 0 || Ws == 0 || D == 0) {
		if (wta_out) STEVI_TRY(wta_dev(c, out, op == BOX_PROD ? STEVI_SCORE : STEVI_COST, *wta_out));   // whatever the staged call does with an empty volume
		return STEVI_OK;
	}
	if (direct) {
		if (D > 65535) return fail(c, STEVI_E_UNSUPPORTED, "cost_volume: more than 65535 disparities");
		STEVI_TRY(pp.stats(c, si, hr, vr, mf, STATS_RAW));
		STEVI_TRY(pp.stats(c, ti, hr, vr, mf, STATS_RAW));
		DirectParams p;
		p.ps = pp.img[si]; p.pt = pp.img[ti]; p.pitch = pp.pitch; p.half = pp.half; p.pad_left = pp.pad_left;
		p.Ms = pp.A[si]; p.Ns = pp.B[si]; p.Mt = pp.A[ti]; p.Nt = pp.B[ti];
		p.H = H; p.Ws = Ws; p.Wt = Wt; p.D = D; p.hr = hr; p.vr = vr; p.sign = r2l ? 1 : -1; p.mf = mf;
		p.zero_mean = mf_zero_mean(mf); p.normalized = mf_normalized(mf); p.dv = 0; p.col_off = 0;
		p.out = out.as<float>(); p.o0 = out.st[0]; p.o1 = out.st[1]; p.o2 = out.st[2];
		dim3 grid(cdiv(Ws, 128), D, H);
		STEVI_LAUNCH(c, k_cv_direct, grid, 128, 0, p);
		if (wta_out) STEVI_TRY(wta_dev(c, out, op == BOX_PROD ? STEVI_SCORE : STEVI_COST, *wta_out));
		return STEVI_OK;
	}
	STEVI_TRY(pp.stats(c, si, hr, vr, mf, op == BOX_PROD ? STATS_PROD_SRC : STATS_DIFF_SRC));
	STEVI_TRY(pp.stats(c, ti, hr, vr, mf, op == BOX_PROD ? STATS_PROD_TGT : STATS_DIFF_TGT));
	BoxParams p;
	p.ps = pp.img[si]; p.pt = pp.img[ti]; p.As = pp.A[si]; p.Bs = pp.B[si]; p.At = pp.A[ti]; p.Bt = pp.B[ti];
	p.pitch = pp.pitch; p.half = pp.half; p.pad_left = pp.pad_left; p.H = H; p.Ws = Ws; p.D = D; p.vr = vr; p.R = R; p.row_shift = ((c->row_origin % R) + R) % R;
	p.Ff = (float)((2 * hr + 1) * (2 * vr + 1));
	p.out = out.as<float>(); p.o0 = out.st[0]; p.o1 = out.st[1]; p.o2 = out.st[2];
	p.keys = nullptr;
	if (wta_out) {
		if (wta_out->ndim != 2 || wta_out->dtype != STEVI_I32 || wta_out->shape[0] != H || wta_out->shape[1] != Ws)
			return fail(c, STEVI_E_SHAPE, "cost_volume: winner map must be (%d,%d) i32", H, Ws);
		STEVI_TRY(arena_alloc_t(c, (size_t)H * Ws, &p.keys));
		STEVI_CUDA(c, cudaMemsetAsync(p.keys, 0, (size_t)H * Ws * sizeof(unsigned long long), c->stream));
	}
	p.vec_store = out.st[1] == 1 && (uintptr_t)out.p % 16 == 0 && out.st[0] % 4 == 0 && out.st[2] % 4 == 0;
	if (p.vec_store && (uintptr_t)out.p % 32 == 0 && out.st[0] % 8 == 0 && out.st[2] % 8 == 0) p.vec_store = 2;
	switch (op) {
	case BOX_PROD: STEVI_TRY(launch_cv_box_prod(c, p, hr, r2l ? 1 : -1)); break;
	case BOX_SQ: STEVI_TRY(launch_cv_box_sq(c, p, hr, r2l ? 1 : -1)); break;
	default: STEVI_TRY(launch_cv_box_abs(c, p, hr, r2l ? 1 : -1)); break;
	}
	if (wta_out) STEVI_TRY(wta_decode_max_keys_dev(c, p.keys, H, Ws, *wta_out));
	return STEVI_OK;
}

} // namespace stevi
