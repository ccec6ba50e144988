// Census transform and Hamming cost volume (SURVEY.md §8(a) rows a1+a3 fused, a4+a5 for CENSUS/HAMMING).
//
// Reference behaviour reproduced here (file:line relative to the LibStevi tree):
//   unfold.h:248-291      window order c = (2*h_r+1)*k + l, zero outside the image
//   census.h:69-115       bit (c-1)%32 of word (c-1)/32 = (feat[0] > feat[c]); feat[0] is the TOP-LEFT window pixel;
//                         Q1: a word is written only when its 32 bits are full
//   cross_correlations.h:194-249  CV(i,j,d) = sum_n popcount(src(i,j,n) ^ tgt(i,j+s*d,n)), target 0 outside the image,
//                         Q2: every target word is rounded through float first (:235-236)
#include "common.cuh"
#include <type_traits>
#include "kernels.cuh"

namespace stevi {

// ---------------------------------------------------------------------------------------------------------------
// census transform: one thread per pixel, 32x8 pixel tile plus the (2*h_r, 2*v_r) apron staged in shared memory
// ---------------------------------------------------------------------------------------------------------------
struct CensusImage { const void* img; int64_t si, sj; int H, W; uint32_t* out; int64_t so0, so1, so2; };
struct CensusPair { CensusImage im[2]; };

// blockIdx.z selects the image, so both images of a pair go through one launch. HS > 0 fixes the window width at
// compile time (the inner loop unrolls and the shared-memory offsets become immediates); HS = 0 takes it from hr.
template<typename T, int HS>
__global__ void __launch_bounds__(256) k_census_transform(const CensusPair pr, int hr, int vr, int n_words, int n_written, int c_end) {
	extern __shared__ unsigned char smem_raw[];
	T* tile = reinterpret_cast<T*>(smem_raw);
#define STEVI_SEL(f) (blockIdx.z ? pr.im[1].f : pr.im[0].f) // no dynamic indexing of kernel parameters (that spills them)
	const T* __restrict__ img = static_cast<const T*>(STEVI_SEL(img));
	const int64_t si = STEVI_SEL(si), sj = STEVI_SEL(sj), so0 = STEVI_SEL(so0), so1 = STEVI_SEL(so1), so2 = STEVI_SEL(so2);
	const int H = STEVI_SEL(H), W = STEVI_SEL(W);
	uint32_t* __restrict__ out = STEVI_SEL(out);
#undef STEVI_SEL
	if ((int)blockIdx.x * 32 >= W || (int)blockIdx.y * 8 >= H) return;
	const int hs = HS > 0 ? HS : 2 * hr + 1, vs = 2 * vr + 1;
	const int tw = 32 + hs - 1, th = 8 + 2 * vr;
	const int j0 = blockIdx.x * 32, i0 = blockIdx.y * 8;
	for (int t = threadIdx.y * 32 + threadIdx.x; t < tw * th; t += 256) {
		int ty = t / tw, tx = t - ty * tw;
		int ii = i0 - vr + ty, jj = j0 - hr + tx;
		T v = T(0);
		if (ii >= 0 && ii < H && jj >= 0 && jj < W) v = img[ii * si + jj * sj];
		tile[t] = v;
	}
	__syncthreads();
	const int i = i0 + threadIdx.y, j = j0 + threadIdx.x;
	if (i >= H || j >= W) return;
	const T* base = tile + threadIdx.y * tw + threadIdx.x;
	const T ref = base[0];
	uint32_t* o = out + i * so0 + j * so1;
	uint32_t word = 0;
	int b = 0, ch = 0, c = 0; // c = hs*k + l
	for (int k = 0; k < vs && c < c_end; k++) { // comparisons past the last written word are never looked at
		const T* rowp = base + k * tw;
#pragma unroll
		for (int l = 0; l < hs; l++, c++) {
			if (c == 0 || c >= c_end) continue;
			word |= (ref > rowp[l] ? 1u : 0u) << b;
			b++;
			if (b == 32) { if (ch < n_written) o[ch * so2] = word; ch++; word = 0; b = 0; }
		}
	}
	// CORRECTED semantics flush the partial word (n_written == n_words); COMPAT leaves it zero
	if (b > 0 && ch < n_written) { o[ch * so2] = word; ch++; }
	for (int n = (ch < n_written ? ch : n_written); n < n_words; n++) o[n * so2] = 0u;
}

// Window fixed at compile time: the comparison index, its bit and its word are constants, so a comparison is a byte load,
// a compare and one predicated OR (the generic kernel above spends more on the bookkeeping than on the comparisons).
// C_END = first comparison index that is not looked at (COMPAT stops at the last full word, Q1).
template<typename T, int HS, int VS, int C_END>
__global__ void __launch_bounds__(256) k_census_window(const CensusPair pr) {
	constexpr int HR = HS / 2, VR = VS / 2;
	constexpr int TW = 32 + HS - 1, TH = 8 + VS - 1;
	constexpr int NW = (HS * VS - 1) / 32 + 1;          // words of the output
	constexpr int NWRITTEN = (C_END - 1 + 31) / 32;     // words that carry comparisons
	__shared__ T tile[TW * TH];
#define STEVI_SEL(f) (blockIdx.z ? pr.im[1].f : pr.im[0].f)
	const T* __restrict__ img = static_cast<const T*>(STEVI_SEL(img));
	const int64_t si = STEVI_SEL(si), sj = STEVI_SEL(sj), so0 = STEVI_SEL(so0), so1 = STEVI_SEL(so1), so2 = STEVI_SEL(so2);
	const int H = STEVI_SEL(H), W = STEVI_SEL(W);
	uint32_t* __restrict__ out = STEVI_SEL(out);
#undef STEVI_SEL
	if ((int)blockIdx.x * 32 >= W || (int)blockIdx.y * 8 >= H) return;
	const int j0 = blockIdx.x * 32, i0 = blockIdx.y * 8;
	for (int t = threadIdx.y * 32 + threadIdx.x; t < TW * TH; t += 256) {
		const int ty = t / TW, tx = t - ty * TW;
		const int ii = i0 - VR + ty, jj = j0 - HR + tx;
		T v = T(0);
		if (ii >= 0 && ii < H && jj >= 0 && jj < W) v = img[ii * si + jj * sj];
		tile[t] = v;
	}
	__syncthreads();
	const int i = i0 + threadIdx.y, j = j0 + threadIdx.x;
	if (i >= H || j >= W) return;
	const T* base = tile + threadIdx.y * TW + threadIdx.x;
	const T ref = base[0];
	uint32_t word[NW];
#pragma unroll
	for (int n = 0; n < NW; n++) word[n] = 0u;
#pragma unroll
	for (int k = 0; k < VS; k++) {
#pragma unroll
		for (int l = 0; l < HS; l++) {
			constexpr int dummy = 0; (void)dummy;
			const int c = HS * k + l;          // compile-time after unrolling
			if (c >= 1 && c < C_END) {
				if (ref > base[k * TW + l]) word[(c - 1) / 32] |= 1u << ((c - 1) % 32);
			}
		}
	}
	uint32_t* o = out + i * so0 + j * so1;
#pragma unroll
	for (int n = 0; n < NW; n++) o[n * so2] = n < NWRITTEN ? word[n] : 0u;
}

int32_t census_n_words(int hr, int vr) { return ((2 * hr + 1) * (2 * vr + 1) - 1) / 32 + 1; }
// number of words that can be non-zero
int32_t census_n_written(int hr, int vr, int semantics) {
	int ncmp = (2 * hr + 1) * (2 * vr + 1) - 1;
	return semantics == STEVI_COMPAT ? ncmp / 32 : (ncmp + 31) / 32;
}

static int32_t census_check(stevi_ctx* c, const View& img, int hr, int vr, const View& out) {
	if (img.ndim != 2) return fail(c, STEVI_E_UNSUPPORTED, "census: only 2-D images are supported");
	if (hr < 0 || vr < 0 || hr > 31 || vr > 31) return fail(c, STEVI_E_ARG, "census: radius out of range");
	const int F = (2 * hr + 1) * (2 * vr + 1);
	if (F <= 1) return fail(c, STEVI_E_SHAPE, "census: a single feature channel has no census (census.h:75-77)");
	const int nw = census_n_words(hr, vr);
	if (out.ndim != 3 || out.shape[0] != img.shape[0] || out.shape[1] != img.shape[1] || out.shape[2] != nw || out.dtype != STEVI_U32)
		return fail(c, STEVI_E_SHAPE, "census: out must be (H,W,%d) u32", nw);
	if (img.dtype != STEVI_U8 && img.dtype != STEVI_F32) return fail(c, STEVI_E_DTYPE, "census: image dtype must be u8 or f32");
	return STEVI_OK;
}

// one launch for up to two images of the same dtype
int32_t census_transform_pair_dev(stevi_ctx* c, const View* imgs, const View* outs, int n, int hr, int vr, int semantics) {
	CensusPair pr;
	int Hm = 0, Wm = 0;
	for (int k = 0; k < n; k++) {
		STEVI_TRY(census_check(c, imgs[k], hr, vr, outs[k]));
		if (imgs[k].dtype != imgs[0].dtype) return fail(c, STEVI_E_DTYPE, "census: images of a pair must share a dtype");
		pr.im[k] = CensusImage{imgs[k].p, imgs[k].st[0], imgs[k].st[1], imgs[k].shape[0], imgs[k].shape[1],
							   outs[k].as<uint32_t>(), outs[k].st[0], outs[k].st[1], outs[k].st[2]};
		Hm = std::max(Hm, imgs[k].shape[0]);
		Wm = std::max(Wm, imgs[k].shape[1]);
	}
	if (n == 1) pr.im[1] = pr.im[0];
	if (Hm == 0 || Wm == 0) return STEVI_OK;
	const int F = (2 * hr + 1) * (2 * vr + 1);
	const int nw = census_n_words(hr, vr), nwr = census_n_written(hr, vr, semantics);
	const int c_end = std::min(F, 1 + 32 * nwr);
	dim3 block(32, 8), grid(cdiv(Wm, 32), cdiv(Hm, 8), n);
	// common windows, both semantics: everything about the window is a compile-time constant
#define STEVI_CENSUS_WINDOW(T, HS, VS)                                                                                  \
	if (2 * hr + 1 == HS && 2 * vr + 1 == VS) {                                                                         \
		constexpr int ncmp = HS * VS - 1;                                                                               \
		if (semantics == STEVI_COMPAT) STEVI_LAUNCH(c, (k_census_window<T, HS, VS, 1 + 32 * (ncmp / 32)>), grid, block, 0, pr); \
		else STEVI_LAUNCH(c, (k_census_window<T, HS, VS, HS * VS>), grid, block, 0, pr);                                \
		return STEVI_OK;                                                                                                \
	}
#define STEVI_CENSUS_WINDOWS(T) \
	STEVI_CENSUS_WINDOW(T, 9, 7) STEVI_CENSUS_WINDOW(T, 9, 9) STEVI_CENSUS_WINDOW(T, 7, 7) STEVI_CENSUS_WINDOW(T, 5, 5) STEVI_CENSUS_WINDOW(T, 7, 5)
	if (imgs[0].dtype == STEVI_U8) { STEVI_CENSUS_WINDOWS(uint8_t) }
	else { STEVI_CENSUS_WINDOWS(float) }
#undef STEVI_CENSUS_WINDOWS
#undef STEVI_CENSUS_WINDOW
#define STEVI_CENSUS_LAUNCH(T, HS) STEVI_LAUNCH(c, (k_census_transform<T, HS>), grid, block, sm, pr, hr, vr, nw, nwr, c_end)
#define STEVI_CENSUS_SWITCH(T) \
	switch (2 * hr + 1) { \
	case 3: STEVI_CENSUS_LAUNCH(T, 3); break; \
	case 5: STEVI_CENSUS_LAUNCH(T, 5); break; \
	case 7: STEVI_CENSUS_LAUNCH(T, 7); break; \
	case 9: STEVI_CENSUS_LAUNCH(T, 9); break; \
	case 11: STEVI_CENSUS_LAUNCH(T, 11); break; \
	default: STEVI_CENSUS_LAUNCH(T, 0); break; \
	}
	if (imgs[0].dtype == STEVI_U8) {
		size_t sm = (size_t)(32 + 2 * hr) * (8 + 2 * vr) * sizeof(uint8_t);
		STEVI_CENSUS_SWITCH(uint8_t)
	} else {
		size_t sm = (size_t)(32 + 2 * hr) * (8 + 2 * vr) * sizeof(float);
		STEVI_CENSUS_SWITCH(float)
	}
#undef STEVI_CENSUS_SWITCH
#undef STEVI_CENSUS_LAUNCH
	return STEVI_OK;
}

int32_t census_transform_dev(stevi_ctx* c, const View& img, int hr, int vr, int semantics, const View& out) {
	return census_transform_pair_dev(c, &img, &out, 1, hr, vr, semantics);
}

} // namespace stevi

// ---------------------------------------------------------------------------------------------------------------
// Hamming cost volume
// ---------------------------------------------------------------------------------------------------------------
namespace stevi {

struct HamParams {
	const uint32_t* src; int64_t s0, s1, s2; // (row, column, word) strides of the source features
	const uint32_t* tgt; int64_t t0, t1, t2;
	int H, Ws, Wt, D, dchunk, n_words;
	void* out; int64_t o0, o1, o2;            // (row, column, disparity) strides of the cost volume
	int32_t* gmap; int pout;
	const int32_t* T; const int32_t* R; unsigned long long* keys;   // MODE 2: path sums in, winner keys out
	uint32_t* kin; uint32_t* kout;            // MODE 4: per pixel min over d of cost * 8192 + (8191 - d), d with j + d < W / >= W
	int round;                                // Q2: round target words through float
	int P;                                    // pitch of one of the 8 interleaved shared-memory lanes
};

// Q2 (cross_correlations.h:235-236): `float t = target; (uint32)t`. A word >= 0xFFFFFF80 rounds to 2^32 whose
// conversion back is undefined in C++; scalar x86-64 code yields 0, which is the documented behaviour here.
__device__ __forceinline__ uint32_t round_through_float(uint32_t w) {
	float f = __uint2float_rn(w);
	return f >= 4294967296.0f ? 0u : __float2uint_rz(f);
}

template<typename TO> __device__ __forceinline__ void store8(TO* p, const uint32_t (&c)[8]);
template<> __device__ __forceinline__ void store8<uint16_t>(uint16_t* p, const uint32_t (&c)[8]) {
	uint4 v;
	v.x = __byte_perm(c[0], c[1], 0x5410); v.y = __byte_perm(c[2], c[3], 0x5410); // one PRMT per pair of costs
	v.z = __byte_perm(c[4], c[5], 0x5410); v.w = __byte_perm(c[6], c[7], 0x5410);
	__stcs(reinterpret_cast<uint4*>(p), v);
}
template<> __device__ __forceinline__ void store8<int32_t>(int32_t* p, const uint32_t (&c)[8]) {
	__stcs(reinterpret_cast<uint4*>(p), make_uint4(c[0], c[1], c[2], c[3]));
	__stcs(reinterpret_cast<uint4*>(p) + 1, make_uint4(c[4], c[5], c[6], c[7]));
}
template<> __device__ __forceinline__ void store8<float>(float* p, const uint32_t (&c)[8]) {
	__stcs(reinterpret_cast<float4*>(p), make_float4((float)c[0], (float)c[1], (float)c[2], (float)c[3]));
	__stcs(reinterpret_cast<float4*>(p) + 1, make_float4((float)c[4], (float)c[5], (float)c[6], (float)c[7]));
}

// One block = one image row x 1024 source columns x one chunk of disparities. A thread owns 8 consecutive columns:
// their source words stay in registers, the 8 target words slide by one column per disparity (one shared-memory
// load per thread and disparity), and each disparity ends in one 16-byte store, so a warp writes 512 contiguous
// bytes of the [H][D][W] volume. Target words of the row are staged once in shared memory, already rounded (Q2),
// in 8 interleaved lanes so that the stride-8 accesses of a warp are bank-conflict free.
//
// MODE 0 stores the cost volume (and, with EMIT_G, the g map of the collapsed SGM on the way). The disparity-only
// census+SGM pipeline never has the volume in memory and runs the kernel twice instead, the costs recomputed from the
// census words both times: MODE 1 = g map only; MODE 2 = the aggregated costs S = (1+R)*C + R*pout*[j+d >= W] - T
// (sgm.cu, collapsed evaluation) and their running arg-min (`<=`: last minimum wins, correlation_base.h:446), merged
// over the disparity chunks through the same 64-bit keys as k_sgm_apply_rows. MODE 3 is MODE 2 with the running arg-min
// packed in one word, (S + T) * 8192 + (8191 - d), so that a cell costs one multiply-add and one min (the host picks it
// when |S + T| < 2^17; D <= 8192 on this path; T is constant along d and re-enters at the end).
// MODE 4 (round 2c) makes the second pass unnecessary: S is, per pixel, an increasing affine map of the cost plus one constant
// on the disparities with j + d >= W, so its arg-min under `<=` (last minimum wins) is decided by two numbers per pixel — the
// smallest cost, with its LARGEST disparity, among the d with j + d < W and among those with j + d >= W. One pass keeps both as
// cost * 8192 + (8191 - d) minima; g follows from them, and after the path sums a thread per pixel picks the winner
// (sgm.cu: k_keys_to_gmap, k_keys_to_disp).
template<int NW, int SIGN, typename TO, bool VEC, bool EMIT_G, int MODE = 0>
__global__ void __launch_bounds__(128) k_hamming_rows(const HamParams p) {
	extern __shared__ uint32_t sm[];
	const int tid = threadIdx.x;
	const int i = blockIdx.y;
	const int j0 = blockIdx.x * 1024;
	const int d0 = blockIdx.z * p.dchunk;          // p.dchunk is a multiple of 8
	const int dn = min(p.D - d0, p.dchunk);        // disparities of this block
	const int dn8 = (dn + 7) & ~7;                 // ... padded to whole groups of 8 (the pad is computed, never stored)
	const int N = 1024 + dn8 - 1;
	// target column held at staging index idx: base + idx; a thread's window for relative disparity dr covers
	// idx = 8*tid + k + dr (SIGN > 0) or 8*tid + k + (dn8 - 1 - dr) (SIGN < 0), k = 0..7
	const int base = SIGN > 0 ? j0 + d0 : j0 - d0 - (dn8 - 1);
	const int P = p.P;
	for (int idx = tid; idx < N; idx += 128) {
		const int col = base + idx;
		const bool ok = col >= 0 && col < p.Wt;
#pragma unroll
		for (int n = 0; n < NW; n++) {
			uint32_t w = ok ? __ldg(p.tgt + i * p.t0 + col * p.t1 + n * p.t2) : 0u;
			if (p.round) w = round_through_float(w);
			sm[n * 8 * P + (idx & 7) * P + (idx >> 3)] = w;
		}
	}
	__syncthreads();
	const int j = j0 + tid * 8;
	if (j >= p.Ws) return;
	uint32_t s[NW][8], w[NW][8];
#pragma unroll
	for (int k = 0; k < 8; k++) {
#pragma unroll
		for (int n = 0; n < NW; n++) s[n][k] = (j + k < p.Ws) ? __ldg(p.src + i * p.s0 + (j + k) * p.s1 + n * p.s2) : 0u;
	}
	// lane (= idx & 7) and group (= idx >> 3) of every window access are known up to a per-thread base, because
	// dn8 and the loop step are multiples of 8: the shared-memory addresses are base + immediate
	const int g8 = dn8 >> 3;
#pragma unroll
	for (int k = 0; k < 8; k++) {
#pragma unroll
		for (int n = 0; n < NW; n++) {
			if (SIGN > 0) w[n][k] = sm[n * 8 * P + k * P + tid];
			else w[n][k] = (k == 0) ? sm[n * 8 * P + 7 * P + tid + g8 - 1] : sm[n * 8 * P + (k - 1) * P + tid + g8];
		}
	}
	int gmin[8];
	if (EMIT_G) {
#pragma unroll
		for (int k = 0; k < 8; k++) gmin[k] = 0x7fffffff;
	}
	const bool full = VEC && (j + 8 <= p.Ws);
	TO* orow = reinterpret_cast<TO*>(p.out) + i * p.o0 + j * p.o1 + (int64_t)d0 * p.o2;
	// Pout only enters g where j + d reaches past the last column: threads that never get there (most of them) track
	// the plain minimum of the costs and double it at the end
	const bool edge = (EMIT_G || MODE >= 2) && (j + 7 + d0 + dn - 1 >= p.Ws);
	int r1[8], rp[8], tneg[8], best[8], bd[8];
	uint32_t kin[8], kout[8];
	if (MODE == 4) {
#pragma unroll
		for (int k = 0; k < 8; k++) kin[k] = kout[k] = 0xffffffffu;
	}
	if (MODE == 2 || MODE == 3) {
#pragma unroll
		for (int k = 0; k < 8; k++) {
			const bool in = j + k < p.Ws;
			const int rr = in ? __ldg(p.R + (int64_t)i * p.Ws + j + k) : 0;
			r1[k] = 1 + rr; rp[k] = rr * p.pout;
			tneg[k] = in ? -__ldg(p.T + (int64_t)i * p.Ws + j + k) : 0;
			best[k] = 0x7fffffff; bd[k] = 0;
			if (MODE == 3) { r1[k] *= 8192; rp[k] *= 8192; }
		}
	}
	auto sweep = [&](auto edge_tag, auto full_tag) {
		constexpr bool EDGE = decltype(edge_tag)::value, FULL = decltype(full_tag)::value;
		TO* o = orow;
		for (int gr = 0; gr < g8; gr++) { // 8 disparities per trip: dr = 8*gr + u
			const int live = dn - 8 * gr;  // < 8 only in a ragged last group
			const uint32_t* nxt = sm + tid + (SIGN > 0 ? gr + 1 : g8 - 1 - gr);
#pragma unroll
			for (int u = 0; u < 8; u++) {
				uint32_t cst[8];
#pragma unroll
				for (int k = 0; k < 8; k++) {
					uint32_t a = 0;
#pragma unroll
					for (int n = 0; n < NW; n++) a += __popc(s[n][k] ^ w[n][k]);
					cst[k] = a;
				}
				if (u < live) {
					if (EMIT_G) {
						const int d = d0 + 8 * gr + u;
#pragma unroll
						for (int k = 0; k < 8; k++) {
							if (EDGE) gmin[k] = min(gmin[k], (int)(2 * cst[k]) + ((j + k + d >= p.Ws) ? p.pout : 0));
							else gmin[k] = min(gmin[k], (int)cst[k]);
						}
					}
					if (MODE == 3) {
						const int d = d0 + 8 * gr + u, inv = 8191 - d;
#pragma unroll
						for (int k = 0; k < 8; k++) {
							int key = r1[k] * (int)cst[k] + inv;
							if (EDGE) key += (j + k + d >= p.Ws) ? rp[k] : 0;
							best[k] = min(best[k], key);
						}
					}
					if (MODE == 4) {
						const int d = d0 + 8 * gr + u;
						const uint32_t inv = 8191u - (uint32_t)d;
#pragma unroll
						for (int k = 0; k < 8; k++) {
							const uint32_t key = cst[k] * 8192u + inv;
							if (EDGE) { if (j + k + d >= p.Ws) kout[k] = min(kout[k], key); else kin[k] = min(kin[k], key); }
							else kin[k] = min(kin[k], key);
						}
					}
					if (MODE == 2) {
						const int d = d0 + 8 * gr + u;
#pragma unroll
						for (int k = 0; k < 8; k++) {
							int sc = r1[k] * (int)cst[k] + tneg[k];
							if (EDGE) sc += (j + k + d >= p.Ws) ? rp[k] : 0;
							if (sc <= best[k]) { best[k] = sc; bd[k] = d; }
						}
					}
					if (MODE == 0) {
						if (FULL) store8<TO>(o, cst);
						else {
#pragma unroll
							for (int k = 0; k < 8; k++) if (j + k < p.Ws) o[k * p.o1] = (TO)cst[k];
						}
					}
				}
				o += p.o2;
				if (gr + 1 < g8 || u < 7) { // slide the target window by one column
#pragma unroll
					for (int n = 0; n < NW; n++) {
						if (SIGN > 0) {
#pragma unroll
							for (int k = 0; k < 7; k++) w[n][k] = w[n][k + 1];
							w[n][7] = nxt[n * 8 * P + u * P];
						} else {
#pragma unroll
							for (int k = 7; k > 0; k--) w[n][k] = w[n][k - 1];
							w[n][0] = (u < 7) ? nxt[n * 8 * P + (6 - u) * P] : nxt[n * 8 * P + 7 * P - 1];
						}
					}
				}
			}
		}
	};
	if (edge) { if (full) sweep(std::true_type{}, std::true_type{}); else sweep(std::true_type{}, std::false_type{}); }
	else {
		if (full) sweep(std::false_type{}, std::true_type{}); else sweep(std::false_type{}, std::false_type{});
		if (EMIT_G) {
#pragma unroll
			for (int k = 0; k < 8; k++) gmin[k] *= 2;
		}
	}
	if (MODE == 3) {
#pragma unroll
		for (int k = 0; k < 8; k++) { bd[k] = 8191 - (best[k] & 8191); best[k] = (best[k] >> 13) + tneg[k]; }
	}
	if (MODE == 4) {
#pragma unroll
		for (int k = 0; k < 8; k++) if (j + k < p.Ws) {
			const int64_t o = ((int64_t)blockIdx.z * p.H + i) * p.Ws + j + k;   // one slice per disparity chunk
			p.kin[o] = kin[k]; p.kout[o] = kout[k];
		}
	}
	if (MODE == 2 || MODE == 3) {
#pragma unroll
		for (int k = 0; k < 8; k++) if (j + k < p.Ws) {
			const unsigned long long key = ((unsigned long long)((uint32_t)best[k] ^ 0x80000000u) << 32) | (0xffffffffu - (uint32_t)bd[k]);
			atomicMin(p.keys + (int64_t)i * p.Ws + j + k, key);   // the host pre-filled the keys with ones
		}
	}
	if (EMIT_G) {
#pragma unroll
		for (int k = 0; k < 8; k++) if (j + k < p.Ws) {
			if (gridDim.z > 1) atomicMin(p.gmap + (int64_t)i * p.Ws + j + k, gmin[k]); // the host pre-filled the map
			else p.gmap[(int64_t)i * p.Ws + j + k] = gmin[k];
		}
	}
}

// any strides, any word count: one thread per source pixel walking the disparities
template<typename TO>
__global__ void __launch_bounds__(256) k_hamming_generic(const HamParams p, int sign) {
	const int j = blockIdx.x * blockDim.x + threadIdx.x;
	const int i = blockIdx.y;
	if (j >= p.Ws) return;
	TO* o = reinterpret_cast<TO*>(p.out) + i * p.o0 + j * p.o1;
	int gmin = 0x7fffffff;
	for (int d = 0; d < p.D; d++) {
		const int jt = j + sign * d;
		const bool ok = jt >= 0 && jt < p.Wt;
		uint32_t a = 0;
		for (int n = 0; n < p.n_words; n++) {
			uint32_t t = ok ? p.tgt[i * p.t0 + jt * p.t1 + n * p.t2] : 0u;
			if (p.round) t = round_through_float(t);
			a += __popc(p.src[i * p.s0 + j * p.s1 + n * p.s2] ^ t);
		}
		o[d * p.o2] = (TO)a;
		gmin = min(gmin, (int)(2 * a) + ((j + d >= p.Ws) ? p.pout : 0));
	}
	if (p.gmap) p.gmap[(int64_t)i * p.Ws + j] = gmin;
}

template<int NW, int SIGN, typename TO>
static int32_t launch_rows(stevi_ctx* c, const HamParams& p, dim3 grid, size_t smem, bool vec) {
	if (p.gmap) {
		if (vec) STEVI_LAUNCH(c, (k_hamming_rows<NW, SIGN, TO, true, true>), grid, 128, smem, p);
		else STEVI_LAUNCH(c, (k_hamming_rows<NW, SIGN, TO, false, true>), grid, 128, smem, p);
	} else {
		if (vec) STEVI_LAUNCH(c, (k_hamming_rows<NW, SIGN, TO, true, false>), grid, 128, smem, p);
		else STEVI_LAUNCH(c, (k_hamming_rows<NW, SIGN, TO, false, false>), grid, 128, smem, p);
	}
	return STEVI_OK;
}

// grid and shared-memory layout of k_hamming_rows
static int32_t hamming_rows_plan(stevi_ctx* c, HamParams& p, dim3& grid, size_t& smem) {
	const int xb = cdiv(p.Ws, 1024);
	int dchunk = (p.D + 7) / 8 * 8;
	// many short blocks (>= 4 waves of 7 resident blocks per SM) so that the last partial wave costs little; more, shorter
	// chunks stage the target row more often and merge more minima (measured: 2 waves 0.212, 4 waves 0.203, 8 waves 0.216 ms at C2)
	const long want = 4L * 7 * c->sm_count;
	if ((long)xb * p.H < want) {
		long parts = (want + (long)xb * p.H - 1) / ((long)xb * p.H);
		int dc = (int)((p.D + parts - 1) / parts);
		dc = std::max(32, (dc + 7) / 8 * 8);
		dchunk = std::min(dchunk, dc);
	}
	// the staged target row (n_words x 8 lanes x P words) has to fit the 48 KB a kernel gets without opting in: long
	// disparity ranges with 3-4 census words are cut into more chunks instead
	auto plan = [&](int dc, int& P) {
		P = (1024 + dc + 8) / 8 + 2;
		while (P % 32 != 4) P++;
		return (size_t)p.n_words * 8 * P * sizeof(uint32_t);
	};
	int P = 0;
	while ((smem = plan(dchunk, P)) > 48 * 1024 && dchunk > 32) dchunk = std::max(32, (dchunk / 2 + 7) / 8 * 8);
	if (p.gmap && dchunk < p.D) // chunks merge their minima with atomicMin
		STEVI_CUDA(c, cudaMemsetAsync(p.gmap, 0x7f, (size_t)p.H * p.Ws * sizeof(int32_t), c->stream));
	p.dchunk = dchunk;
	p.P = P;
	grid = dim3(xb, p.H, cdiv(p.D, dchunk));
	return STEVI_OK;
}

template<typename TO>
static int32_t hamming_typed(stevi_ctx* c, HamParams p, int sign) {
	const bool rows_ok = p.o1 == 1 && p.n_words >= 1 && p.n_words <= 4 && p.D <= 8192;
	if (!rows_ok) {
		dim3 grid(cdiv(p.Ws, 256), p.H);
		STEVI_LAUNCH(c, k_hamming_generic<TO>, grid, 256, 0, p, sign);
		return STEVI_OK;
	}
	dim3 grid;
	size_t smem = 0;
	STEVI_TRY(hamming_rows_plan(c, p, grid, smem));
	const size_t esz = sizeof(TO);
	const bool vec = ((uintptr_t)p.out % 16 == 0) && ((p.o0 * esz) % 16 == 0) && ((p.o2 * esz) % 16 == 0);
#define STEVI_HAM_CASE(NW) \
	case NW: return sign > 0 ? launch_rows<NW, 1, TO>(c, p, grid, smem, vec) : launch_rows<NW, -1, TO>(c, p, grid, smem, vec);
	switch (p.n_words) {
		STEVI_HAM_CASE(1) STEVI_HAM_CASE(2) STEVI_HAM_CASE(3) STEVI_HAM_CASE(4)
	}
#undef STEVI_HAM_CASE
	return fail(c, STEVI_E_UNSUPPORTED, "hamming: unreachable");
}

static int32_t ham_params(stevi_ctx* c, const HamSource& h, HamParams& p) {
	const View& fl = h.fl; const View& fr = h.fr;
	if (fl.ndim != 3 || fr.ndim != 3 || fl.dtype != STEVI_U32 || fr.dtype != STEVI_U32)
		return fail(c, STEVI_E_DTYPE, "hamming: features must be (H,W,n) u32");
	if (fl.shape[0] != fr.shape[0] || fl.shape[2] != fr.shape[2])
		return fail(c, STEVI_E_SHAPE, "hamming: feature volumes disagree (cross_correlations.h:209-211)");
	const bool r2l = h.direction == STEVI_RIGHT_TO_LEFT;
	const View& src = r2l ? fr : fl;
	const View& tgt = r2l ? fl : fr;
	p.src = src.as<uint32_t>(); p.s0 = src.st[0]; p.s1 = src.st[1]; p.s2 = src.st[2];
	p.tgt = tgt.as<uint32_t>(); p.t0 = tgt.st[0]; p.t1 = tgt.st[1]; p.t2 = tgt.st[2];
	p.H = src.shape[0]; p.Ws = src.shape[1]; p.Wt = tgt.shape[1]; p.D = h.D; p.dchunk = h.D;
	p.n_words = std::min(h.n_cmp, src.shape[2]);
	p.out = nullptr; p.o0 = p.o2 = 0; p.o1 = 1;
	p.gmap = nullptr; p.pout = 0; p.T = p.R = nullptr; p.keys = nullptr; p.kin = p.kout = nullptr; p.round = h.round ? 1 : 0; p.P = 0;
	return STEVI_OK;
}

bool hamming_fused_capable(const HamSource& h) {
	const int nw = std::min(h.n_cmp, h.fl.shape[2]);
	return nw >= 1 && nw <= 4 && h.D >= 1 && h.D <= 8192 && h.fl.shape[0] >= 1 && h.fl.shape[0] <= 65535 && h.fl.shape[1] >= 1 && h.fr.shape[1] >= 1;
}

template<int MODE>
static int32_t hamming_fused_launch(stevi_ctx* c, HamParams& p, int sign) {
	dim3 grid;
	size_t smem = 0;
	STEVI_TRY(hamming_rows_plan(c, p, grid, smem));
#define STEVI_HAM_CASE(NW)                                                                                         \
	case NW:                                                                                                       \
		if (sign > 0) STEVI_LAUNCH(c, (k_hamming_rows<NW, 1, uint16_t, true, MODE == 1, MODE>), grid, 128, smem, p);  \
		else STEVI_LAUNCH(c, (k_hamming_rows<NW, -1, uint16_t, true, MODE == 1, MODE>), grid, 128, smem, p);          \
		return STEVI_OK;
	switch (p.n_words) {
		STEVI_HAM_CASE(1) STEVI_HAM_CASE(2) STEVI_HAM_CASE(3) STEVI_HAM_CASE(4)
	}
#undef STEVI_HAM_CASE
	return fail(c, STEVI_E_UNSUPPORTED, "hamming: unreachable");
}

int32_t hamming_gmap_dev(stevi_ctx* c, const HamSource& h, int32_t* gmap, int pout_i) {
	HamParams p;
	STEVI_TRY(ham_params(c, h, p));
	p.gmap = gmap; p.pout = pout_i;
	return hamming_fused_launch<1>(c, p, h.direction == STEVI_RIGHT_TO_LEFT ? 1 : -1);
}

int32_t hamming_minkeys_dev(stevi_ctx* c, const HamSource& h, uint32_t** kin, uint32_t** kout, int* nz) {
	HamParams p;
	STEVI_TRY(ham_params(c, h, p));
	dim3 grid;
	size_t smem = 0;
	STEVI_TRY(hamming_rows_plan(c, p, grid, smem));   // (no side effects: no map is attached yet)
	const size_t npix = (size_t)p.H * p.Ws;
	STEVI_TRY(arena_alloc_t(c, npix * grid.z, kin));
	STEVI_TRY(arena_alloc_t(c, npix * grid.z, kout));
	p.kin = *kin; p.kout = *kout;
	*nz = (int)grid.z;
	const int sign = h.direction == STEVI_RIGHT_TO_LEFT ? 1 : -1;
#define STEVI_HAM_CASE(NW)                                                                                   \
	case NW:                                                                                                 \
		if (sign > 0) STEVI_LAUNCH(c, (k_hamming_rows<NW, 1, uint16_t, true, false, 4>), grid, 128, smem, p);   \
		else STEVI_LAUNCH(c, (k_hamming_rows<NW, -1, uint16_t, true, false, 4>), grid, 128, smem, p);           \
		return STEVI_OK;
	switch (p.n_words) {
		STEVI_HAM_CASE(1) STEVI_HAM_CASE(2) STEVI_HAM_CASE(3) STEVI_HAM_CASE(4)
	}
#undef STEVI_HAM_CASE
	return fail(c, STEVI_E_UNSUPPORTED, "hamming: unreachable");
}

int32_t hamming_sgm_wta_dev(stevi_ctx* c, const HamSource& h, const int32_t* T, const int32_t* R, int pout_i, int r_max,
							unsigned long long* keys) {
	HamParams p;
	STEVI_TRY(ham_params(c, h, p));
	p.T = T; p.R = R; p.keys = keys; p.pout = pout_i;
	const int sign = h.direction == STEVI_RIGHT_TO_LEFT ? 1 : -1;
	// (1+R)*C + R*pout with R <= r_max passes and C <= 32 bits per word
	const long bound = (1L + r_max) * 32 * p.n_words + (long)r_max * labs((long)pout_i);
	if (bound < (1L << 17)) return hamming_fused_launch<3>(c, p, sign);
	return hamming_fused_launch<2>(c, p, sign);
}

int32_t hamming_cv_dev(stevi_ctx* c, const View& fl, const View& fr, int n_cmp, int D, int direction,
					   bool round_target, const View& out, int32_t* gmap, int pout_i) {
	if (fl.ndim != 3 || fr.ndim != 3 || fl.dtype != STEVI_U32 || fr.dtype != STEVI_U32)
		return fail(c, STEVI_E_DTYPE, "hamming: features must be (H,W,n) u32");
	if (fl.shape[0] != fr.shape[0] || fl.shape[2] != fr.shape[2])
		return fail(c, STEVI_E_SHAPE, "hamming: feature volumes disagree (cross_correlations.h:209-211)");
	if (D < 0) return fail(c, STEVI_E_ARG, "hamming: negative disparity width");
	const bool r2l = direction == STEVI_RIGHT_TO_LEFT;
	const View& src = r2l ? fr : fl;
	const View& tgt = r2l ? fl : fr;
	HamParams p;
	p.src = src.as<uint32_t>(); p.s0 = src.st[0]; p.s1 = src.st[1]; p.s2 = src.st[2];
	p.tgt = tgt.as<uint32_t>(); p.t0 = tgt.st[0]; p.t1 = tgt.st[1]; p.t2 = tgt.st[2];
	p.H = src.shape[0]; p.Ws = src.shape[1]; p.Wt = tgt.shape[1]; p.D = D; p.dchunk = D;
	p.n_words = std::min(n_cmp, src.shape[2]);
	p.out = out.p; p.o0 = out.st[0]; p.o1 = out.st[1]; p.o2 = out.st[2];
	p.gmap = gmap; p.pout = pout_i; p.round = round_target ? 1 : 0; p.P = 0;
	p.T = p.R = nullptr; p.keys = nullptr; p.kin = p.kout = nullptr;
	if (out.ndim != 3 || out.shape[0] != p.H || out.shape[1] != p.Ws || out.shape[2] != D)
		return fail(c, STEVI_E_SHAPE, "hamming: out must be (%d,%d,%d)", p.H, p.Ws, D);
	if (p.H == 0 || p.Ws == 0 || D == 0) return STEVI_OK;
	const int sign = r2l ? 1 : -1;
	switch (out.dtype) {
	case STEVI_U16: return hamming_typed<uint16_t>(c, p, sign);
	case STEVI_I32: return hamming_typed<int32_t>(c, p, sign);
	case STEVI_F32: return hamming_typed<float>(c, p, sign);
	default: return fail(c, STEVI_E_DTYPE, "hamming: cost volume dtype must be u16, i32 or f32");
	}
}

} // namespace stevi
