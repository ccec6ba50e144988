// On-demand cost evaluation and PatchMatch.
//
//   correlation/on_demand_cost_volume.h:35-330   GenericOnDemandCostVolume: costs of single (pixel, disparity) cells computed when asked
//                                                for, and its bulk call truncatedCostVolume(disp, radius) (:187-310)
//   correlation/patchmatch.h:166-556             patchMatchTestCost, patchMatchPropagate, patchMatchSearch, patchMatchImpl, patchMatch
//
// The reference keeps a lazily filled cost volume on the host and asks it for one cell at a time; on the GPU a cell is simply
// evaluated where it is needed — a comparison of two F-vectors of the processed feature volumes, summed in the reference's channel
// order — and nothing of the volume is ever stored.
//
// PatchMatch follows the reference's own structure, which is already line-parallel: a propagation pass scans every image row
// along the columns (rows independent of each other: a thread per row), then every column along the rows (a thread per column),
// each pixel testing the disparity of its predecessor and keeping it when it compares at least as well (`>=` / `<=`, :211-219);
// the four scan directions alternate over the iterations (:463-482); the random search (:233-369) is per pixel: nRandomSearch
// candidates drawn over the whole range and pulled towards the pixel's disparity by (k+1)/(nRandomSearch+1). With
// nRandomSearch = 0 and a given initial map the result is a deterministic function of the features and is compared bit for bit
// with the reference; the random draws come from a counter-based generator (the reference seeds from std::random_device, so its
// own runs differ from each other) and are compared statistically.
#include "kernels.cuh"

namespace stevi {
namespace {

enum { BOX_PROD = 0, BOX_SQ = 1, BOX_ABS = 2 };   // the comparison: products, squared differences, absolute differences

struct Pm {
	const float* fs; const float* ft;   // processed feature volumes, dense [H][W][Fp]: vectors padded with zeros to a multiple of 4
	int Hs, Ws, Ht, Wt, F, Fp;          // (16-byte loads; a zero pair adds an exact zero to every sum)
	int dims;                           // search dimensions: 1 (columns) or 2 (rows, columns)
	int lo0, up0, lo1, up1;             // dims == 1: lo1 / up1 only
	int32_t* disp; int64_t d0s, d1s, d2s;
	float* cur; unsigned char* cur_ok;  // cost of the pixel's disparity, and whether that disparity names a cell at all
	int op, score;
	int* changes; int* stop;
};

// on_demand_cost_volume.h:105-181: no value outside the search range or the target image
__device__ __forceinline__ bool pm_cost(const Pm& p, int i, int j, int d0, int d1, float& c) {
	if (p.dims == 2 && (d0 < p.lo0 || d0 > p.up0)) return false;
	if (d1 < p.lo1 || d1 > p.up1) return false;
	const int ti = i + (p.dims == 2 ? d0 : 0), tj = j + d1;
	if (ti < 0 || ti >= p.Ht || tj < 0 || tj >= p.Wt) return false;
	const float4* s = reinterpret_cast<const float4*>(p.fs + ((int64_t)i * p.Ws + j) * p.Fp);
	const float4* t = reinterpret_cast<const float4*>(p.ft + ((int64_t)ti * p.Wt + tj) * p.Fp);
	float acc = 0.f;
	// the reference's sequential sum over the channels (matching_costs.h:59-156). The loads of a chunk of 32 channels of both
	// vectors are issued together before any of them is used: a vector spans one or two 128-byte lines, and taken one load at a
	// time in a rolled loop each line's miss would be waited for in turn
	const int n4 = p.Fp / 4;
	for (int k0 = 0; k0 < n4; k0 += 8) {
		float4 a[8], b[8];
#pragma unroll
		for (int k = 0; k < 8; k++) if (k0 + k < n4) { a[k] = __ldg(s + k0 + k); b[k] = __ldg(t + k0 + k); }
#pragma unroll
		for (int k = 0; k < 8; k++) {
			if (k0 + k >= n4) break;
			const float sa[4] = {a[k].x, a[k].y, a[k].z, a[k].w}, tb[4] = {b[k].x, b[k].y, b[k].z, b[k].w};
#pragma unroll
			for (int q = 0; q < 4; q++) {
				if (p.op == BOX_PROD) acc = __fadd_rn(acc, __fmul_rn(sa[q], tb[q]));
				else if (p.op == BOX_SQ) { const float e = __fsub_rn(sa[q], tb[q]); acc = __fadd_rn(acc, __fmul_rn(e, e)); }
				else acc = __fadd_rn(acc, fabsf(__fsub_rn(sa[q], tb[q])));
			}
		}
	}
	c = acc;
	return true;
}

// patchmatch.h:166-231. Comparing with a disparity that names no cell is comparing with an empty std::optional: any value is
// `>=` it and none is `<=` it.
__device__ __forceinline__ int pm_test(const Pm& p, int i, int j, int d0, int d1) {
	float c;
	if (!pm_cost(p, i, j, d0, d1, c)) return 0;
	const int64_t px = (int64_t)i * p.Ws + j;
	const bool ok = p.cur_ok[px] != 0;
	const float cur = p.cur[px];
	const bool keep = p.score ? (!ok || c >= cur) : (ok && c <= cur);
	if (!keep) return 0;
	int32_t* d = p.disp + i * p.d0s + j * p.d1s;
	if (p.dims == 2) { d[0] = d0; d[p.d2s] = d1; } else d[0] = d1;
	p.cur[px] = c; p.cur_ok[px] = 1;
	return 1;
}

__device__ __forceinline__ void pm_read(const Pm& p, int i, int j, int& d0, int& d1) {
	const int32_t* d = p.disp + i * p.d0s + j * p.d1s;
	if (p.dims == 2) { d0 = d[0]; d1 = d[p.d2s]; } else { d0 = 0; d1 = d[0]; }
}

__global__ void __launch_bounds__(128) k_pm_current(const Pm p) {
	const int j = blockIdx.x * 128 + threadIdx.x, i = blockIdx.y;
	if (j >= p.Ws) return;
	int d0, d1;
	pm_read(p, i, j, d0, d1);
	float c = 0.f;
	const bool ok = pm_cost(p, i, j, d0, d1, c);
	p.cur[(int64_t)i * p.Ws + j] = c;
	p.cur_ok[(int64_t)i * p.Ws + j] = ok ? 1 : 0;
}

__device__ __forceinline__ unsigned long long pm_mix(unsigned long long z) {
	z += 0x9E3779B97F4A7C15ull;
	z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
	z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
	return z ^ (z >> 31);
}
// uniform integer in [lo, up]
__device__ __forceinline__ int pm_uniform(unsigned long long key, int lo, int up) {
	const unsigned long long range = (unsigned long long)(up - lo) + 1ull;
	return lo + (int)(pm_mix(key) % range);
}

// randomDispInit with a search range, patchmatch.h:57-160
__global__ void __launch_bounds__(128) k_pm_init(const Pm p, unsigned long long seed) {
	const int j = blockIdx.x * 128 + threadIdx.x, i = blockIdx.y;
	if (j >= p.Ws) return;
	const unsigned long long px = ((unsigned long long)i * p.Ws + j) * 2ull;
	int32_t* d = p.disp + i * p.d0s + j * p.d1s;
	if (p.dims == 2) { d[0] = pm_uniform(seed ^ pm_mix(px), p.lo0, p.up0); d[p.d2s] = pm_uniform(seed ^ pm_mix(px + 1), p.lo1, p.up1); }
	else d[0] = pm_uniform(seed ^ pm_mix(px), p.lo1, p.up1);
}

// pulls the feature vector that starts at q towards the SM (one request per 128-byte line)
__device__ __forceinline__ void pm_prefetch(const float* q, int F) {
	for (int k = 0; k < F; k += 32) asm volatile("prefetch.global.L1 [%0];" ::"l"(q + k));
	asm volatile("prefetch.global.L1 [%0];" ::"l"(q + F - 1));
}

// what a scan knows about a pixel before it reaches it: its disparity, the cost of that disparity and whether that cost exists
struct PmPixel { int d0, d1; float cur; bool ok; };
__device__ __forceinline__ PmPixel pm_load_pixel(const Pm& p, int i, int j) {
	PmPixel x;
	pm_read(p, i, j, x.d0, x.d1);
	const int64_t px = (int64_t)i * p.Ws + j;
	x.cur = p.cur[px]; x.ok = p.cur_ok[px] != 0;
	return x;
}

// One step of a scan: pixel (i, j), whose state `me` was loaded a step ago, tests the disparity (d0, d1) its predecessor ended
// with. The step is a dependent chain — the candidate's target vector can only be fetched once the predecessor is settled — so
// everything that does not depend on it is asked for a step ahead: the next pixel's own state (`next`, loaded here, used by the
// next step: a warp issues in order, a load consumed in the same step would expose its whole latency), the next pixel's source
// vector, and the target vector the next pixel will need if this one keeps the disparity it has (after the first sweeps the
// common case).
__device__ __forceinline__ int pm_scan_step(const Pm& p, int i, int j, int ni, int nj, const PmPixel& me, PmPixel& next, int& d0, int& d1) {
	const bool has_next = ni >= 0 && ni < p.Hs && nj >= 0 && nj < p.Ws;
	if (has_next) {
		next = pm_load_pixel(p, ni, nj);
		pm_prefetch(p.fs + ((int64_t)ni * p.Ws + nj) * p.Fp, p.Fp);
		const int ti = ni + (p.dims == 2 ? me.d0 : 0), tj = nj + me.d1;
		if (ti >= 0 && ti < p.Ht && tj >= 0 && tj < p.Wt) pm_prefetch(p.ft + ((int64_t)ti * p.Wt + tj) * p.Fp, p.Fp);
	}
	float c;
	bool keep = false;
	if (pm_cost(p, i, j, d0, d1, c)) keep = p.score ? (!me.ok || c >= me.cur) : (me.ok && c <= me.cur);   // patchmatch.h:211-219
	if (keep) {
		int32_t* d = p.disp + i * p.d0s + j * p.d1s;
		if (p.dims == 2) { d[0] = d0; d[p.d2s] = d1; } else d[0] = d1;
		const int64_t px = (int64_t)i * p.Ws + j;
		p.cur[px] = c; p.cur_ok[px] = 1;
		return 1;
	}
	d0 = me.d0; d1 = me.d1;   // the pixel keeps its own disparity: that is what its successor will test
	return 0;
}

// line scans of patchMatchPropagate, :391-415: a thread per row, along the columns in the direction inc
__global__ void __launch_bounds__(64) k_pm_scan_rows(const Pm p, int inc) {
	if (*p.stop) return;
	const int i = blockIdx.x * 64 + threadIdx.x;
	if (i >= p.Hs) return;
	int n = 0;
	// backward scans end BEFORE index 0: initialAndFinalPos<-1> gives {size - 1, 0} and the loop runs while j != final
	// (utils/propagation_direction.h:45-60)
	const int first = inc > 0 ? 0 : p.Ws - 1;
	int d0, d1;
	pm_read(p, i, first, d0, d1);   // the first pixel of the line has no predecessor and is only a source
	PmPixel me = {0, 0, 0.f, false}, next = me;
	if (first + inc >= 0 && first + inc < p.Ws) me = pm_load_pixel(p, i, first + inc);
	for (int j = first + inc; inc > 0 ? j < p.Ws : j > 0; j += inc) {
		n += pm_scan_step(p, i, j, i, j + inc, me, next, d0, d1);
		me = next;
	}
	if (n) atomicAdd(p.changes, n);
}

// column scans, :417-442: a thread per column, along the rows
__global__ void __launch_bounds__(64) k_pm_scan_cols(const Pm p, int inc) {
	if (*p.stop) return;
	const int j = blockIdx.x * 64 + threadIdx.x;
	if (j >= p.Ws) return;
	int n = 0;
	const int first = inc > 0 ? 0 : p.Hs - 1;
	int d0, d1;
	pm_read(p, first, j, d0, d1);
	PmPixel me = {0, 0, 0.f, false}, next = me;
	if (first + inc >= 0 && first + inc < p.Hs) me = pm_load_pixel(p, first + inc, j);
	for (int i = first + inc; inc > 0 ? i < p.Hs : i > 0; i += inc) {
		n += pm_scan_step(p, i, j, i + inc, j, me, next, d0, d1);
		me = next;
	}
	if (n) atomicAdd(p.changes, n);
}

// patchMatchSearch, :233-369. Like the reference, the count of the LAST candidate is what the pixel reports (n_chang is
// overwritten per candidate, :356).
__global__ void __launch_bounds__(128) k_pm_search(const Pm p, int n_random, unsigned long long seed) {
	if (*p.stop) return;
	const int j = blockIdx.x * 128 + threadIdx.x, i = blockIdx.y;
	if (j >= p.Ws) return;
	int b0, b1;
	pm_read(p, i, j, b0, b1);
	const unsigned long long px = ((unsigned long long)i * p.Ws + j) * (unsigned long long)(2 * n_random);
	int n = 0;
	for (int k = 0; k < n_random; k++) {
		int d0 = 0, d1;
		if (p.dims == 2) { d0 = pm_uniform(seed ^ pm_mix(px + 2 * k), p.lo0, p.up0); d1 = pm_uniform(seed ^ pm_mix(px + 2 * k + 1), p.lo1, p.up1); }
		else d1 = pm_uniform(seed ^ pm_mix(px + 2 * k), p.lo1, p.up1);
		int e0 = d0 - b0, e1 = d1 - b1;
		e1 = e1 * (k + 1) / (n_random + 1);
		if (p.dims == 2) e0 = e0 * (k + 1) / (n_random + 1);
		d0 = b0 + e0; d1 = b1 + e1;
		if (p.dims == 1) { if (d1 == b1) d1 = b1 + 1; }
		else if (d0 == b0 && d1 == b1) { d0 = b0 + 1; d1 = b1 + 1; }
		n = pm_test(p, i, j, d0, d1);
	}
	if (n) atomicAdd(p.changes, n);
}

__global__ void __launch_bounds__(128) k_pm_copy(const int32_t* __restrict__ src, int64_t s0, int64_t s1, int64_t s2, const Pm p) {
	const int j = blockIdx.x * 128 + threadIdx.x, i = blockIdx.y;
	if (j >= p.Ws) return;
	for (int k = 0; k < p.dims; k++) p.disp[i * p.d0s + j * p.d1s + k * p.d2s] = src[i * s0 + j * s1 + k * s2];
}

// end of an iteration: no change at all ends the run (patchMatchImpl, :489-491)
__global__ void k_pm_end_iteration(int* changes, int* stop) {
	if (*changes == 0) *stop = 1;
	*changes = 0;
}

// GenericOnDemandCostVolume::truncatedCostVolume, on_demand_cost_volume.h:187-310. The reference turns the disparity into a
// cost-volume index, moves by the cell offset and hands that INDEX to costValue as a disparity (:224-241): with a search range
// that does not start at 0 the window is shifted by the lower offset. Reproduced (COMPAT).
struct OdTcv {
	Pm p;
	float* out; int64_t o0, o1, o2, o3;
	int radius; float missing;
};
__global__ void __launch_bounds__(128) k_od_tcv(const OdTcv q) {
	const Pm& p = q.p;
	const int j = blockIdx.x * 128 + threadIdx.x, i = blockIdx.y;
	if (j >= p.Ws) return;
	const int n = 2 * q.radius + 1;
	int d0, d1;
	pm_read(p, i, j, d0, d1);
	float* o = q.out + i * q.o0 + j * q.o1;
	if (p.dims == 1) {
		for (int k = 0; k < n; k++) {
			float c;
			o[k * q.o2] = pm_cost(p, i, j, 0, (d1 - p.lo1) + k - q.radius, c) ? c : q.missing;
		}
	} else {
		for (int a = 0; a < n; a++) for (int b = 0; b < n; b++) {
			float c;
			o[a * q.o2 + b * q.o3] = pm_cost(p, i, j, (d0 - p.lo0) + a - q.radius, (d1 - p.lo1) + b - q.radius, c) ? c : q.missing;
		}
	}
}

inline int op_of(int mf) {
	switch (mf) {
	case STEVI_MF_CC: case STEVI_MF_NCC: case STEVI_MF_ZCC: case STEVI_MF_ZNCC: return BOX_PROD;
	case STEVI_MF_SSD: case STEVI_MF_ZSSD: return BOX_SQ;
	default: return BOX_ABS;
	}
}

// processed features of both volumes, dense; the reference processes them in patchMatch (:524-525) and again in the constructor
// of the on-demand volume when the function is normalised or zero-mean (on_demand_cost_volume.h:63-68): `passes` = 2 reproduces it
int32_t prepare(stevi_ctx* c, const char* who, int mf, const View& fs, const View& ft, int dims, const int32_t* search, int passes, Pm& p) {
	if (mf < STEVI_MF_CC || mf > STEVI_MF_ZSAD) return fail(c, STEVI_E_UNSUPPORTED, "%s: matching function %d", who, mf);
	if (fs.ndim != 3 || ft.ndim != 3 || fs.dtype != STEVI_F32 || ft.dtype != STEVI_F32) return fail(c, STEVI_E_DTYPE, "%s: feature volumes must be (H,W,F) f32", who);
	if (dims != 1 && dims != 2) return fail(c, STEVI_E_ARG, "%s: the search has 1 or 2 dimensions", who);
	if (!search) return fail(c, STEVI_E_ARG, "%s: search range missing", who);
	// patchmatch.h:531-539: empty array
	if (fs.shape[2] != ft.shape[2] || (dims == 1 && fs.shape[0] != ft.shape[0])) return fail(c, STEVI_E_SHAPE, "%s: the feature volumes do not match", who);
	if (fs.shape[0] > 65535) return fail(c, STEVI_E_UNSUPPORTED, "%s: more than 65535 rows", who);
	p = Pm{};
	p.Hs = fs.shape[0]; p.Ws = fs.shape[1]; p.Ht = ft.shape[0]; p.Wt = ft.shape[1]; p.F = fs.shape[2]; p.Fp = (p.F + 3) / 4 * 4;
	p.dims = dims;
	if (dims == 1) { p.lo1 = search[0]; p.up1 = search[1]; }
	else { p.lo0 = search[0]; p.up0 = search[1]; p.lo1 = search[2]; p.up1 = search[3]; }
	p.op = op_of(mf);
	p.score = p.op == BOX_PROD;
	const bool processed = mf == STEVI_MF_NCC || mf == STEVI_MF_ZNCC || mf == STEVI_MF_ZCC || mf == STEVI_MF_ZSSD || mf == STEVI_MF_ZSAD;
	const View* in[2] = {&fs, &ft};
	const float* res[2];
	for (int k = 0; k < 2; k++) {
		View cur = *in[k];
		const int n = processed ? passes : 0;
		for (int pass = 0; pass < std::max(n, 1); pass++) {
			View dst = cur;
			float* buf = nullptr;
			const size_t padded = (size_t)cur.shape[0] * cur.shape[1] * p.Fp;
			STEVI_TRY(arena_alloc_t(c, padded, &buf));
			if (p.Fp != p.F) STEVI_CUDA(c, cudaMemsetAsync(buf, 0, padded * sizeof(float), c->stream));
			dst.p = buf; dst.st[0] = (int64_t)cur.shape[1] * p.Fp; dst.st[1] = p.Fp; dst.st[2] = 1;
			// CC / SSD / SAD: features_for_matchfunc is the identity, used here as the copy into a dense volume
			STEVI_TRY(features_for_matchfunc_dev(c, n ? mf : STEVI_MF_CC, cur, dst));
			cur = dst;
		}
		res[k] = cur.as<float>();
	}
	p.fs = res[0]; p.ft = res[1];
	return STEVI_OK;
}

} // namespace

int32_t patch_match_dev(stevi_ctx* c, int mf, const View& fs, const View& ft, int dims, const int32_t* search, int n_iter, int n_random,
						unsigned long long seed, const View* init, const View& disp) {
	Pm p;
	STEVI_TRY(prepare(c, "patch_match", mf, fs, ft, dims, search, 2, p));
	if (n_iter < 0 || n_random < 0) return fail(c, STEVI_E_ARG, "patch_match: negative iteration counts");
	if (disp.ndim != 3 || disp.dtype != STEVI_I32 || disp.shape[0] != p.Hs || disp.shape[1] != p.Ws || disp.shape[2] != dims)
		return fail(c, STEVI_E_SHAPE, "patch_match: disp must be (%d,%d,%d) i32", p.Hs, p.Ws, dims);
	if (init && (init->ndim != 3 || init->dtype != STEVI_I32 || init->shape[0] != p.Hs || init->shape[1] != p.Ws || init->shape[2] != dims))
		return fail(c, STEVI_E_SHAPE, "patch_match: the initial map must be (%d,%d,%d) i32", p.Hs, p.Ws, dims);
	if (p.Hs == 0 || p.Ws == 0) return STEVI_OK;
	p.disp = disp.as<int32_t>(); p.d0s = disp.st[0]; p.d1s = disp.st[1]; p.d2s = disp.st[2];
	STEVI_TRY(arena_alloc_t(c, (size_t)p.Hs * p.Ws, &p.cur));
	STEVI_TRY(arena_alloc_t(c, (size_t)p.Hs * p.Ws, &p.cur_ok));
	int* flags = nullptr;
	STEVI_TRY(arena_alloc_t(c, 2, &flags));
	STEVI_CUDA(c, cudaMemsetAsync(flags, 0, 2 * sizeof(int), c->stream));
	p.changes = flags; p.stop = flags + 1;
	dim3 gpix(cdiv(p.Ws, 128), p.Hs);
	if (init) STEVI_LAUNCH(c, k_pm_copy, gpix, 128, 0, init->as<int32_t>(), init->st[0], init->st[1], init->st[2], p);
	else STEVI_LAUNCH(c, k_pm_init, gpix, 128, 0, p, seed);
	STEVI_LAUNCH(c, k_pm_current, gpix, 128, 0, p);
	static const int inc[4][2] = {{1, 1}, {1, -1}, {-1, 1}, {-1, -1}};   // utils/propagation_direction.h:65-86, patchmatch.h:463-482
	for (int it = 0; it < n_iter; it++) {
		STEVI_LAUNCH(c, k_pm_scan_rows, cdiv(p.Hs, 64), 64, 0, p, inc[it % 4][1]);
		STEVI_LAUNCH(c, k_pm_scan_cols, cdiv(p.Ws, 64), 64, 0, p, inc[it % 4][0]);
		if (n_random > 0) STEVI_LAUNCH(c, k_pm_search, gpix, 128, 0, p, n_random, seed + 0x632BE59BD9B4E019ull * (unsigned long long)(it + 1));
		STEVI_LAUNCH(c, k_pm_end_iteration, 1, 1, 0, p.changes, p.stop);
	}
	return STEVI_OK;
}

int32_t on_demand_tcv_dev(stevi_ctx* c, int mf, const View& fs, const View& ft, int dims, const int32_t* search, const View& disp, int radius,
						  const View& out) {
	OdTcv q;
	STEVI_TRY(prepare(c, "on_demand_truncated_cost_volume", mf, fs, ft, dims, search, 1, q.p));
	Pm& p = q.p;
	if (radius < 0 || radius > 64) return fail(c, STEVI_E_ARG, "on_demand_truncated_cost_volume: radius");
	const int n = 2 * radius + 1;
	// 1-D search: the reference's overload takes the (H,W) map (on_demand_cost_volume.h:253-310); 2-D: (H,W,2) (:187-251)
	const bool d_ok = disp.dtype == STEVI_I32 && disp.shape[0] == p.Hs && disp.shape[1] == p.Ws &&
					  ((dims == 1 && (disp.ndim == 2 || (disp.ndim == 3 && disp.shape[2] == 1))) || (dims == 2 && disp.ndim == 3 && disp.shape[2] == 2));
	if (!d_ok) return fail(c, STEVI_E_SHAPE, "on_demand_truncated_cost_volume: disp must be (%d,%d%s) i32", p.Hs, p.Ws, dims == 2 ? ",2" : "");
	const bool o_ok = out.dtype == STEVI_F32 && out.ndim == 2 + dims && out.shape[0] == p.Hs && out.shape[1] == p.Ws && out.shape[2] == n && (dims == 1 || out.shape[3] == n);
	if (!o_ok) return fail(c, STEVI_E_SHAPE, "on_demand_truncated_cost_volume: out must be (%d,%d,%d%s) f32", p.Hs, p.Ws, n, dims == 2 ? ",2r+1" : "");
	if (p.Hs == 0 || p.Ws == 0) return STEVI_OK;
	p.disp = disp.as<int32_t>(); p.d0s = disp.st[0]; p.d1s = disp.st[1]; p.d2s = disp.ndim == 3 ? disp.st[2] : 0;
	q.out = out.as<float>(); q.o0 = out.st[0]; q.o1 = out.st[1]; q.o2 = out.st[2]; q.o3 = dims == 2 ? out.st[3] : 0;
	q.radius = radius;
	// defaultCvValForMatchFunc, matching_costs.h:706-713: numeric_limits<float>::max() for costs, ::min() (the smallest POSITIVE
	// normal, as the reference writes it) for scores
	q.missing = p.score ? 1.17549435e-38f : 3.402823466e+38f;
	dim3 gpix(cdiv(p.Ws, 128), p.Hs);
	STEVI_LAUNCH(c, k_od_tcv, gpix, 128, 0, q);
	return STEVI_OK;
}

} // namespace stevi
