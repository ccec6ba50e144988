// Winner-take-all disparity extraction and index -> disparity conversion.
//
//   correlation_base.h:427-464  extractSelectedIndex: best = vol(i,j,0); for d = 1..D-1: Cost `<=`, Score `>=`
//                               => the LAST extremum wins (Q5); NaN never replaces, a NaN at d = 0 sticks
//   correlation_base.h:511-532  selectedIndexToDisp: disp = deltaSign*idx + offset, deltaSign = +1 for RightToLeft
#include "common.cuh"
#include "kernels.cuh"

namespace stevi {

template<typename T, bool SCORE> __device__ __forceinline__ bool wins(T v, T best) { return SCORE ? (v >= best) : (v <= best); }

template<typename T, int V> struct VecOf;
template<> struct VecOf<float, 4> { typedef float4 type; };
template<> struct VecOf<int32_t, 4> { typedef int4 type; };
template<> struct VecOf<uint16_t, 8> { typedef uint4 type; };

template<typename T, int V> __device__ __forceinline__ void unpack(const typename VecOf<T, V>::type& r, T (&v)[V]);
template<> __device__ __forceinline__ void unpack<float, 4>(const float4& r, float (&v)[4]) { v[0] = r.x; v[1] = r.y; v[2] = r.z; v[3] = r.w; }
template<> __device__ __forceinline__ void unpack<int32_t, 4>(const int4& r, int32_t (&v)[4]) { v[0] = r.x; v[1] = r.y; v[2] = r.z; v[3] = r.w; }
template<> __device__ __forceinline__ void unpack<uint16_t, 8>(const uint4& r, uint16_t (&v)[8]) {
	v[0] = r.x & 0xffff; v[1] = r.x >> 16; v[2] = r.y & 0xffff; v[3] = r.y >> 16;
	v[4] = r.z & 0xffff; v[5] = r.z >> 16; v[6] = r.w & 0xffff; v[7] = r.w >> 16;
}

// Volume with a pixel axis of stride 1 (the reference's [H][D][W] cost volumes, default-stride SGM volumes): a thread
// owns V consecutive pixels (one 16-byte load per disparity plane) and walks d exactly like the reference loop.
// axes: a = the other pixel axis (blockIdx.y), b = the stride-1 pixel axis.
template<typename T, int V, bool SCORE>
__global__ void __launch_bounds__(128) k_wta_planes(const T* __restrict__ vol, int64_t sa, int64_t sd, int na, int nb, int D,
													int32_t* __restrict__ out, int64_t oa, int64_t ob) {
	typedef typename VecOf<T, V>::type vec_t;
	const int b0 = (blockIdx.x * 128 + threadIdx.x) * V;
	const int a = blockIdx.y;
	if (b0 >= nb) return;
	const T* p = vol + a * sa + b0;
	T best[V];
	int bd[V];
	if (b0 + V <= nb) {
		unpack<T, V>(__ldcs(reinterpret_cast<const vec_t*>(p)), best);
#pragma unroll
		for (int k = 0; k < V; k++) bd[k] = 0;
#pragma unroll 8
		for (int d = 1; d < D; d++) {
			T v[V];
			unpack<T, V>(__ldcs(reinterpret_cast<const vec_t*>(p + d * sd)), v);
#pragma unroll
			for (int k = 0; k < V; k++) if (wins<T, SCORE>(v[k], best[k])) { best[k] = v[k]; bd[k] = d; }
		}
#pragma unroll
		for (int k = 0; k < V; k++) out[a * oa + (b0 + k) * ob] = bd[k];
	} else {
		for (int k = 0; b0 + k < nb; k++) {
			T bst = p[k];
			int bdd = 0;
			for (int d = 1; d < D; d++) { T v = p[k + d * sd]; if (wins<T, SCORE>(v, bst)) { bst = v; bdd = d; } }
			out[a * oa + (b0 + k) * ob] = bdd;
		}
	}
}

// any strides: one thread per pixel
template<typename T, bool SCORE>
__global__ void __launch_bounds__(128) k_wta_generic(const T* __restrict__ vol, int64_t s0, int64_t s1, int64_t sd, int H, int W, int D,
													 int32_t* __restrict__ out, int64_t o0, int64_t o1) {
	const int j = blockIdx.x * 128 + threadIdx.x;
	const int i = blockIdx.y;
	if (j >= W) return;
	const T* p = vol + i * s0 + j * s1;
	T best = p[0];
	int bd = 0;
#pragma unroll 4
	for (int d = 1; d < D; d++) { T v = p[d * sd]; if (wins<T, SCORE>(v, best)) { best = v; bd = d; } }
	out[i * o0 + j * o1] = bd;
}

__device__ __forceinline__ float shfl_xor_t(float v, int o) { return __shfl_xor_sync(0xffffffffu, v, o); }
__device__ __forceinline__ int32_t shfl_xor_t(int32_t v, int o) { return __shfl_xor_sync(0xffffffffu, v, o); }
__device__ __forceinline__ uint16_t shfl_xor_t(uint16_t v, int o) { return (uint16_t)__shfl_xor_sync(0xffffffffu, (int)v, o); }

// disparity axis of stride 1 (dense C-order volumes): one warp per pixel, lanes stride over d; the warp reduction
// keeps the reference's tie rule (equal values -> larger d) and its NaN rule (a NaN at d = 0 sticks).
template<typename T, bool SCORE>
__global__ void __launch_bounds__(256) k_wta_dfast(const T* __restrict__ vol, int64_t s0, int64_t s1, int H, int W, int D,
												   int32_t* __restrict__ out, int64_t o0, int64_t o1) {
	const int lane = threadIdx.x & 31;
	const int64_t pix = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
	if (pix >= (int64_t)H * W) return;
	const int i = (int)(pix / W), j = (int)(pix - (int64_t)i * W);
	const T* p = vol + i * s0 + j * s1;
	bool have = false;
	T best = T(0);
	int bd = -1;
	for (int d = lane; d < D; d += 32) {
		T v = p[d];
		if (v != v) continue; // NaN never wins past d = 0
		if (!have || wins<T, SCORE>(v, best)) { best = v; bd = d; have = true; }
	}
#pragma unroll
	for (int o = 16; o > 0; o >>= 1) {
		T ov = shfl_xor_t(best, o);
		int od = __shfl_xor_sync(0xffffffffu, bd, o);
		bool oh = od >= 0;
		bool take = oh && (!have || (ov == best ? od > bd : wins<T, SCORE>(ov, best)));
		if (take) { best = ov; bd = od; have = true; }
	}
	if (lane == 0) {
		T first = p[0];
		out[i * o0 + j * o1] = (first != first || bd < 0) ? 0 : bd;
	}
}

template<typename T, int V, bool SCORE>
static int32_t wta_typed(stevi_ctx* c, const View& vol, const View& out) {
	const int H = vol.shape[0], W = vol.shape[1], D = vol.shape[2];
	const T* p = vol.as<T>();
	int32_t* o = out.as<int32_t>();
	const size_t esz = sizeof(T);
	if (D > 1 && (vol.st[1] == 1 || vol.st[0] == 1)) {
		const int b = vol.st[1] == 1 ? 1 : 0, a = 1 - b;
		const bool aligned = ((uintptr_t)p % 16 == 0) && ((vol.st[a] * esz) % 16 == 0) && ((vol.st[2] * esz) % 16 == 0);
		if (aligned && vol.shape[a] <= 65535) {
			dim3 grid(cdiv(vol.shape[b], 128 * V), vol.shape[a]);
			STEVI_LAUNCH(c, (k_wta_planes<T, V, SCORE>), grid, 128, 0, p, vol.st[a], vol.st[2], vol.shape[a], vol.shape[b], D,
						 o, out.st[a], out.st[b]);
			return STEVI_OK;
		}
	}
	if (vol.st[2] == 1 && D >= 32) {
		const int64_t npix = (int64_t)H * W;
		STEVI_LAUNCH(c, (k_wta_dfast<T, SCORE>), (unsigned)((npix + 7) / 8), 256, 0, p, vol.st[0], vol.st[1], H, W, D, o, out.st[0], out.st[1]);
		return STEVI_OK;
	}
	dim3 grid(cdiv(W, 128), H);
	STEVI_LAUNCH(c, (k_wta_generic<T, SCORE>), grid, 128, 0, p, vol.st[0], vol.st[1], vol.st[2], H, W, D, o, out.st[0], out.st[1]);
	return STEVI_OK;
}

int32_t wta_dev(stevi_ctx* c, const View& vol, int strategy, const View& out) {
	if (vol.ndim != 3) return fail(c, STEVI_E_SHAPE, "wta: volume must be (H,W,D)");
	if (out.ndim != 2 || out.shape[0] != vol.shape[0] || out.shape[1] != vol.shape[1] || out.dtype != STEVI_I32)
		return fail(c, STEVI_E_SHAPE, "wta: out must be (H,W) i32");
	if (vol.shape[0] == 0 || vol.shape[1] == 0) return STEVI_OK;
	if (vol.shape[2] < 1) return fail(c, STEVI_E_SHAPE, "wta: empty disparity axis");
	if (vol.shape[0] > 65535) return fail(c, STEVI_E_UNSUPPORTED, "wta: more than 65535 rows");
	const bool score = strategy == STEVI_SCORE;
	switch (vol.dtype) {
	case STEVI_F32: return score ? wta_typed<float, 4, true>(c, vol, out) : wta_typed<float, 4, false>(c, vol, out);
	case STEVI_I32: return score ? wta_typed<int32_t, 4, true>(c, vol, out) : wta_typed<int32_t, 4, false>(c, vol, out);
	case STEVI_U16: return score ? wta_typed<uint16_t, 8, true>(c, vol, out) : wta_typed<uint16_t, 8, false>(c, vol, out);
	default: return fail(c, STEVI_E_DTYPE, "wta: volume dtype must be u16, i32 or f32");
	}
}

// extractSelected2dIndex (correlation_base.h:466-508): a thread per pixel scans (d1, d2) in row-major order starting from
// the value of cell (0,0), `<=` / `>=`: the last extremum wins, a NaN at (0,0) sticks (no comparison against it is true)
template<bool SCORE>
__global__ void __launch_bounds__(128) k_wta_2d(const float* __restrict__ v, int64_t s0, int64_t s1, int64_t s2, int64_t s3, int H, int W,
												int D1, int D2, int32_t* __restrict__ out, int64_t o0, int64_t o1, int64_t o2) {
	const int j = blockIdx.x * 128 + threadIdx.x, i = blockIdx.y;
	if (j >= W) return;
	const float* p = v + i * s0 + j * s1;
	float best = p[0];
	int b1 = 0, b2 = 0;
	for (int a = 0; a < D1; a++)
		for (int b = 0; b < D2; b++) {
			const float x = p[a * s2 + b * s3];
			if (SCORE ? (x >= best) : (x <= best)) { best = x; b1 = a; b2 = b; }
		}
	out[i * o0 + j * o1] = b1;
	out[i * o0 + j * o1 + o2] = b2;
}

int32_t wta2d_dev(stevi_ctx* c, const View& vol, int strategy, const View& out) {
	if (vol.ndim != 4 || vol.dtype != STEVI_F32) return fail(c, STEVI_E_SHAPE, "wta_2d: volume must be (H,W,D1,D2) f32");
	if (out.ndim != 3 || out.shape[0] != vol.shape[0] || out.shape[1] != vol.shape[1] || out.shape[2] != 2 || out.dtype != STEVI_I32)
		return fail(c, STEVI_E_SHAPE, "wta_2d: out must be (H,W,2) i32");
	if (strategy != STEVI_COST && strategy != STEVI_SCORE) return fail(c, STEVI_E_ARG, "wta_2d: bad strategy");
	const int H = vol.shape[0], W = vol.shape[1];
	if (H == 0 || W == 0) return STEVI_OK;
	if (vol.shape[2] < 1 || vol.shape[3] < 1) return fail(c, STEVI_E_SHAPE, "wta_2d: empty search axis");
	if (H > 65535) return fail(c, STEVI_E_UNSUPPORTED, "wta_2d: more than 65535 rows");
	dim3 grid(cdiv(W, 128), H);
	if (strategy == STEVI_SCORE)
		STEVI_LAUNCH(c, k_wta_2d<true>, grid, 128, 0, vol.as<float>(), vol.st[0], vol.st[1], vol.st[2], vol.st[3], H, W, vol.shape[2], vol.shape[3],
					 out.as<int32_t>(), out.st[0], out.st[1], out.st[2]);
	else
		STEVI_LAUNCH(c, k_wta_2d<false>, grid, 128, 0, vol.as<float>(), vol.st[0], vol.st[1], vol.st[2], vol.st[3], H, W, vol.shape[2], vol.shape[3],
					 out.as<int32_t>(), out.st[0], out.st[1], out.st[2]);
	return STEVI_OK;
}

// selected2dIndexToDisp (correlation_base.h:534-555): index + lower offset of its axis
__global__ void __launch_bounds__(256) k_index2d_to_disp(const int32_t* __restrict__ idx, int64_t s0, int64_t s1, int64_t s2, int H, int W, int lo0, int lo1,
														  int32_t* __restrict__ out, int64_t o0, int64_t o1, int64_t o2) {
	const int j = blockIdx.x * 256 + threadIdx.x, i = blockIdx.y;
	if (j >= W) return;
	out[i * o0 + j * o1] = idx[i * s0 + j * s1] + lo0;
	out[i * o0 + j * o1 + o2] = idx[i * s0 + j * s1 + s2] + lo1;
}

int32_t index2d_to_disp_dev(stevi_ctx* c, const View& idx, const int32_t* search, const View& out) {
	if (!search) return fail(c, STEVI_E_ARG, "index2d_to_disp: null search window");
	if (idx.ndim != 3 || idx.shape[2] != 2 || idx.dtype != STEVI_I32 || out.ndim != 3 || out.dtype != STEVI_I32 ||
		out.shape[0] != idx.shape[0] || out.shape[1] != idx.shape[1] || out.shape[2] != 2)
		return fail(c, STEVI_E_SHAPE, "index2d_to_disp: idx and out must be (H,W,2) i32");
	const int H = idx.shape[0], W = idx.shape[1];
	if (H == 0 || W == 0) return STEVI_OK;
	if (H > 65535) return fail(c, STEVI_E_UNSUPPORTED, "index2d_to_disp: more than 65535 rows");
	dim3 grid(cdiv(W, 256), H);
	STEVI_LAUNCH(c, k_index2d_to_disp, grid, 256, 0, idx.as<int32_t>(), idx.st[0], idx.st[1], idx.st[2], H, W, search[0], search[2],
				 out.as<int32_t>(), out.st[0], out.st[1], out.st[2]);
	return STEVI_OK;
}

__global__ void __launch_bounds__(256) k_wta_decode_max_keys(const unsigned long long* __restrict__ keys, int H, int W,
															 int32_t* __restrict__ out, int64_t o0, int64_t o1) {
	const int j = blockIdx.x * 256 + threadIdx.x, i = blockIdx.y;
	if (j >= W) return;
	out[i * o0 + j * o1] = (int32_t)(unsigned)(keys[(int64_t)i * W + j] & 0xffffffffull);
}

int32_t wta_decode_max_keys_dev(stevi_ctx* c, const unsigned long long* keys, int H, int W, const View& out) {
	if (H == 0 || W == 0) return STEVI_OK;
	dim3 grid(cdiv(W, 256), H);
	STEVI_LAUNCH(c, k_wta_decode_max_keys, grid, 256, 0, keys, H, W, out.as<int32_t>(), out.st[0], out.st[1]);
	return STEVI_OK;
}

__global__ void k_index_to_disp(const int32_t* __restrict__ idx, int64_t s0, int64_t s1, int H, int W, int sign, int offset,
								int32_t* __restrict__ out, int64_t o0, int64_t o1) {
	const int j = blockIdx.x * blockDim.x + threadIdx.x;
	const int i = blockIdx.y;
	if (j >= W) return;
	out[i * o0 + j * o1] = sign * idx[i * s0 + j * s1] + offset;
}

int32_t index_to_disp_dev(stevi_ctx* c, const View& idx, int direction, int offset, const View& out) {
	if (idx.ndim != 2 || out.ndim != 2 || idx.dtype != STEVI_I32 || out.dtype != STEVI_I32 ||
		idx.shape[0] != out.shape[0] || idx.shape[1] != out.shape[1])
		return fail(c, STEVI_E_SHAPE, "index_to_disp: (H,W) i32 in and out");
	if (idx.shape[0] == 0 || idx.shape[1] == 0) return STEVI_OK;
	if (idx.shape[0] > 65535) return fail(c, STEVI_E_UNSUPPORTED, "index_to_disp: more than 65535 rows");
	dim3 grid(cdiv(idx.shape[1], 256), idx.shape[0]);
	STEVI_LAUNCH(c, k_index_to_disp, grid, 256, 0, idx.as<int32_t>(), idx.st[0], idx.st[1], idx.shape[0], idx.shape[1],
				 direction == STEVI_RIGHT_TO_LEFT ? 1 : -1, offset, out.as<int32_t>(), out.st[0], out.st[1]);
	return STEVI_OK;
}

} // namespace stevi
