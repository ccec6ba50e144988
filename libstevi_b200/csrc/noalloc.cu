// stevi_census_cost_volume_noalloc: the census cost volume of a pair (censusTransform2D on both images, then the Hamming
// cost volume — what stevi_cost_volume(STEVI_MF_CENSUS, ..) does) for callers that may not allocate: code under CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc. It has no counterpart in the reference.
// The stage's scratch memory — the census words of the two images, which the regular entry
// point takes from the context's workspace arena (api.cu: census_cv_dev / alloc_census) — lives in a file-scope __device__
// array of STEVI_NOALLOC_SCRATCH_BYTES instead. Every GPU that loads the library holds that array until the process exits.
//
// The entry point calls no cudaMalloc / cudaFree / allocating routine: it does not touch the arena (no begin_call), takes
// device-resident operands only (host operands would need staging buffers) and refuses contexts with per-kernel timing on
// (that creates events). All calls share the array and it keeps the previous call's contents: the part a call uses is cleared
// first, and two calls that overlap on the GPU (different streams, no ordering between them) race on it.
#include "common.cuh"
#include "kernels.cuh"
#include <mutex>

namespace stevi {

namespace {

// In host code this name is only a stand-in for the device symbol: its address is NOT a device address. Host code gets the
// device address with cudaGetSymbolAddress (scratch_of_current_device); device code below names the array directly.
static __device__ __align__(256) unsigned char g_noalloc_scratch[STEVI_NOALLOC_SCRATCH_BYTES];

constexpr int MAX_DEVICES = 64;
std::mutex g_mu;
unsigned char* g_scratch_addr[MAX_DEVICES];   // per device, resolved once with that device current

int32_t scratch_of_current_device(stevi_ctx* c, unsigned char** out) {
	if (c->device < 0 || c->device >= MAX_DEVICES) return fail(c, STEVI_E_UNSUPPORTED, "noalloc: device ordinal %d", c->device);
	std::lock_guard<std::mutex> lock(g_mu);
	if (!g_scratch_addr[c->device]) {
		void* p = nullptr;
		STEVI_CUDA(c, cudaGetSymbolAddress(&p, g_noalloc_scratch));
		if (!p || (uintptr_t)p % 256) return fail(c, STEVI_E_CUDA, "noalloc: the scratch symbol resolved to %p", p);
		g_scratch_addr[c->device] = static_cast<unsigned char*>(p);
	}
	*out = g_scratch_addr[c->device];
	return STEVI_OK;
}

// zeroes the first n16 16-byte units of the scratch array (named directly: this is device code)
__global__ void __launch_bounds__(256) k_noalloc_clear(size_t n16) {
	uint4* s = reinterpret_cast<uint4*>(g_noalloc_scratch);
	for (size_t k = (size_t)blockIdx.x * 256 + threadIdx.x; k < n16; k += (size_t)gridDim.x * 256) s[k] = make_uint4(0u, 0u, 0u, 0u);
}

int32_t device_view(stevi_ctx* c, const stevi_array* a, View& v, const char* name) {
	if (!a || !a->data) return fail(c, STEVI_E_ARG, "%s: null array", name);
	if (a->mem != STEVI_DEVICE)
		return fail(c, STEVI_E_ARG, "%s: the no-allocation entry point takes device memory only (a host operand needs a staging buffer)", name);
	if (a->ndim < 1 || a->ndim > STEVI_MAX_DIMS) return fail(c, STEVI_E_ARG, "%s: ndim %d", name, a->ndim);
	if (dtype_size(a->dtype) == 0) return fail(c, STEVI_E_DTYPE, "%s: unknown dtype %d", name, a->dtype);
	v.p = a->data; v.dtype = a->dtype; v.ndim = a->ndim;
	for (int k = 0; k < a->ndim; k++) {
		if (a->shape[k] < 0) return fail(c, STEVI_E_SHAPE, "%s: negative extent", name);
		if (a->strides[k] < 0) return fail(c, STEVI_E_UNSUPPORTED, "%s: negative strides are not supported", name);
		v.shape[k] = a->shape[k]; v.st[k] = a->strides[k];
	}
	return STEVI_OK;
}

size_t round256(size_t b) { return (b + 255) / 256 * 256; }

} // namespace

} // namespace stevi

using namespace stevi;

extern "C" int32_t stevi_census_cost_volume_noalloc(stevi_ctx* c, const stevi_array* img_l, const stevi_array* img_r, int32_t hr, int32_t vr,
													int32_t D, int32_t direction, int32_t semantics, stevi_array* out) {
	if (!c) return STEVI_E_ARG;
	{
		cudaError_t e = cudaSetDevice(c->device);
		if (e != cudaSuccess) return fail(c, STEVI_E_CUDA, "cudaSetDevice(%d): %s", c->device, cudaGetErrorString(e));
	}
	if (c->timing) return fail(c, STEVI_E_UNSUPPORTED, "noalloc: per-kernel timing creates events; switch it off (stevi_ctx_set_timing) for this entry point");
	View l, r, o;
	STEVI_TRY(device_view(c, img_l, l, "img_l"));
	STEVI_TRY(device_view(c, img_r, r, "img_r"));
	STEVI_TRY(device_view(c, out, o, "out"));
	if (l.ndim != 2 || r.ndim != 2) return fail(c, STEVI_E_UNSUPPORTED, "images must be 2-D (H,W)");
	if (l.dtype != r.dtype || (l.dtype != STEVI_U8 && l.dtype != STEVI_F32)) return fail(c, STEVI_E_DTYPE, "images must both be u8 or both f32");
	if (l.shape[0] != r.shape[0]) return fail(c, STEVI_E_SHAPE, "image heights differ (cross_correlations.h:751)");
	if (hr < 0 || vr < 0 || hr > 127 || vr > 127) return fail(c, STEVI_E_ARG, "noalloc: window radii (%d,%d)", hr, vr);
	semantics &= 1;
	const int nw = census_n_words(hr, vr);
	const int H = l.shape[0];
	const size_t bytes_l = round256((size_t)H * l.shape[1] * nw * sizeof(uint32_t));
	const size_t bytes_r = round256((size_t)H * r.shape[1] * nw * sizeof(uint32_t));
	if (bytes_l + bytes_r > (size_t)STEVI_NOALLOC_SCRATCH_BYTES)
		return fail(c, STEVI_E_UNSUPPORTED, "noalloc: the census words of this pair need %zu bytes of scratch memory, the static array holds %zu "
					"(STEVI_NOALLOC_SCRATCH_BYTES): H * (Wl + Wr) * %d words * 4 bytes must fit; use stevi_cost_volume for larger pairs",
					bytes_l + bytes_r, (size_t)STEVI_NOALLOC_SCRATCH_BYTES, nw);
	unsigned char* scratch = nullptr;
	STEVI_TRY(scratch_of_current_device(c, &scratch));
	if (bytes_l + bytes_r) {
		const size_t n16 = (bytes_l + bytes_r) / 16;
		const int grid = (int)std::min<size_t>((n16 + 255) / 256, (size_t)c->sm_count * 8);
		STEVI_LAUNCH(c, k_noalloc_clear, grid, 256, 0, n16);   // the array keeps the last call's contents
	}
	// the word volumes in the layout of api.cu's alloc_census: (H,W,nw) with the word axis slowest
	View cl, cr;
	const View* src[2] = {&l, &r};
	View* dst[2] = {&cl, &cr};
	unsigned char* at[2] = {scratch, scratch + bytes_l};
	for (int k = 0; k < 2; k++) {
		View& v = *dst[k];
		v.p = at[k]; v.dtype = STEVI_U32; v.ndim = 3;
		v.shape[0] = H; v.shape[1] = src[k]->shape[1]; v.shape[2] = nw;
		v.st[0] = src[k]->shape[1]; v.st[1] = 1; v.st[2] = (int64_t)H * src[k]->shape[1];
	}
	const View imgs[2] = {l, r}, cens[2] = {cl, cr};
	STEVI_TRY(census_transform_pair_dev(c, imgs, cens, 2, hr, vr, semantics));
	return hamming_cv_dev(c, cl, cr, census_n_written(hr, vr, semantics), D, direction, semantics == STEVI_COMPAT, o, nullptr, 0);
}
