// Internal plumbing shared by the kernels' host wrappers: context, workspace arena, host-operand staging,
// launch bookkeeping. Nothing in here touches the CPU oracle; there is no CPU fallback anywhere in csrc/.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdarg>
#include <string>
#include <vector>
#include <algorithm>

#include "../../include/stevi_b200/c_api.h"

struct stevi_ctx {
	int device = 0;
	cudaStream_t stream = nullptr;
	bool own_stream = false;
	std::string err;
	int64_t launches = 0;
	int sm_count = 148;
	bool deferred_host = false; // stevi_ctx_set_deferred_host: calls with host operands do not wait for their copies
	int row_origin = 0; // global row of row 0 of the images handed to the next calls (stevi_ctx_set_row_origin)
	// grow-only bump arena; reset at the start of every API call (coalesced into one block when it had to grow)
	struct Block { char* p; size_t cap; };
	std::vector<Block> blocks;
	size_t cur = 0;   // index of the block being bumped
	size_t off = 0;   // offset inside it
	size_t high = 0;  // bytes handed out during the current call
	// optional per-kernel timing (stevi_ctx_set_timing): one CUDA event pair per launch on the ctx stream
	bool timing = false;
	struct Timed { const char* name; cudaEvent_t e0, e1; };
	std::vector<Timed> timed;
	std::vector<cudaEvent_t> event_pool;
	std::string report;
	// device-side fault flag (page-locked, mapped): a kernel that would otherwise wait forever on another block raises it
	// and carries on; checked whenever the host synchronises the stream
	int* fault_host = nullptr;
	int* fault_dev = nullptr;
	// strip-to-strip exchange buffer of the general SGM sweeps (sgm_sweep.cu): grow-only, holds {value, tag} pairs; the epoch
	// makes the tags of every launch unique
	void* sgm_xchg = nullptr;
	size_t sgm_xchg_bytes = 0;
	unsigned sgm_epoch = 0;
	// disparity maps going to host memory travel as the narrowest exact type (u8 / u16) and are widened into the caller's
	// int32 array by the host once the copy has landed (ctx.cu: Stager::out_index): tasks waiting for that, and the
	// page-locked staging buffers they use (a free list; a buffer is busy from its call to the synchronisation that widens it)
	// (round 2c: the widening is done by a small pool of worker threads; a task is handed to them by the first API call of the
	// process that finds its copy landed (an event after the copy, polled on entry to every call), at the latest by the
	// synchronising call, which only waits for tasks that are not finished yet)
	struct Widen { void* staged; int32_t* dst; size_t n; int bytes_per; int sign, offset; size_t slot; void* task; };
	std::vector<Widen> widen;
	struct Pinned { void* p; size_t cap; bool busy; };
	std::vector<Pinned> pinned;
	bool narrow_wire = true;
};

namespace stevi {

inline int32_t fail(stevi_ctx* c, int32_t code, const char* fmt, ...) {
	char buf[512];
	va_list ap;
	va_start(ap, fmt);
	vsnprintf(buf, sizeof buf, fmt, ap);
	va_end(ap);
	if (c) c->err = buf;
	return code;
}

// a failed runtime call leaves its code in the thread's "last error"; it is cleared here so that a recoverable failure (an
// allocation that did not fit, a launch configuration that was refused) does not fail every later call of this host thread
#define STEVI_CUDA(c, expr) do { cudaError_t e__ = (expr); if (e__ != cudaSuccess) { (void)cudaGetLastError(); \
	return stevi::fail((c), STEVI_E_CUDA, "%s:%d %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e__)); } } while (0)

#define STEVI_TRY(expr) do { int32_t s__ = (expr); if (s__ != STEVI_OK) return s__; } while (0)

// kernel<<<grid, block, smem, ctx->stream>>>(args...) with launch counting and error check
#define STEVI_LAUNCH(c, kern, grid, block, smem, ...) do { \
	if ((c)->timing) stevi::timing_begin((c), #kern); \
	kern<<<(grid), (block), (smem), (c)->stream>>>(__VA_ARGS__); \
	if ((c)->timing) stevi::timing_end((c)); \
	(c)->launches++; \
	STEVI_CUDA((c), cudaGetLastError()); } while (0)

void timing_begin(stevi_ctx* c, const char* name);
void timing_end(stevi_ctx* c);

inline size_t dtype_size(int dt) {
	switch (dt) {
	case STEVI_U8: return 1;
	case STEVI_U16: return 2;
	case STEVI_I32: case STEVI_U32: case STEVI_F32: return 4;
	default: return 0;
	}
}

// ---- arena ---------------------------------------------------------------------------------------------------
int32_t arena_reset(stevi_ctx* c);
int32_t arena_alloc(stevi_ctx* c, size_t bytes, void** out);
template<typename T> inline int32_t arena_alloc_t(stevi_ctx* c, size_t n, T** out) {
	void* p = nullptr;
	int32_t s = arena_alloc(c, n * sizeof(T), &p);
	*out = static_cast<T*>(p);
	return s;
}

// ---- device view of an operand ----------------------------------------------------------------------------------
struct View {
	void* p = nullptr;
	int dtype = 0, ndim = 0;
	int shape[STEVI_MAX_DIMS] = {0, 0, 0, 0};
	int64_t st[STEVI_MAX_DIMS] = {0, 0, 0, 0};
	template<typename T> T* as() const { return static_cast<T*>(p); }
	size_t count() const { size_t n = 1; for (int k = 0; k < ndim; k++) n *= (size_t)shape[k]; return n; }
	size_t span() const { // elements between the first and one past the last addressed element
		size_t s = 1;
		for (int k = 0; k < ndim; k++) { if (shape[k] == 0) return 0; s += (size_t)(shape[k] - 1) * (size_t)st[k]; }
		return s;
	}
	bool dense() const { return span() == count(); }
};

// Brings host operands to the device for the duration of one API call and results back at the end.
class Stager {
public:
	explicit Stager(stevi_ctx* c) : c_(c) {}
	int32_t in(const stevi_array* a, View& v, const char* name);
	int32_t out(const stevi_array* a, View& v, const char* name);
	// an (H,W) i32 result whose values are sign * k + offset with k in [0, n_values): a host array receives k as u8 / u16
	// and is widened by the host after the copy (4x / 2x fewer bytes over PCIe); anything else behaves like out()
	int32_t out_index(const stevi_array* a, View& v, int n_values, int sign, int offset, const char* name);
	// issues the D2H copies; synchronises the stream when any operand lived in host memory
	int32_t finish();
private:
	int32_t check(const stevi_array* a, const char* name);
	struct Back { void* host; void* dev; size_t bytes; };
	// final != nullptr: the caller's own array is u8 / u16 (device, or host: no widening), the narrowed map goes straight there
	struct Narrow { int32_t* host; const int32_t* dev; size_t n; int bytes_per; int sign, offset; void* final; int final_mem; };
	stevi_ctx* c_;
	std::vector<Back> back_;
	std::vector<Narrow> narrow_;
	bool any_host_ = false;
};

// narrow maps of ANY context of this device whose copies have landed go to the widening workers (ctx.cu); cheap when none is pending
void poll_landed_maps(int device);

inline int32_t begin_call(stevi_ctx* c) {
	if (!c) return STEVI_E_ARG;
	cudaError_t e = cudaSetDevice(c->device);
	if (e != cudaSuccess) return fail(c, STEVI_E_CUDA, "cudaSetDevice(%d): %s", c->device, cudaGetErrorString(e));
	poll_landed_maps(c->device);
	return arena_reset(c);
}

inline int cdiv(int a, int b) { return (a + b - 1) / b; }

// after a stream synchronisation: widen the narrow disparity maps that have landed (ctx.cu)
void finish_host(stevi_ctx* c);

// after a stream synchronisation: did a kernel raise the fault flag?
inline int32_t check_fault(stevi_ctx* c) {
	if (c->fault_host && *c->fault_host) {
		*c->fault_host = 0;
		return fail(c, STEVI_E_CUDA, "a kernel gave up waiting for another thread block (device-side fault flag): results of this stream are invalid");
	}
	return STEVI_OK;
}

} // namespace stevi
