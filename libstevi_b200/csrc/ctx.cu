// Context, workspace arena and host-operand staging for the C ABI (include/stevi_b200/c_api.h).
#include "common.cuh"
#include <immintrin.h>
#include <atomic>
#include <condition_variable>
#include <deque>
#include <mutex>
#include <thread>

namespace stevi {

static const size_t ARENA_ALIGN = 256;

int32_t arena_reset(stevi_ctx* c) {
	if (c->blocks.size() > 1) {
		// the previous call outgrew the arena: replace the pieces by one block that fits all of it
		size_t total = 0;
		for (auto& b : c->blocks) total += b.cap;
		STEVI_CUDA(c, cudaStreamSynchronize(c->stream));
		for (auto& b : c->blocks) cudaFree(b.p);
		c->blocks.clear();
		char* p = nullptr;
		STEVI_CUDA(c, cudaMalloc(&p, total));
		c->blocks.push_back({p, total});
	}
	c->cur = 0;
	c->off = 0;
	c->high = 0;
	return STEVI_OK;
}

int32_t arena_alloc(stevi_ctx* c, size_t bytes, void** out) {
	bytes = (bytes + ARENA_ALIGN - 1) / ARENA_ALIGN * ARENA_ALIGN;
	if (bytes == 0) bytes = ARENA_ALIGN;
	while (c->cur < c->blocks.size()) {
		auto& b = c->blocks[c->cur];
		if (c->off + bytes <= b.cap) {
			*out = b.p + c->off;
			c->off += bytes;
			c->high += bytes;
			return STEVI_OK;
		}
		c->cur++;
		c->off = 0;
	}
	size_t cap = std::max(bytes, (size_t)(64u << 20));
	char* p = nullptr;
	STEVI_CUDA(c, cudaMalloc(&p, cap));
	c->blocks.push_back({p, cap});
	c->cur = c->blocks.size() - 1;
	c->off = bytes;
	c->high += bytes;
	*out = p;
	return STEVI_OK;
}

static cudaEvent_t take_event(stevi_ctx* c) {
	cudaEvent_t e = nullptr;
	if (!c->event_pool.empty()) { e = c->event_pool.back(); c->event_pool.pop_back(); }
	else cudaEventCreate(&e);
	return e;
}

void timing_begin(stevi_ctx* c, const char* name) {
	stevi_ctx::Timed t{name, take_event(c), take_event(c)};
	cudaEventRecord(t.e0, c->stream);
	c->timed.push_back(t);
}

void timing_end(stevi_ctx* c) { cudaEventRecord(c->timed.back().e1, c->stream); }

int32_t Stager::check(const stevi_array* a, const char* name) {
	if (!a || !a->data) return fail(c_, STEVI_E_ARG, "%s: null array", name);
	if (a->ndim < 1 || a->ndim > STEVI_MAX_DIMS) return fail(c_, STEVI_E_ARG, "%s: ndim %d", name, a->ndim);
	if (dtype_size(a->dtype) == 0) return fail(c_, STEVI_E_DTYPE, "%s: unknown dtype %d", name, a->dtype);
	for (int k = 0; k < a->ndim; k++) {
		if (a->shape[k] < 0) return fail(c_, STEVI_E_SHAPE, "%s: negative extent", name);
		if (a->strides[k] < 0) return fail(c_, STEVI_E_UNSUPPORTED, "%s: negative strides are not supported", name);
	}
	return STEVI_OK;
}

static void fill_view(const stevi_array* a, View& v) {
	v.p = a->data;
	v.dtype = a->dtype;
	v.ndim = a->ndim;
	for (int k = 0; k < a->ndim; k++) { v.shape[k] = a->shape[k]; v.st[k] = a->strides[k]; }
}

int32_t Stager::in(const stevi_array* a, View& v, const char* name) {
	STEVI_TRY(check(a, name));
	fill_view(a, v);
	if (a->mem == STEVI_DEVICE) return STEVI_OK;
	any_host_ = true;
	size_t bytes = v.span() * dtype_size(v.dtype);
	void* d = nullptr;
	STEVI_TRY(arena_alloc(c_, bytes, &d));
	if (bytes) STEVI_CUDA(c_, cudaMemcpyAsync(d, a->data, bytes, cudaMemcpyHostToDevice, c_->stream));
	v.p = d;
	return STEVI_OK;
}

int32_t Stager::out(const stevi_array* a, View& v, const char* name) {
	STEVI_TRY(check(a, name));
	fill_view(a, v);
	if (a->mem == STEVI_DEVICE) return STEVI_OK;
	any_host_ = true;
	size_t bytes = v.span() * dtype_size(v.dtype);
	void* d = nullptr;
	STEVI_TRY(arena_alloc(c_, bytes, &d));
	// a strided result leaves gaps: carry the caller's bytes through so the copy back does not clobber them
	if (bytes && !v.dense()) STEVI_CUDA(c_, cudaMemcpyAsync(d, a->data, bytes, cudaMemcpyHostToDevice, c_->stream));
	v.p = d;
	back_.push_back({a->data, d, bytes});
	return STEVI_OK;
}

template<typename TN>
__global__ void __launch_bounds__(256) k_narrow_index(const int32_t* __restrict__ src, TN* __restrict__ dst, size_t n, int sign, int offset) {
	const size_t k = (size_t)blockIdx.x * 256 + threadIdx.x;
	if (k < n) dst[k] = (TN)((src[k] - offset) * sign);
}

int32_t Stager::out_index(const stevi_array* a, View& v, int n_values, int sign, int offset, const char* name) {
	STEVI_TRY(check(a, name));
	const bool dense2 = a->ndim == 2 && a->strides[1] == 1 && a->strides[0] == a->shape[1];
	if (a->dtype == STEVI_U8 || a->dtype == STEVI_U16) {
		// the caller takes the map in a narrow type (non-negative indices only): computed as int32, narrowed into its array
		if (!dense2 || sign != 1 || offset != 0 || n_values < 1 || n_values > (a->dtype == STEVI_U8 ? 256 : 65536))
			return fail(c_, STEVI_E_DTYPE, "%s: a u8 / u16 map must be dense, hold non-negative values (RIGHT_TO_LEFT, offset 0) and fit its type", name);
		fill_view(a, v);
		const size_t n = v.count();
		void* d = nullptr;
		STEVI_TRY(arena_alloc(c_, n * sizeof(int32_t), &d));
		v.p = d; v.dtype = STEVI_I32;
		if (a->mem != STEVI_DEVICE) any_host_ = true;
		narrow_.push_back({nullptr, static_cast<const int32_t*>(d), n, a->dtype == STEVI_U8 ? 1 : 2, 1, 0, a->data, a->mem});
		return STEVI_OK;
	}
	if (a->mem == STEVI_DEVICE || !c_->narrow_wire || a->dtype != STEVI_I32 || !dense2 || n_values > 65536 || n_values < 1 ||
		(sign != 1 && sign != -1))
		return out(a, v, name);
	fill_view(a, v);
	any_host_ = true;
	const size_t n = v.count();
	void* d = nullptr;
	STEVI_TRY(arena_alloc(c_, n * sizeof(int32_t), &d));
	v.p = d;
	narrow_.push_back({static_cast<int32_t*>(a->data), static_cast<const int32_t*>(d), n, n_values <= 256 ? 1 : 2, sign, offset, nullptr, 0});
	return STEVI_OK;
}

static int host_threads() {
	static int n = [] { const char* e = getenv("STEVI_HOST_THREADS"); int v = e ? atoi(e) : 4; return v < 1 ? 1 : v; }();
	return n;
}

// dst[k] = sign * src[k] + offset. The int32 map is written with streaming stores: it is 4x / 2x the bytes that crossed PCIe and
// is not read again by this code, so it should not pull its cache lines in first.
template<typename TN>
static void widen_range(const TN* src, int32_t* dst, long n, int sign, int offset) {
	long k = 0;
	while (k < n && ((uintptr_t)(dst + k) & 31)) { dst[k] = sign * (int32_t)src[k] + offset; k++; }
	const __m256i vo = _mm256_set1_epi32(offset);
	for (; k + 8 <= n; k += 8) {
		__m256i w;
		if (sizeof(TN) == 1) w = _mm256_cvtepu8_epi32(_mm_loadl_epi64(reinterpret_cast<const __m128i*>(src + k)));
		else w = _mm256_cvtepu16_epi32(_mm_loadu_si128(reinterpret_cast<const __m128i*>(src + k)));
		if (sign < 0) w = _mm256_sub_epi32(_mm256_setzero_si256(), w);
		_mm256_stream_si256(reinterpret_cast<__m256i*>(dst + k), _mm256_add_epi32(w, vo));
	}
	for (; k < n; k++) dst[k] = sign * (int32_t)src[k] + offset;
	_mm_sfence();
}

// ---- widening off the caller's thread ----------------------------------------------------------------------------------------
// A process-wide pool of host_threads() workers (they inherit the affinity mask of the thread that makes the first call, i.e. the
// rank's core slice). A task is one narrow map; it is cut into chunks, one queue entry each. Who makes the entries: see
// poll_landed_maps below.
struct WidenTask {
	const void* staged; int32_t* dst; long n; int bytes_per, sign, offset, chunks;
	cudaEvent_t landed; int device;   // recorded after the device-to-host copy
	std::atomic<int> remaining;
	std::atomic<int> submitted;   // handed to the pool once, by the host function or (if the stream never ran it) by the synchronising call
};

class WidenPool {
public:
	static WidenPool& get() { static WidenPool* p = new WidenPool(); return *p; }   // never destroyed: the workers outlive static teardown
	void submit(WidenTask* t) {
		{
			std::lock_guard<std::mutex> lk(m_);
			for (int k = 0; k < t->chunks; k++) q_.push_back({t, k});
		}
		if (t->chunks > 1) cv_.notify_all(); else cv_.notify_one();
	}
	// a thread that waits for a task takes chunks itself instead of spinning beside the workers (a rank may own 4 cores)
	bool run_one() {
		std::pair<WidenTask*, int> job;
		{
			std::lock_guard<std::mutex> lk(m_);
			if (q_.empty()) return false;
			job = q_.front();
			q_.pop_front();
		}
		work(job);
		return true;
	}
private:
	WidenPool() {
		const int n = host_threads();
		for (int k = 0; k < n; k++) std::thread([this] { run(); }).detach();
	}
	static void work(const std::pair<WidenTask*, int>& job) {
		WidenTask* t = job.first;
		const long a = t->n * job.second / t->chunks / 8 * 8, b = job.second + 1 == t->chunks ? t->n : t->n * (job.second + 1) / t->chunks / 8 * 8;
		if (t->bytes_per == 1) widen_range(static_cast<const uint8_t*>(t->staged) + a, t->dst + a, b - a, t->sign, t->offset);
		else widen_range(static_cast<const uint16_t*>(t->staged) + a, t->dst + a, b - a, t->sign, t->offset);
		t->remaining.fetch_sub(1, std::memory_order_release);
	}
	void run() {
		for (;;) {
			std::pair<WidenTask*, int> job;
			{
				std::unique_lock<std::mutex> lk(m_);
				cv_.wait(lk, [this] { return !q_.empty(); });
				job = q_.front();
				q_.pop_front();
			}
			work(job);
		}
	}
	std::mutex m_;
	std::condition_variable cv_;
	std::deque<std::pair<WidenTask*, int>> q_;
};

static void hand_over_once(WidenTask* t) { if (t->submitted.exchange(1) == 0) WidenPool::get().submit(t); }

// Who notices that a copy has landed: measured at C2, 4 pairs in flight (ms per pair, same box): OpenMP region in the synchronising
// call 0.33; a host function in the stream (cudaLaunchHostFunc) 0.52 — the stream stalls around it; the pool fed by the
// synchronising call 0.28; and, built here, every API call of the process polls the events of the pending maps of its device on
// entry, so that the workers widen pair k while the calling thread enqueues pair k+1.
static std::mutex g_pending_m;
static std::vector<WidenTask*> g_pending;
static std::atomic<int> g_pending_n{0};

void poll_landed_maps(int device) {
	if (g_pending_n.load(std::memory_order_relaxed) == 0) return;   // (a miss only delays the hand-over to the next call)
	std::lock_guard<std::mutex> lk(g_pending_m);
	for (size_t k = 0; k < g_pending.size();) {
		WidenTask* t = g_pending[k];
		if (t->device == device && cudaEventQuery(t->landed) == cudaSuccess) {
			hand_over_once(t);
			g_pending[k] = g_pending.back();
			g_pending.pop_back();
			g_pending_n.fetch_sub(1, std::memory_order_relaxed);
		} else k++;
	}
	(void)cudaGetLastError();   // cudaErrorNotReady is not an error
}

// after a stream synchronisation: every host function of the stream has run, i.e. every task has been handed to the pool; wait
// for the ones still being widened
void finish_host(stevi_ctx* c) {
	for (auto& w : c->widen) {
		WidenTask* t = static_cast<WidenTask*>(w.task);
		{
			std::lock_guard<std::mutex> lk(g_pending_m);
			for (size_t k = 0; k < g_pending.size(); k++)
				if (g_pending[k] == t) { g_pending[k] = g_pending.back(); g_pending.pop_back(); g_pending_n.fetch_sub(1, std::memory_order_relaxed); break; }
		}
		hand_over_once(t);
		while (t->remaining.load(std::memory_order_acquire) != 0)
			if (!WidenPool::get().run_one()) std::this_thread::yield();
		if (t->landed) cudaEventDestroy(t->landed);
		delete t;
		c->pinned[w.slot].busy = false;
	}
	c->widen.clear();
}

int32_t Stager::finish() {
	for (auto& b : back_)
		if (b.bytes) STEVI_CUDA(c_, cudaMemcpyAsync(b.host, b.dev, b.bytes, cudaMemcpyDeviceToHost, c_->stream));
	back_.clear();
	for (auto& nw : narrow_) {
		if (!nw.n) continue;
		const size_t bytes = nw.n * nw.bytes_per;
		void* dn = nullptr;
		if (nw.final && nw.final_mem == STEVI_DEVICE) dn = nw.final;
		else STEVI_TRY(arena_alloc(c_, bytes, &dn));
		const unsigned blocks = (unsigned)((nw.n + 255) / 256);
		if (nw.bytes_per == 1) STEVI_LAUNCH(c_, k_narrow_index<uint8_t>, blocks, 256, 0, nw.dev, static_cast<uint8_t*>(dn), nw.n, nw.sign, nw.offset);
		else STEVI_LAUNCH(c_, k_narrow_index<uint16_t>, blocks, 256, 0, nw.dev, static_cast<uint16_t*>(dn), nw.n, nw.sign, nw.offset);
		if (nw.final) {   // the caller's array is narrow itself
			if (nw.final_mem != STEVI_DEVICE) STEVI_CUDA(c_, cudaMemcpyAsync(nw.final, dn, bytes, cudaMemcpyDeviceToHost, c_->stream));
			continue;
		}
		// a page-locked staging buffer that no pending task uses
		size_t slot = c_->pinned.size();
		for (size_t k = 0; k < c_->pinned.size(); k++)
			if (!c_->pinned[k].busy && c_->pinned[k].cap >= bytes) { slot = k; break; }
		if (slot == c_->pinned.size()) {
			void* hp = nullptr;
			STEVI_CUDA(c_, cudaHostAlloc(&hp, bytes, cudaHostAllocDefault));
			c_->pinned.push_back({hp, bytes, false});
		}
		c_->pinned[slot].busy = true;
		STEVI_CUDA(c_, cudaMemcpyAsync(c_->pinned[slot].p, dn, bytes, cudaMemcpyDeviceToHost, c_->stream));
		(void)WidenPool::get();   // the workers are started by an API-calling thread (they inherit ITS affinity mask, not the driver's callback thread's)
		WidenTask* t = new WidenTask();
		t->staged = c_->pinned[slot].p; t->dst = nw.host; t->n = (long)nw.n; t->bytes_per = nw.bytes_per; t->sign = nw.sign; t->offset = nw.offset;
		t->chunks = (int)std::max(1L, std::min(2L * host_threads(), t->n / 65536));   // (more chunks than workers: the waiting thread takes some)
		t->remaining.store(t->chunks, std::memory_order_relaxed);
		t->submitted.store(0, std::memory_order_relaxed);
		c_->widen.push_back({c_->pinned[slot].p, nw.host, nw.n, nw.bytes_per, nw.sign, nw.offset, slot, t});
		t->device = c_->device; t->landed = nullptr;
		static const bool no_poll = getenv("STEVI_WIDEN_NO_POLL") != nullptr;   // measurement aid: hand over at the synchronisation only
		if (no_poll) {
		} else if (cudaEventCreateWithFlags(&t->landed, cudaEventDisableTiming) == cudaSuccess && cudaEventRecord(t->landed, c_->stream) == cudaSuccess) {
			std::lock_guard<std::mutex> lk(g_pending_m);
			g_pending.push_back(t);
			g_pending_n.fetch_add(1, std::memory_order_relaxed);
		} else (void)cudaGetLastError();   // (then the synchronising call hands the task over)
	}
	narrow_.clear();
	if (any_host_ && !c_->deferred_host) {
		STEVI_CUDA(c_, cudaStreamSynchronize(c_->stream));
		finish_host(c_);
		return check_fault(c_);
	}
	return STEVI_OK;
}

} // namespace stevi

using namespace stevi;

extern "C" {

const char* stevi_version(void) { return "stevi_b200 0.1 (sm_100a)"; }

int32_t stevi_device_count(void) {
	int n = 0;
	if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
	return n;
}

int32_t stevi_ctx_create(int32_t device, void* stream, stevi_ctx** out) {
	if (!out) return STEVI_E_ARG;
	*out = nullptr;
	int n = 0;
	cudaError_t e = cudaGetDeviceCount(&n);
	if (e != cudaSuccess || n == 0) { cudaGetLastError(); return STEVI_E_CUDA; } // no device: fail loudly, no fallback
	if (device < 0 || device >= n) return STEVI_E_ARG;
	if (cudaSetDevice(device) != cudaSuccess) return STEVI_E_CUDA;
	stevi_ctx* c = new stevi_ctx();
	c->device = device;
	if (stream) { c->stream = static_cast<cudaStream_t>(stream); c->own_stream = false; }
	else {
		if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) { delete c; return STEVI_E_CUDA; }
		c->own_stream = true;
	}
	cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, device);
	if (cudaHostAlloc((void**)&c->fault_host, sizeof(int), cudaHostAllocMapped) != cudaSuccess ||
		cudaHostGetDevicePointer((void**)&c->fault_dev, c->fault_host, 0) != cudaSuccess) {
		cudaGetLastError();
		if (c->own_stream) cudaStreamDestroy(c->stream);
		delete c;
		return STEVI_E_CUDA;
	}
	*c->fault_host = 0;
	*out = c;
	return STEVI_OK;
}

int32_t stevi_ctx_destroy(stevi_ctx* c) {
	if (!c) return STEVI_OK;
	cudaSetDevice(c->device);
	cudaStreamSynchronize(c->stream);
	for (auto& b : c->blocks) cudaFree(b.p);
	for (auto& t : c->timed) { cudaEventDestroy(t.e0); cudaEventDestroy(t.e1); }
	for (auto e : c->event_pool) cudaEventDestroy(e);
	if (c->own_stream) cudaStreamDestroy(c->stream);
	if (c->fault_host) cudaFreeHost(c->fault_host);
	if (c->sgm_xchg) cudaFree(c->sgm_xchg);
	finish_host(c);   // (the stream has been synchronised: every pending map has been handed to the workers; wait for them)
	for (auto& pb : c->pinned) cudaFreeHost(pb.p);
	delete c;
	return STEVI_OK;
}

int32_t stevi_ctx_set_stream(stevi_ctx* c, void* stream) {
	if (!c) return STEVI_E_ARG;
	if (stream && static_cast<cudaStream_t>(stream) == c->stream) return STEVI_OK;
	STEVI_CUDA(c, cudaStreamSynchronize(c->stream));   // the workspace is reused by the next call, on the new stream
	if (c->own_stream) { cudaStreamDestroy(c->stream); c->own_stream = false; }
	if (stream) c->stream = static_cast<cudaStream_t>(stream);
	else {
		STEVI_CUDA(c, cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
		c->own_stream = true;
	}
	return STEVI_OK;
}

int32_t stevi_ctx_set_deferred_host(stevi_ctx* c, int32_t enable) {
	if (!c) return STEVI_E_ARG;
	c->deferred_host = enable != 0;
	return STEVI_OK;
}

int32_t stevi_ctx_set_row_origin(stevi_ctx* c, int32_t row) {
	if (!c) return STEVI_E_ARG;
	c->row_origin = row;
	return STEVI_OK;
}

void* stevi_ctx_stream(stevi_ctx* c) { return c ? (void*)c->stream : nullptr; }

int32_t stevi_ctx_synchronize(stevi_ctx* c) {
	if (!c) return STEVI_E_ARG;
	STEVI_CUDA(c, cudaStreamSynchronize(c->stream));
	finish_host(c);
	return check_fault(c);
}

int32_t stevi_ctx_set_narrow_wire(stevi_ctx* c, int32_t enable) {
	if (!c) return STEVI_E_ARG;
	c->narrow_wire = enable != 0;
	return STEVI_OK;
}

const char* stevi_last_error(stevi_ctx* c) { return c ? c->err.c_str() : "null context"; }
int64_t stevi_ctx_launch_count(stevi_ctx* c) { return c ? c->launches : 0; }
int64_t stevi_ctx_workspace_bytes(stevi_ctx* c) {
	if (!c) return 0;
	size_t t = c->sgm_xchg_bytes;
	for (auto& b : c->blocks) t += b.cap;
	return (int64_t)t;
}

int32_t stevi_ctx_set_timing(stevi_ctx* c, int32_t enable) {
	if (!c) return STEVI_E_ARG;
	c->timing = enable != 0;
	return STEVI_OK;
}

// "kernel name\tms\n" per launch recorded since the last report, in launch order; clears the record
const char* stevi_ctx_timing_report(stevi_ctx* c) {
	if (!c) return "";
	cudaStreamSynchronize(c->stream);
	c->report.clear();
	char buf[64];
	for (auto& t : c->timed) {
		float ms = 0.f;
		cudaEventElapsedTime(&ms, t.e0, t.e1);
		c->report += t.name;
		snprintf(buf, sizeof buf, "\t%.6f\n", ms);
		c->report += buf;
		c->event_pool.push_back(t.e0);
		c->event_pool.push_back(t.e1);
	}
	c->timed.clear();
	return c->report.c_str();
}

int32_t stevi_host_alloc(size_t bytes, void** out) {
	if (!out) return STEVI_E_ARG;
	if (cudaHostAlloc(out, bytes, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); *out = nullptr; return STEVI_E_CUDA; }
	return STEVI_OK;
}
int32_t stevi_host_free(void* p) {
	if (p && cudaFreeHost(p) != cudaSuccess) { cudaGetLastError(); return STEVI_E_CUDA; }
	return STEVI_OK;
}

} // extern "C"
