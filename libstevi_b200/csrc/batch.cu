// Batches of independent stereo pairs over several contexts / devices behind the C ABI: what
// api.HostPairStream + sharding.shard_pairs give a Python caller, for a C or C++ caller. The reference processes a batch as a
// loop of calls on the host cores (test/benchmarks/benchmarkCostVolumeWorkflow.cpp:175-260); here the loop keeps one pair in flight
// per context: contexts on the same device overlap the copies of one pair with the kernels of another, contexts on different
// devices run side by side, each device driven by its own host thread. No data-path collective: pairs are independent.
#include "common.cuh"
#include <thread>
#include <atomic>
#include <map>

extern "C" int32_t stevi_batch_census_sgm(stevi_ctx* const* ctxs, int32_t n_ctx, const stevi_pair* pairs, int32_t n_pairs,
										   int32_t hr, int32_t vr, int32_t D, int32_t direction, int32_t ndir, float P1, float P2, float Pout,
										   const stevi_margins* margins, int32_t semantics) {
	if (!ctxs || n_ctx < 1 || (!pairs && n_pairs > 0) || n_pairs < 0) return STEVI_E_ARG;
	for (int k = 0; k < n_ctx; k++) if (!ctxs[k]) return STEVI_E_ARG;
	// contexts grouped by device; pair k goes to device group k mod n_groups and, inside it, round robin over its contexts
	std::map<int, std::vector<stevi_ctx*>> groups;
	for (int k = 0; k < n_ctx; k++) groups[ctxs[k]->device].push_back(ctxs[k]);
	std::vector<std::vector<stevi_ctx*>> g;
	for (auto& kv : groups) g.push_back(kv.second);
	const int ng = (int)g.size();
	std::atomic<int32_t> status(STEVI_OK);
	std::vector<bool> was_deferred(n_ctx);
	for (int k = 0; k < n_ctx; k++) { was_deferred[k] = ctxs[k]->deferred_host; ctxs[k]->deferred_host = true; }
	auto worker = [&](int gi) {
		auto& cs = g[gi];
		size_t next = 0;
		for (int k = gi; k < n_pairs && status.load() == STEVI_OK; k += ng) {
			stevi_ctx* c = cs[next];
			next = (next + 1) % cs.size();
			int32_t s = stevi_ctx_synchronize(c);   // the pair this context carried before: its result is complete now
			if (s == STEVI_OK)
				s = stevi_census_sgm_pipeline(c, pairs[k].img_l, pairs[k].img_r, hr, vr, D, direction, ndir, P1, P2, Pout, margins, semantics,
											  nullptr, nullptr, pairs[k].disp);
			if (s != STEVI_OK) { int32_t ok = STEVI_OK; status.compare_exchange_strong(ok, s); }
		}
		for (auto* c : cs) {
			const int32_t s = stevi_ctx_synchronize(c);
			if (s != STEVI_OK) { int32_t ok = STEVI_OK; status.compare_exchange_strong(ok, s); }
		}
	};
	if (ng == 1) worker(0);
	else {
		std::vector<std::thread> th;
		for (int gi = 0; gi < ng; gi++) th.emplace_back(worker, gi);
		for (auto& t : th) t.join();
	}
	for (int k = 0; k < n_ctx; k++) ctxs[k]->deferred_host = was_deferred[k];
	return status.load();
}
