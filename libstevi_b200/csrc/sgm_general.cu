// General SGM recurrence (placeholder until the warp-per-line kernel lands).
#include "common.cuh"
#include "kernels.cuh"
namespace stevi {
int32_t sgm_general(stevi_ctx* c, const View& cv, int ndir, int strategy, float P1, float P2, float Pout,
					const stevi_margins& mg, int semantics, const View& out) {
	return fail(c, STEVI_E_UNSUPPORTED, "sgm: general recurrence not built yet");
}
}
