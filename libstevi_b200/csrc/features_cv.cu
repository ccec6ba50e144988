// Feature-level entry points of the cost-volume path, for callers that bring their own feature volumes (compressed unfolds,
// multi-channel images, learned features):
//   features_for_matchfunc_dev   getFeatureVolumeForMatchFunc<matchFunc>(feature_vol) (cross_correlations.h:645-722) on a ready-made
//                                (H,W,F) float volume: v | v - mean | v / norm | (v - mean) / norm in the reference's operation order
//                                (mean = sum * (1/F), correlation_base.h:1116-1130; norm = sqrt(sum of squares), cross_correlations.h:81-99;
//                                no epsilon: a constant vector gives NaN, Q6);
//   aggregate_cost_dev           aggregateCost<matchFunc> (cross_correlations.h:194-249, searchOffset<1> overload :251-308):
//                                cv(i,j,d) = compare(source(i,j,.), target(i, j + sign*(d + lower), .)), target features 0 outside the
//                                image, the F terms accumulated in channel order with separately rounded operations.
// featureVolume2CostVolume (:724-738) and the compressor / multi-channel overloads of unfoldBasedCostVolume (:741-791) are these
// two after stevi_unfold / stevi_unfold_compressed (api.cu).
#include "common.cuh"
#include "kernels.cuh"

namespace stevi {

namespace {

// thread per pixel; the feature axis is walked three times (mean, norm, output) like the reference's three passes
__global__ void __launch_bounds__(128) k_features_for_matchfunc(const float* __restrict__ in, int64_t s0, int64_t s1, int64_t s2, int H, int W, int F,
																	 int zero_mean, int normalized, float* __restrict__ out, int64_t o0, int64_t o1, int64_t o2) {
	const int j = blockIdx.x * 128 + threadIdx.x, i = blockIdx.y;
	if (j >= W) return;
	const float* v = in + i * s0 + j * s1;
	float mean = 0.f;
	if (zero_mean) {
		for (int c = 0; c < F; c++) mean = __fadd_rn(mean, v[c * s2]);
		mean = __fmul_rn(mean, 1.f / (float)F);
	}
	float norm = 1.f;
	if (normalized) {
		float n = 0.f;
		for (int c = 0; c < F; c++) { const float t = zero_mean ? __fsub_rn(v[c * s2], mean) : v[c * s2]; n = __fadd_rn(n, __fmul_rn(t, t)); }
		norm = sqrtf(n);
	}
	float* o = out + i * o0 + j * o1;
	for (int c = 0; c < F; c++) {
		float t = zero_mean ? __fsub_rn(v[c * s2], mean) : v[c * s2];
		if (normalized) t = __fdiv_rn(t, norm);
		o[c * o2] = t;
	}
}

enum { AG_PROD = 0, AG_SQ = 1, AG_ABS = 2 };

struct AggFeat {
	const float* src; int64_t a0, a1, a2;
	const float* tgt; int64_t b0, b1, b2;
	float* out; int64_t o0, o1, o2;
	int H, Ws, Wt, F, D, sign, lower;
};

// Block = 64 source pixels of one row x a chunk of 8 disparities: the F features of the 64 source pixels and of the 64 + 7 + ...
// target pixels the chunk touches are staged in shared memory once (features fastest), then a thread (pixel, disparity) walks
// its two vectors. One global read per feature and pixel instead of one per feature and cell.
template<int OP>
__global__ void __launch_bounds__(256) k_aggregate_features(const AggFeat p) {
	extern __shared__ float sm[];
	constexpr int PX = 64, DCH = 4;   // 64 pixels x 4 disparities per pass; the block covers blockDim.y... disparities [d0, d0 + DSPAN)
	const int DSPAN = 32;
	const int i = blockIdx.y, j0 = blockIdx.x * PX, d0 = blockIdx.z * DSPAN;
	const int dn = min(DSPAN, p.D - d0);
	const int Fp = p.F | 1;                       // odd pitch: pixels of a warp hit different banks
	float* ssrc = sm;                             // [PX][Fp]
	float* stgt = sm + PX * Fp;                   // [PX + DSPAN - 1][Fp], target column t0 + k
	// target columns needed: j + sign * (d + lower) for j in [j0, j0+PX), d in [d0, d0+dn)
	const int tlo = p.sign > 0 ? j0 + (d0 + p.lower) : j0 - (d0 + dn - 1 + p.lower);
	const int nt = PX + dn - 1;
	for (int e = threadIdx.x; e < PX * p.F; e += 256) {
		const int px = e / p.F, c = e - px * p.F;
		const int j = j0 + px;
		ssrc[px * Fp + c] = j < p.Ws ? p.src[i * p.a0 + j * p.a1 + c * p.a2] : 0.f;
	}
	for (int e = threadIdx.x; e < nt * p.F; e += 256) {
		const int k = e / p.F, c = e - k * p.F;
		const int jt = tlo + k;
		stgt[k * Fp + c] = (jt >= 0 && jt < p.Wt) ? p.tgt[i * p.b0 + jt * p.b1 + c * p.b2] : 0.f;   // valueOrAlt(..., 0), :235
	}
	__syncthreads();
	const int px = threadIdx.x & 63, dq = threadIdx.x >> 6;
	const int j = j0 + px;
	if (j >= p.Ws) return;
	const float* s = ssrc + px * Fp;
	for (int dd = dq; dd < dn; dd += DCH) {
		const int d = d0 + dd;
		const int jt = j + p.sign * (d + p.lower);
		const float* t = stgt + (jt - tlo) * Fp;
		float acc = 0.f;
		for (int c = 0; c < p.F; c++) {
			if (OP == AG_PROD) acc = __fadd_rn(acc, __fmul_rn(s[c], t[c]));
			else if (OP == AG_SQ) { const float e = __fsub_rn(s[c], t[c]); acc = __fadd_rn(acc, __fmul_rn(e, e)); }
			else acc = __fadd_rn(acc, fabsf(__fsub_rn(s[c], t[c])));
		}
		p.out[i * p.o0 + j * p.o1 + d * p.o2] = acc;
	}
}

} // namespace

static int32_t traits_of(stevi_ctx* c, int mf, bool& zm, bool& nm, int& op) {
	zm = nm = false;
	switch (mf) {
	case STEVI_MF_CC: op = AG_PROD; break;
	case STEVI_MF_NCC: op = AG_PROD; nm = true; break;
	case STEVI_MF_ZCC: op = AG_PROD; zm = true; break;
	case STEVI_MF_ZNCC: op = AG_PROD; zm = nm = true; break;
	case STEVI_MF_SSD: op = AG_SQ; break;
	case STEVI_MF_ZSSD: op = AG_SQ; zm = true; break;
	case STEVI_MF_SAD: op = AG_ABS; break;
	case STEVI_MF_ZSAD: op = AG_ABS; zm = true; break;
	default: return fail(c, STEVI_E_UNSUPPORTED, "matching function %d has no float feature form here (census: stevi_census_features + "
						 "stevi_hamming_cost_volume; MEDAD / ZMEDAD / KERMI are not provided)", mf);
	}
	return STEVI_OK;
}

int32_t features_for_matchfunc_dev(stevi_ctx* c, int mf, const View& in, const View& out) {
	bool zm, nm; int op;
	STEVI_TRY(traits_of(c, mf, zm, nm, op));
	if (in.ndim != 3 || in.dtype != STEVI_F32) return fail(c, STEVI_E_DTYPE, "features_for_matchfunc: the feature volume must be (H,W,F) f32");
	const int H = in.shape[0], W = in.shape[1], F = in.shape[2];
	if (out.ndim != 3 || out.dtype != STEVI_F32 || out.shape[0] != H || out.shape[1] != W || out.shape[2] != F)
		return fail(c, STEVI_E_SHAPE, "features_for_matchfunc: out must be (%d,%d,%d) f32", H, W, F);
	if (H == 0 || W == 0 || F == 0) return STEVI_OK;
	if (H > 65535) return fail(c, STEVI_E_UNSUPPORTED, "features_for_matchfunc: more than 65535 rows");
	dim3 grid(cdiv(W, 128), H);
	STEVI_LAUNCH(c, k_features_for_matchfunc, grid, 128, 0, in.as<float>(), in.st[0], in.st[1], in.st[2], H, W, F, zm ? 1 : 0, nm ? 1 : 0,
				 out.as<float>(), out.st[0], out.st[1], out.st[2]);
	return STEVI_OK;
}

int32_t aggregate_cost_dev(stevi_ctx* c, int mf, const View& fl, const View& fr, int lower, int D, int direction, const View& out) {
	bool zm, nm; int op;
	STEVI_TRY(traits_of(c, mf, zm, nm, op));
	if (fl.ndim != 3 || fr.ndim != 3 || fl.dtype != STEVI_F32 || fr.dtype != STEVI_F32)
		return fail(c, STEVI_E_DTYPE, "aggregate_cost: feature volumes must be (H,W,F) f32");
	// cross_correlations.h:209-211 (heights) and :753-759 (channels): the reference returns an empty array
	if (fl.shape[0] != fr.shape[0] || fl.shape[2] != fr.shape[2])
		return fail(c, STEVI_E_SHAPE, "aggregate_cost: feature volumes disagree in height or feature count");
	if (D < 1) return fail(c, STEVI_E_ARG, "aggregate_cost: the search range is empty");
	const bool r2l = direction == STEVI_RIGHT_TO_LEFT;
	const View& src = r2l ? fr : fl;
	const View& tgt = r2l ? fl : fr;
	const int H = src.shape[0], Ws = src.shape[1], F = src.shape[2];
	if (out.ndim != 3 || out.dtype != STEVI_F32 || out.shape[0] != H || out.shape[1] != Ws || out.shape[2] != D)
		return fail(c, STEVI_E_SHAPE, "aggregate_cost: out must be (%d,%d,%d) f32", H, Ws, D);
	if (H == 0 || Ws == 0) return STEVI_OK;
	if (H > 65535 || cdiv(D, 32) > 65535) return fail(c, STEVI_E_UNSUPPORTED, "aggregate_cost: more than 65535 rows or disparity chunks");
	AggFeat p;
	p.src = src.as<float>(); p.a0 = src.st[0]; p.a1 = src.st[1]; p.a2 = src.st[2];
	p.tgt = tgt.as<float>(); p.b0 = tgt.st[0]; p.b1 = tgt.st[1]; p.b2 = tgt.st[2];
	p.out = out.as<float>(); p.o0 = out.st[0]; p.o1 = out.st[1]; p.o2 = out.st[2];
	p.H = H; p.Ws = Ws; p.Wt = tgt.shape[1]; p.F = F; p.D = D; p.sign = r2l ? 1 : -1; p.lower = lower;
	const size_t smem = (size_t)(64 + 64 + 31) * (F | 1) * sizeof(float);
	if (smem > (200u << 10)) return fail(c, STEVI_E_UNSUPPORTED, "aggregate_cost: more than %d features per pixel", (int)((200u << 10) / (159 * 4)));
	dim3 grid(cdiv(Ws, 64), H, cdiv(D, 32));
#define STEVI_AG(OP) do { \
	if (smem > (48u << 10)) STEVI_CUDA(c, cudaFuncSetAttribute(k_aggregate_features<OP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
	STEVI_LAUNCH(c, k_aggregate_features<OP>, grid, 256, smem, p); } while (0)
	if (op == AG_PROD) STEVI_AG(AG_PROD); else if (op == AG_SQ) STEVI_AG(AG_SQ); else STEVI_AG(AG_ABS);
#undef STEVI_AG
	return STEVI_OK;
}

} // namespace stevi
