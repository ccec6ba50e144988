// Materialised feature volumes. The cost-volume kernels never build these —
// they exist for callers of the reference's feature-level API.
//
//   unfold.h:248-344      unfold: out(i,j,c) = img(i+k-pad_top, j+l-pad_left[, ch]), 0 outside; c = channelFromCord (:171-191)
//   unfold.h:36-137       UnFoldCompressor: super-pixel means, features in ascending label order, weight 1/count
//   unfold.h:346-471      compressed unfold: out(i,j,ch*nF+f) += weight*img(...) in index order
//   census.h:69-115       censusFeatures on an (H,W,F) volume (Q1: only full 32-bit words are stored)
//   cross_correlations.h:645-722  getFeatureVolumeForMatchFunc: (v-mean)/norm | v-mean | v/norm | v
#include "common.cuh"
#include "kernels.cuh"
#include <map>

namespace stevi {

namespace {

__host__ __device__ inline int channel_from_cord(int k, int l, int ch, int hs, int vs, int nch, int orientation) {
	switch (orientation) {
	case 0: return nch * hs * k + nch * l + ch;
	case 1: return nch * vs * (hs - l - 1) + nch * k + ch;
	case 2: return nch * hs * (vs - k - 1) + nch * (hs - l - 1) + ch;
	default: return nch * vs * l + nch * (vs - k - 1) + ch;
	}
}

struct Unfold {
	const void* img; int64_t s0, s1, s2; int H, W, C;
	void* out; int64_t o0, o1, o2; int oH, oW;
	int hs, vs, pad_left, pad_top, orientation;
};

template<typename TI, typename TO>
__global__ void __launch_bounds__(128) k_unfold(const Unfold p) {
	const int j = blockIdx.x * 128 + threadIdx.x, i = blockIdx.y;
	if (j >= p.oW) return;
	int z = blockIdx.z;                         // (k, l, ch)
	const int ch = z % p.C; z /= p.C;
	const int l = z % p.hs, k = z / p.hs;
	const int ii = i + k - p.pad_top, jj = j + l - p.pad_left;
	TI v = TI(0);
	if (ii >= 0 && ii < p.H && jj >= 0 && jj < p.W) v = static_cast<const TI*>(p.img)[ii * p.s0 + jj * p.s1 + ch * p.s2];
	const int c = channel_from_cord(k, l, ch, p.hs, p.vs, p.C, p.orientation);
	static_cast<TO*>(p.out)[i * p.o0 + j * p.o1 + c * p.o2] = (TO)v;
}

struct CompIndex { int k, l, f; float w; };

template<typename TI>
__global__ void __launch_bounds__(128) k_unfold_compressed(const Unfold p, const CompIndex* __restrict__ idx, int n_idx, int nF, int top, int left) {
	const int j = blockIdx.x * 128 + threadIdx.x, i = blockIdx.y, ch = blockIdx.z;
	if (j >= p.oW) return;
	float* o = static_cast<float*>(p.out) + i * p.o0 + j * p.o1;
	for (int f = 0; f < nF; f++) o[(ch * nF + f) * p.o2] = 0.f;   // features without pixels cannot occur, kept for safety
	float acc = 0.f;
	int cur = -1;
	for (int n = 0; n < n_idx; n++) {
		const CompIndex e = idx[n];
		if (e.f != cur) { if (cur >= 0) o[(ch * nF + cur) * p.o2] = acc; cur = e.f; acc = 0.f; }
		const int ii = i + e.k + top - p.pad_top, jj = j + e.l + left - p.pad_left;
		TI v = TI(0);
		if (ii >= 0 && ii < p.H && jj >= 0 && jj < p.W) v = static_cast<const TI*>(p.img)[ii * p.s0 + jj * p.s1 + ch * p.s2];
		acc = __fadd_rn(acc, __fmul_rn(e.w, (float)v));
	}
	if (cur >= 0) o[(ch * nF + cur) * p.o2] = acc;
}

template<typename T>
__global__ void __launch_bounds__(128) k_census_features(const T* __restrict__ fv, int64_t s0, int64_t s1, int64_t s2, int H, int W, int F,
														 int n_words, int n_written, uint32_t* __restrict__ out, int64_t o0, int64_t o1, int64_t o2) {
	const int j = blockIdx.x * 128 + threadIdx.x, i = blockIdx.y;
	if (j >= W) return;
	const T* f = fv + i * s0 + j * s1;
	uint32_t* o = out + i * o0 + j * o1;
	const T ref = f[0];
	uint32_t word = 0;
	int b = 0, w = 0;
	for (int c = 1; c < F; c++) {
		word |= (ref > f[c * s2] ? 1u : 0u) << b;
		if (++b == 32) { if (w < n_written) o[w * o2] = word; w++; word = 0; b = 0; }
	}
	if (b > 0 && w < n_written) { o[w * o2] = word; w++; }      // corrected semantics flush the partial word
	for (; w < n_words; w++) o[w * o2] = 0u;
}

// getFeatureVolumeForMatchFunc on a float image: thread per pixel, reference summation order
__global__ void __launch_bounds__(128) k_feature_volume(const float* __restrict__ img, int64_t s0, int64_t s1, int H, int W, int hr, int vr,
														int zero_mean, int normalized, float* __restrict__ out, int64_t o0, int64_t o1, int64_t o2) {
	const int j = blockIdx.x * 128 + threadIdx.x, i = blockIdx.y;
	if (j >= W) return;
	const int hs = 2 * hr + 1, vs = 2 * vr + 1;
	auto px = [&](int k, int l) -> float {
		const int ii = i - vr + k, jj = j - hr + l;
		return (ii >= 0 && ii < H && jj >= 0 && jj < W) ? img[ii * s0 + jj * s1] : 0.f;
	};
	float mean = 0.f;
	if (zero_mean) {
		for (int k = 0; k < vs; k++) for (int l = 0; l < hs; l++) mean = __fadd_rn(mean, px(k, l));
		mean = __fmul_rn(mean, 1.f / (float)(hs * vs));
	}
	float norm = 1.f;
	if (normalized) {
		float n = 0.f;
		for (int k = 0; k < vs; k++) for (int l = 0; l < hs; l++) { const float t = zero_mean ? __fsub_rn(px(k, l), mean) : px(k, l); n = __fadd_rn(n, __fmul_rn(t, t)); }
		norm = sqrtf(n);
	}
	float* o = out + i * o0 + j * o1;
	for (int k = 0; k < vs; k++) for (int l = 0; l < hs; l++) {
		float t = zero_mean ? __fsub_rn(px(k, l), mean) : px(k, l);
		if (normalized) t = __fdiv_rn(t, norm);
		o[(hs * k + l) * o2] = t;
	}
}

} // namespace

static int32_t fill_unfold(stevi_ctx* c, const View& img, const View& out, int hs, int vs, const int32_t* pad, int auto_l, int auto_t,
						   int auto_r, int auto_b, int n_feat_per_channel, Unfold& p) {
	if (img.ndim != 2 && img.ndim != 3) return fail(c, STEVI_E_SHAPE, "unfold: image must be (H,W) or (H,W,C)");
	const bool is_auto = !pad || pad[0] < 0 || pad[1] < 0 || pad[2] < 0 || pad[3] < 0;   // PaddingMargins(): auto
	const int pl = is_auto ? auto_l : pad[0], pt = is_auto ? auto_t : pad[1], pr = is_auto ? auto_r : pad[2], pb = is_auto ? auto_b : pad[3];
	p.img = img.p; p.s0 = img.st[0]; p.s1 = img.st[1]; p.s2 = img.ndim == 3 ? img.st[2] : 0;
	p.H = img.shape[0]; p.W = img.shape[1]; p.C = img.ndim == 3 ? img.shape[2] : 1;
	p.oH = p.H - vs + pt + pb + 1; p.oW = p.W - hs + pl + pr + 1;
	p.hs = hs; p.vs = vs; p.pad_left = pl; p.pad_top = pt;
	if (p.oH < 0 || p.oW < 0) return fail(c, STEVI_E_SHAPE, "unfold: window larger than the padded image");
	if (out.ndim != 3 || out.shape[0] != p.oH || out.shape[1] != p.oW || out.shape[2] != n_feat_per_channel * p.C)
		return fail(c, STEVI_E_SHAPE, "unfold: out must be (%d,%d,%d)", p.oH, p.oW, n_feat_per_channel * p.C);
	p.out = out.p; p.o0 = out.st[0]; p.o1 = out.st[1]; p.o2 = out.st[2];
	if (p.oH > 65535) return fail(c, STEVI_E_UNSUPPORTED, "unfold: more than 65535 rows");
	return STEVI_OK;
}

int32_t unfold_dev(stevi_ctx* c, const View& img, int hr, int vr, const int32_t* pad, int orientation, const View& out) {
	if (hr < 0 || vr < 0 || hr > 127 || vr > 127 || orientation < 0 || orientation > 3) return fail(c, STEVI_E_ARG, "unfold: radii / orientation");
	Unfold p;
	const int hs = 2 * hr + 1, vs = 2 * vr + 1;
	STEVI_TRY(fill_unfold(c, img, out, hs, vs, pad, hr, vr, hr, vr, hs * vs, p));
	p.orientation = orientation;
	if (p.oH == 0 || p.oW == 0) return STEVI_OK;
	if ((int64_t)hs * vs * p.C > 65535) return fail(c, STEVI_E_UNSUPPORTED, "unfold: more than 65535 features");
	dim3 grid(cdiv(p.oW, 128), p.oH, hs * vs * p.C);
	if (img.dtype == STEVI_U8 && out.dtype == STEVI_U8) { STEVI_LAUNCH(c, (k_unfold<uint8_t, uint8_t>), grid, 128, 0, p); }
	else if (img.dtype == STEVI_U8 && out.dtype == STEVI_F32) { STEVI_LAUNCH(c, (k_unfold<uint8_t, float>), grid, 128, 0, p); }
	else if (img.dtype == STEVI_F32 && out.dtype == STEVI_F32) { STEVI_LAUNCH(c, (k_unfold<float, float>), grid, 128, 0, p); }
	else return fail(c, STEVI_E_DTYPE, "unfold: u8->u8, u8->f32 or f32->f32");
	return STEVI_OK;
}

// mask: host memory, (mh, mw) int32 with element strides m0, m1 (a small configuration array, like the reference's masks)
int32_t unfold_compressed_dev(stevi_ctx* c, const int32_t* mask, int mh, int mw, int64_t m0, int64_t m1, const View& img,
							  const int32_t* pad, const View& out) {
	if (!mask || mh < 1 || mw < 1) return fail(c, STEVI_E_ARG, "unfold_compressed: mask");
	// UnFoldCompressor's constructor, unfold.h:48-122
	const int v_off = mh / 2, h_off = mw / 2;
	int minH = 0, maxH = 0, minW = 0, maxW = 0;
	std::map<int, int> count;
	for (int i = 0; i < mh; i++) for (int j = 0; j < mw; j++) {
		const int feat = mask[i * m0 + j * m1];
		if (feat <= 0) continue;
		minH = std::min(minH, i - v_off); maxH = std::max(maxH, i - v_off);
		minW = std::min(minW, j - h_off); maxW = std::max(maxW, j - h_off);
		count[feat]++;
	}
	const int nF = (int)count.size();
	std::vector<CompIndex> idx;
	int f = 0;
	for (auto& kv : count) {   // std::map iterates in ascending label order = the reference's sorted feats
		for (int i = 0; i < mh; i++) for (int j = 0; j < mw; j++)
			if (mask[i * m0 + j * m1] == kv.first) idx.push_back({i - v_off, j - h_off, f, static_cast<float>(1. / kv.second)});
		f++;
	}
	Unfold p;
	STEVI_TRY(fill_unfold(c, img, out, maxW - minW + 1, maxH - minH + 1, pad, -minW, -minH, maxW, maxH, nF, p));
	if (out.dtype != STEVI_F32) return fail(c, STEVI_E_DTYPE, "unfold_compressed: out must be f32");
	if (p.oH == 0 || p.oW == 0 || nF == 0) return STEVI_OK;
	CompIndex* d_idx = nullptr;
	STEVI_TRY(arena_alloc_t(c, idx.size(), &d_idx));
	// the index list lives in pageable host memory that dies with this call: the copy below is synchronous with respect to the host
	STEVI_CUDA(c, cudaMemcpyAsync(d_idx, idx.data(), idx.size() * sizeof(CompIndex), cudaMemcpyHostToDevice, c->stream));
	STEVI_CUDA(c, cudaStreamSynchronize(c->stream));
	dim3 grid(cdiv(p.oW, 128), p.oH, p.C);
	if (img.dtype == STEVI_U8) { STEVI_LAUNCH(c, k_unfold_compressed<uint8_t>, grid, 128, 0, p, d_idx, (int)idx.size(), nF, -minH, -minW); }
	else if (img.dtype == STEVI_F32) { STEVI_LAUNCH(c, k_unfold_compressed<float>, grid, 128, 0, p, d_idx, (int)idx.size(), nF, -minH, -minW); }
	else return fail(c, STEVI_E_DTYPE, "unfold_compressed: image must be u8 or f32");
	return STEVI_OK;
}

int32_t census_features_dev(stevi_ctx* c, const View& fv, int semantics, const View& out) {
	if (fv.ndim != 3) return fail(c, STEVI_E_SHAPE, "census_features: feature volume must be (H,W,F)");
	const int H = fv.shape[0], W = fv.shape[1], F = fv.shape[2];
	if (F <= 1) return fail(c, STEVI_E_SHAPE, "census_features: fewer than 2 features (census.h:75-77)");
	const int nw = (F - 1) / 32 + 1;
	const int n_written = semantics == STEVI_CORRECTED ? nw : (F - 1) / 32;
	if (out.ndim != 3 || out.dtype != STEVI_U32 || out.shape[0] != H || out.shape[1] != W || out.shape[2] != nw)
		return fail(c, STEVI_E_SHAPE, "census_features: out must be (%d,%d,%d) u32", H, W, nw);
	if (H == 0 || W == 0) return STEVI_OK;
	if (H > 65535) return fail(c, STEVI_E_UNSUPPORTED, "census_features: more than 65535 rows");
	dim3 grid(cdiv(W, 128), H);
	if (fv.dtype == STEVI_U8) { STEVI_LAUNCH(c, k_census_features<uint8_t>, grid, 128, 0, fv.as<uint8_t>(), fv.st[0], fv.st[1], fv.st[2], H, W, F, nw, n_written, out.as<uint32_t>(), out.st[0], out.st[1], out.st[2]); }
	else if (fv.dtype == STEVI_F32) { STEVI_LAUNCH(c, k_census_features<float>, grid, 128, 0, fv.as<float>(), fv.st[0], fv.st[1], fv.st[2], H, W, F, nw, n_written, out.as<uint32_t>(), out.st[0], out.st[1], out.st[2]); }
	else return fail(c, STEVI_E_DTYPE, "census_features: features must be u8 or f32");
	return STEVI_OK;
}

int32_t feature_volume_dev(stevi_ctx* c, int mf, const View& img, int hr, int vr, const View& out) {
	bool zm = false, nm = false;
	switch (mf) {
	case STEVI_MF_CC: case STEVI_MF_SSD: case STEVI_MF_SAD: break;
	case STEVI_MF_NCC: nm = true; break;
	case STEVI_MF_ZNCC: zm = nm = true; break;
	case STEVI_MF_ZCC: case STEVI_MF_ZSSD: case STEVI_MF_ZSAD: zm = true; break;
	default: return fail(c, STEVI_E_UNSUPPORTED, "feature_volume: matching function %d (census features: stevi_census_transform)", mf);
	}
	if (img.ndim != 2 || img.dtype != STEVI_F32) return fail(c, STEVI_E_UNSUPPORTED, "feature_volume: image must be (H,W) f32");
	if (hr < 0 || vr < 0 || hr > 127 || vr > 127) return fail(c, STEVI_E_ARG, "feature_volume: radii");
	const int H = img.shape[0], W = img.shape[1], F = (2 * hr + 1) * (2 * vr + 1);
	if (out.ndim != 3 || out.dtype != STEVI_F32 || out.shape[0] != H || out.shape[1] != W || out.shape[2] != F)
		return fail(c, STEVI_E_SHAPE, "feature_volume: out must be (%d,%d,%d) f32", H, W, F);
	if (H == 0 || W == 0) return STEVI_OK;
	if (H > 65535) return fail(c, STEVI_E_UNSUPPORTED, "feature_volume: more than 65535 rows");
	dim3 grid(cdiv(W, 128), H);
	STEVI_LAUNCH(c, k_feature_volume, grid, 128, 0, img.as<float>(), img.st[0], img.st[1], H, W, hr, vr, zm ? 1 : 0, nm ? 1 : 0,
				 out.as<float>(), out.st[0], out.st[1], out.st[2]);
	return STEVI_OK;
}

} // namespace stevi
