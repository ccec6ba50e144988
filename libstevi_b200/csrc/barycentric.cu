// Feature-domain subpixel refinement (correlation/image_based_refinement.h): the reference fits, per
// pixel, the best affine combination of the target feature vectors around the winner to the source feature vector
// (Optimization::affineBestL2Approximation / sphericalAffineBestApproximation / affineBestL1Approximation on F x (K+1) Eigen
// matrices, solved with FullPivHouseholderQR) and moves the disparity by the barycentre of the coefficients.
//
// Here eight lanes share a pixel: they stream the F features of the K+2 vectors once (32-byte sectors per vector), accumulate
// the Gram matrix of [targets, source] in fp64 (products of floats are exact in fp64), fold it with three shuffle steps, and
// every lane solves the K x K normal equations (Gaussian elimination with full pivoting, fp64; rank-deficient systems get the
// basic solution with the free unknowns at 0, like the rank-revealing QR). The least-squares coefficients, the scores of the
// interpolated vectors and the base score all follow from that one Gram matrix, so the feature volumes are read once.
// Same estimator as the reference, other rounding: results are held to a tolerance (tests/test_gpu_image_refinement.py).
#include "kernels.cuh"
#include <cfloat>

namespace stevi {
namespace {

constexpr int GL = 8;        // lanes per pixel
constexpr int BLOCK = 256;   // 32 pixels per block
constexpr int MAXK = 8;      // unknowns of the largest system (Queen neighbourhood around the centre)
constexpr int MAXN = MAXK + 2;

enum { FAM_SPHERICAL = 0, FAM_L2 = 1, FAM_L1 = 2, FAM_NONE = 3 };

__host__ __device__ inline int family_of(int mf) {
	switch (mf) {
	case STEVI_MF_NCC: case STEVI_MF_ZNCC: return FAM_SPHERICAL;  // matching_costs.h:435-439, :538-542
	case STEVI_MF_SSD: case STEVI_MF_ZSSD: return FAM_L2;         // :475-479, :561-565
	case STEVI_MF_SAD: case STEVI_MF_ZSAD: return FAM_L1;         // :498-502, :584-588
	default: return FAM_NONE;                                     // CC, ZCC: no barycentricBestApproximation in their traits
	}
}

__device__ __forceinline__ double gsum(double v) {
#pragma unroll
	for (int o = 1; o < GL; o <<= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
	return v;
}
__device__ __forceinline__ float gmin(float v) {
#pragma unroll
	for (int o = 1; o < GL; o <<= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
	return v;
}

// Gram matrix of N strided vectors, summed over the group; G is the full symmetric N x N matrix (leading dimension MAXN)
template<int N>
__device__ __forceinline__ void gram(const float* const (&p)[N], const int64_t (&st)[N], int F, int lane, double (*G)[MAXN]) {
	double acc[N * (N + 1) / 2];
#pragma unroll
	for (int e = 0; e < N * (N + 1) / 2; e++) acc[e] = 0.0;
	for (int c = lane; c < F; c += GL) {
		double v[N];
#pragma unroll
		for (int k = 0; k < N; k++) v[k] = (double)p[k][c * st[k]];
		int e = 0;
#pragma unroll
		for (int a = 0; a < N; a++)
#pragma unroll
			for (int b = a; b < N; b++, e++) acc[e] = fma(v[a], v[b], acc[e]);
	}
	int e = 0;
#pragma unroll
	for (int a = 0; a < N; a++)
#pragma unroll
		for (int b = a; b < N; b++, e++) { const double s = gsum(acc[e]); G[a][b] = s; G[b][a] = s; }
}

// A x = b for a symmetric positive semi-definite A (a Gram matrix), N known at compile time so that everything stays in
// registers: Gaussian elimination without exchanges (safe on Gram matrices); a pivot that is negligible against the largest
// diagonal entry marks a dependent column, whose unknown stays 0 — the basic solution a rank-revealing QR gives. A and b are destroyed.
template<int N>
__device__ __forceinline__ void solve_gram(double (&A)[N][N], double (&b)[N], double (&x)[N]) {
	double maxd = 0.0;
#pragma unroll
	for (int k = 0; k < N; k++) maxd = fmax(maxd, fabs(A[k][k]));
	const double thr = maxd * ((double)FLT_EPSILON * N) * ((double)FLT_EPSILON * N);
	double inv[N];
#pragma unroll
	for (int k = 0; k < N; k++) {
		inv[k] = fabs(A[k][k]) > thr ? 1.0 / A[k][k] : 0.0;
#pragma unroll
		for (int i = k + 1; i < N; i++) {
			const double f = A[i][k] * inv[k];
#pragma unroll
			for (int j = k + 1; j < N; j++) A[i][j] -= f * A[k][j];
			b[i] -= f * b[k];
		}
	}
#pragma unroll
	for (int k = N - 1; k >= 0; k--) {
		double s = b[k];
#pragma unroll
		for (int j = k + 1; j < N; j++) s -= A[k][j] * x[j];
		x[k] = s * inv[k];
	}
}

// Coefficients of the best affine combination of the columns cols[0..K] of the Gram matrix (the last one is the origin of the
// affine space: optimization/affine_utils.h:44-57 with referenceId = -1) to the source vector ib.
// L2: l2optimization.h:48-63. Spherical: sphericaloptimization.h:35-63, every product it forms expressed through G.
template<int K>
__device__ __forceinline__ void bary_coeffs(int family, const int* cols, int ib, const double (*G)[MAXN], float* coeffs) {
	double g[K + 1][K + 1], gb[K + 1];   // Gram matrix of the K+1 columns, and their products with the source
#pragma unroll
	for (int a = 0; a <= K; a++) {
		gb[a] = G[cols[a]][ib];
#pragma unroll
		for (int b = 0; b <= K; b++) g[a][b] = G[cols[a]][cols[b]];
	}
	const double gKK = g[K][K];
	double P[K][K], q[K], r[K], alpha[K];
#pragma unroll
	for (int a = 0; a < K; a++) {
		q[a] = g[a][K] - gKK; // M^T a_K
#pragma unroll
		for (int b = 0; b < K; b++) P[a][b] = g[a][b] - g[a][K] - g[b][K] + gKK;
	}
	if (family == FAM_L2) {
#pragma unroll
		for (int a = 0; a < K; a++) r[a] = (gb[a] - gb[K]) - q[a];
		solve_gram<K>(P, r, alpha);
	} else {
		// bPerp = A lstsq(A, b)
		double ga[K + 1][K + 1], ra[K + 1], xa[K + 1];
#pragma unroll
		for (int a = 0; a <= K; a++) {
			ra[a] = gb[a];
#pragma unroll
			for (int b = 0; b <= K; b++) ga[a][b] = g[a][b];
		}
		solve_gram<K + 1>(ga, ra, xa);
		// aPerp = a_K - M lstsq(M, a_K)
		double P2[K][K], y[K];
#pragma unroll
		for (int a = 0; a < K; a++) {
			r[a] = q[a];
#pragma unroll
			for (int b = 0; b < K; b++) P2[a][b] = P[a][b];
		}
		solve_gram<K>(P2, r, y);
		double aa = gKK, ab = 0.0, w[K];
#pragma unroll
		for (int a = 0; a < K; a++) {
			double s = 0.0;
#pragma unroll
			for (int b = 0; b < K; b++) s += P[a][b] * y[b];
			aa += y[a] * (s - 2.0 * q[a]);
			w[a] = 0.0;
		}
#pragma unroll
		for (int i = 0; i <= K; i++) {
			double u = g[K][i]; // aPerp . a_i
#pragma unroll
			for (int a = 0; a < K; a++) { const double m = g[a][i] - g[K][i]; u -= y[a] * m; w[a] += xa[i] * m; }
			ab += xa[i] * u;
		}
		const double gg = aa / ab;
#pragma unroll
		for (int a = 0; a < K; a++) r[a] = gg * w[a] - q[a];
		solve_gram<K>(P, r, alpha);
	}
	float sum = 0.f;
#pragma unroll
	for (int a = 0; a < K; a++) { coeffs[a] = (float)alpha[a]; sum += coeffs[a]; }
	coeffs[K] = 1.f - sum; // affine_utils.h:83-97
}
// featureComparison(source / |source|, A coeffs / |A coeffs|) for the normalised functions, SumSquareDiff(source, A coeffs)
// otherwise (image_based_refinement.h:293-306), from the Gram matrix
__device__ double score_of(int family, int K, const int* cols, int ib, const double (*G)[MAXN], const float* coeffs) {
	double num = 0.0, quad = 0.0;
	for (int a = 0; a <= K; a++) {
		num += (double)coeffs[a] * G[cols[a]][ib];
		for (int b = 0; b <= K; b++) quad += (double)coeffs[a] * (double)coeffs[b] * G[cols[a]][cols[b]];
	}
	if (family == FAM_SPHERICAL) return num / (sqrt(G[ib][ib]) * sqrt(quad));
	return G[ib][ib] - 2.0 * num + quad;
}
__device__ __forceinline__ bool better(int family, double cand, double best) { return family == FAM_SPHERICAL ? cand > best : cand < best; }

struct Bary {
	const float* src; int64_t s0, s1, s2;
	const float* tgt; int64_t t0, t1, t2;
	const int32_t* idx; int64_t i0, i1, i2;
	float* out; int64_t o0, o1, o2;
	int H, W, F, sign, family, disp_width, radius;
	int lo0, up0, lo1, up1, contiguity, symmetric;
};

// refineBarycentricDisp, image_based_refinement.h:184-355 (NCC, ZNCC, SSD, ZSSD)
__global__ void __launch_bounds__(BLOCK) k_bary_1d(const Bary p) {
	const int lane = threadIdx.x & (GL - 1);
	const int j = blockIdx.x * (BLOCK / GL) + threadIdx.x / GL, i = blockIdx.y;
	const bool in = j < p.W;
	const int d = in ? p.idx[i * p.i0 + j * p.i1] : 0;
	const int jd = j + p.sign * d;
	const bool work = in && !(jd < 1 || jd + 1 >= p.W); // :214
	const float* t = p.tgt + i * p.t0 + (int64_t)(work ? jd : 0) * p.t1;
	const float* const vec[4] = {t - (work ? p.t1 : 0), t, t + (work ? p.t1 : 0), p.src + i * p.s0 + (int64_t)(in ? j : 0) * p.s1};
	const int64_t st[4] = {p.t2, p.t2, p.t2, p.s2};
	double G[MAXN][MAXN];
	gram<4>(vec, st, work ? p.F : 0, lane, G);
	// lane 0 fits the pair (target(jd), target(jd+1)), lane 1 the pair (target(jd-1), target(jd)); lane 0 decides
	const int cp[2] = {1, 2}, cm[2] = {0, 1}, c0[1] = {1};
	float frac = 0.f;
	double sc = 0.0;
	bool adm = false;
	if (work && lane < 2) {
		const int* cols = lane == 0 ? cp : cm;
		float k[2];
		bary_coeffs<1>(p.family, cols, 3, G, k);
		frac = lane == 0 ? k[1] : k[0];
		adm = frac > 0.f && frac < 1.f;
		if (adm) sc = score_of(p.family, 1, cols, 3, G, k);
	}
	const bool adm_m = __shfl_sync(0xffffffffu, (int)adm, 1, GL) != 0;
	const double sc_m = __shfl_sync(0xffffffffu, sc, 1, GL);
	const float frac_m = __shfl_sync(0xffffffffu, frac, 1, GL);
	if (!in || lane != 0) return;
	float delta = 0.f;
	if (work) {
		const float one = 1.f;
		double score = score_of(p.family, 0, c0, 3, G, &one);
		if (adm && better(p.family, sc, score)) { score = sc; delta = frac; }
		if (adm_m && better(p.family, sc_m, score)) { score = sc_m; delta = -frac_m; }
	}
	p.out[i * p.o0 + j * p.o1] = __fadd_rn((float)d, __fmul_rn((float)p.sign, delta)); // :347
}

// refineBarycentricDisp for SAD / ZSAD: with two columns affineBestL1Approximation is the one-unknown overload of
// leastAbsoluteDifferences (l1optimization.h:257-283), the weighted median of the break points (b - a1)/(a0 - a1) with weights
// |a0 - a1| (weighted_median_index, :33-60: the first break point, in ascending order, at which twice the accumulated weight
// passes the total). Its not-a-number test looks at t_breaks.back() — the last feature only — so only that entry is reset to
// (-inf, weight 0); other 0/0 entries stay NaN in the reference's std::sort (unspecified there); here they never count.
// Shared memory: break points and weights of the block's 32 pixels.
__device__ float l1_two_columns(const float* a0, const float* a1, const float* b, int64_t ts, int64_t bs, int F, int lane, float* tb, float* tw) {
	for (int c = lane; c < F; c += GL) {
		const float o = __fsub_rn(a0[c * ts], a1[c * ts]);
		float w = fabsf(o), t = __fdiv_rn(__fsub_rn(b[c * bs], a1[c * ts]), o);
		if (c == F - 1 && (isnan(t) || isinf(t))) { w = 0.f; t = __int_as_float(0xff800000); }
		tb[c] = t; tw[c] = w;
	}
	__syncwarp();
	double tot = 0.0;
	for (int c = lane; c < F; c += GL) tot += (double)tw[c];
	const float total = (float)gsum(tot);
	float best = __int_as_float(0x7f800000);
	for (int c = lane; c < F; c += GL) {
		const float t = tb[c];
		double acc = 0.0;
		for (int e = 0; e < F; e++) acc += tb[e] <= t ? (double)tw[e] : 0.0;
		if (2.f * (float)acc > total) best = fminf(best, t);
	}
	best = gmin(best);
	__syncwarp();
	return best == __int_as_float(0x7f800000) ? __int_as_float(0x7fc00000) : best;
}

__global__ void __launch_bounds__(BLOCK) k_bary_1d_l1(const Bary p) {
	extern __shared__ float sm[];
	const int lane = threadIdx.x & (GL - 1), g = threadIdx.x / GL;
	float* tb = sm + (size_t)g * 2 * p.F;
	float* tw = tb + p.F;
	const int j = blockIdx.x * (BLOCK / GL) + g, i = blockIdx.y;
	const bool in = j < p.W;
	const int d = in ? p.idx[i * p.i0 + j * p.i1] : 0;
	const int jd = j + p.sign * d;
	const bool work = in && !(jd < 1 || jd + 1 >= p.W);
	const int F = work ? p.F : 0;
	const float* t0 = p.tgt + i * p.t0 + (int64_t)(work ? jd : 0) * p.t1;
	const float* tm = t0 - (work ? p.t1 : 0), * tp = t0 + (work ? p.t1 : 0);
	const float* b = p.src + i * p.s0 + (int64_t)(in ? j : 0) * p.s1;
	// plus: A = [t0, tp], origin tp; minus: A = [tm, t0], origin t0
	const float ap = l1_two_columns(t0, tp, b, p.t2, p.s2, F, lane, tb, tw);
	const float am = l1_two_columns(tm, t0, b, p.t2, p.s2, F, lane, tb, tw);
	const float kp[2] = {ap, __fsub_rn(1.f, ap)}, km[2] = {am, __fsub_rn(1.f, am)};
	double s0 = 0.0, sp = 0.0, smn = 0.0;
	for (int c = lane; c < F; c += GL) {
		const float vb = b[c * p.s2], v0 = t0[c * p.t2], vm = tm[c * p.t2], vp = tp[c * p.t2];
		s0 += (double)fabsf(__fsub_rn(vb, v0));
		sp += (double)fabsf(__fsub_rn(vb, __fadd_rn(__fmul_rn(v0, kp[0]), __fmul_rn(vp, kp[1]))));
		smn += (double)fabsf(__fsub_rn(vb, __fadd_rn(__fmul_rn(vm, km[0]), __fmul_rn(v0, km[1]))));
	}
	s0 = gsum(s0); sp = gsum(sp); smn = gsum(smn);
	if (!in || lane != 0) return;
	float delta = 0.f;
	if (work) {
		double score = s0;
		const float dplus = kp[1], dminus = km[0];
		if (dplus > 0.f && dplus < 1.f && sp < score) { score = sp; delta = dplus; }
		if (dminus > 0.f && dminus < 1.f && smn < score) { score = smn; delta = -dminus; }
	}
	p.out[i * p.o0 + j * p.o1] = __fadd_rn((float)d, __fmul_rn((float)p.sign, delta));
}

// refineBarycentricSymmetricDisp<matchFunc, R>, image_based_refinement.h:31-105: one system over the 2R+1 target columns
template<int R>
__global__ void __launch_bounds__(BLOCK) k_bary_sym_1d(const Bary p) {
	constexpr int N = 2 * R + 2;
	const int lane = threadIdx.x & (GL - 1);
	const int j = blockIdx.x * (BLOCK / GL) + threadIdx.x / GL, i = blockIdx.y;
	const bool in = j < p.W;
	const int d = in ? p.idx[i * p.i0 + j * p.i1] : 0;
	const int jd = j + p.sign * d;
	// :64-69 (the first test can only fire on the last column)
	const bool work = in && !(j + 1 >= p.W) && !(jd - R < 0 || jd + 1 > p.W - R) && !(d == 0 || d + 1 >= p.disp_width);
	const float* t = p.tgt + i * p.t0 + (int64_t)(work ? jd : 0) * p.t1;
	const float* vec[N];
	int64_t st[N];
#pragma unroll
	for (int k = 0; k < N - 1; k++) { vec[k] = t + (work ? (k - R) * p.t1 : 0); st[k] = p.t2; }
	vec[N - 1] = p.src + i * p.s0 + (int64_t)(in ? j : 0) * p.s1; st[N - 1] = p.s2;
	double G[MAXN][MAXN];
	gram<N>(vec, st, work ? p.F : 0, lane, G);
	if (!in || lane != 0) return;
	float res = (float)d;
	if (work) {
		int cols[N - 1];
		float co[N - 1];
		for (int k = 0; k < N - 1; k++) cols[k] = k;
		bary_coeffs<N - 2>(p.family, cols, N - 1, G, co);
		float delta = 0.f;
		for (int k = 0; k < N - 1; k++) delta = __fadd_rn(delta, __fmul_rn(co[k], (float)(k - R))); // :88-91
		if (fabsf(delta) < 1.f) res = __fadd_rn((float)d, delta);
	}
	p.out[i * p.o0 + j * p.o1] = res;
}

// refineBarycentric2dDisp (image_based_refinement.h:675-880) and refineBarycentricSymmetric2dDisp (:1068-1241): the Gram
// matrix of the 3 x 3 target neighbourhood (index 3*(di+1) + (dj+1)) and the source serves every system of the pixel.
__constant__ int c_corner[3][3][2] = {{{1, 1}, {1, 0}, {0, 1}}, {{1, 0}, {0, 1}, {0, 0}}, {{1, 1}, {0, 0}, {0, 0}}};          // utils/contiguity.h:139-164
__constant__ int c_ncorner[3] = {3, 2, 1};
__constant__ int c_dirs[3][8][2] = {{{1, 1}, {1, 0}, {1, -1}, {0, 1}, {0, -1}, {-1, 1}, {-1, 0}, {-1, -1}},                 // :98-135
									 {{1, 0}, {0, 1}, {0, -1}, {-1, 0}, {0, 0}, {0, 0}, {0, 0}, {0, 0}},
									 {{1, 1}, {1, -1}, {-1, 1}, {-1, -1}, {0, 0}, {0, 0}, {0, 0}, {0, 0}}};
__constant__ int c_ndirs[3] = {8, 4, 4};

template<int K>
__global__ void __launch_bounds__(BLOCK) k_bary_2d(const Bary p) {
	const int lane = threadIdx.x & (GL - 1);
	const int j = blockIdx.x * (BLOCK / GL) + threadIdx.x / GL, i = blockIdx.y;
	const bool in = j < p.W;
	const int d0 = in ? p.idx[i * p.i0 + j * p.i1] : 0, d1 = in ? p.idx[i * p.i0 + j * p.i1 + p.i2] : 0;
	const int id = i + d0, jd = j + d1;
	const bool work = in && !(id < 1 || id + 1 >= p.H) && !(jd < 1 || jd + 1 >= p.W) && !(d0 < p.lo0 || d0 > p.up0) && !(d1 < p.lo1 || d1 > p.up1); // :711-722
	const float* t = p.tgt + (int64_t)(work ? id : 0) * p.t0 + (int64_t)(work ? jd : 0) * p.t1;
	const float* vec[10];
	int64_t st[10];
#pragma unroll
	for (int a = 0; a < 3; a++)
#pragma unroll
		for (int b = 0; b < 3; b++) { vec[3 * a + b] = t + (work ? (a - 1) * p.t0 + (b - 1) * p.t1 : 0); st[3 * a + b] = p.t2; }
	vec[9] = p.src + i * p.s0 + (int64_t)(in ? j : 0) * p.s1; st[9] = p.s2;
	double G[MAXN][MAXN];
	gram<10>(vec, st, work ? p.F : 0, lane, G);
	// one system per lane (the four quadrants; a single one for the symmetric form); lane 0 takes them in the reference's order
	const int nsys = p.symmetric ? 1 : 4;
	float t0 = 0.f, t1 = 0.f;
	double sc = 0.0;
	bool adm = false;
	if (work && lane < nsys) {
		int cols[K + 1];
		float co[K + 1];
		const int s = lane;
		const int sx = p.symmetric ? 1 : (s < 2 ? 1 : -1), sy = p.symmetric ? 1 : ((s & 1) ? -1 : 1); // dirShifts = {1,-1}, x outer (:783-790)
		for (int k = 0; k < K; k++) {
			const int di = (p.symmetric ? c_dirs[p.contiguity][k][0] : c_corner[p.contiguity][k][0]) * sx;
			const int dj = (p.symmetric ? c_dirs[p.contiguity][k][1] : c_corner[p.contiguity][k][1]) * sy;
			cols[k] = 3 * (di + 1) + (dj + 1);
		}
		cols[K] = 4;
		bary_coeffs<K>(p.family, cols, 9, G, co);
		for (int k = 0; k < K; k++) {
			const int di = cols[k] / 3 - 1, dj = cols[k] % 3 - 1;
			t0 = __fadd_rn(t0, __fmul_rn(co[k], (float)di));
			t1 = __fadd_rn(t1, __fmul_rn(co[k], (float)dj));
		}
		adm = fabsf(t0) <= 1.f && fabsf(t1) <= 1.f; // :834
		if (adm) sc = score_of(p.family, K, cols, 9, G, co);
	}
	float dd0 = 0.f, dd1 = 0.f;
	double score = 0.0;
	if (work && lane == 0) { const int centre[1] = {4}; const float one = 1.f; score = score_of(p.family, 0, centre, 9, G, &one); }
	for (int q = 0; q < nsys; q++) {
		const bool a = __shfl_sync(0xffffffffu, (int)adm, q, GL) != 0;
		const double v = __shfl_sync(0xffffffffu, sc, q, GL);
		const float e0 = __shfl_sync(0xffffffffu, t0, q, GL), e1 = __shfl_sync(0xffffffffu, t1, q, GL);
		if (a && better(p.family, v, score)) { score = v; dd0 = e0; dd1 = e1; }
	}
	if (!in || lane != 0) return;
	float* o = p.out + i * p.o0 + j * p.o1;
	o[0] = __fadd_rn((float)d0, dd0);
	o[p.o2] = __fadd_rn((float)d1, dd1);
}

// refineCostSymmetricDisp, image_based_refinement.h:1591-1690: the parabola through the winner's costs, averaged with the
// parabola through the comparisons of the half-pixel source vector (mean of the source and its neighbour on the side the first
// estimate points away from) with the three target vectors.
struct CostSym {
	Bary b;
	const float* cv; int64_t c0, c1, c2;
	int cvW, cvD, mf;
};

__global__ void __launch_bounds__(BLOCK) k_cost_symmetric(const CostSym q) {
	const Bary& p = q.b;
	const int lane = threadIdx.x & (GL - 1);
	const int j = blockIdx.x * (BLOCK / GL) + threadIdx.x / GL, i = blockIdx.y;
	const bool in = j < p.W;
	const int d = in ? p.idx[i * p.i0 + j * p.i1] : 0;
	const int jd = j + p.sign * d;
	float delta = 0.f;
	int dir = 1;
	const bool first = in && j > 1 && j + 1 < q.cvW && d > 0 && d + 1 < q.cvD; // :1623
	if (first) {
		const float* c = q.cv + i * q.c0 + j * q.c1 + (int64_t)d * q.c2;
		const float cm1 = c[-q.c2], c0 = c[0], c1 = c[q.c2];
		delta = __fdiv_rn(__fsub_rn(cm1, c1), __fmul_rn(2.f, __fadd_rn(__fsub_rn(c1, __fmul_rn(2.f, c0)), cm1)));
		if (delta > 0.f) dir = -1;
	}
	const bool second = first && jd + 1 < q.cvW && jd > 1; // :1637
	const int F = second ? p.F : 0;
	const float* s = p.src + i * p.s0 + (int64_t)(in ? j : 0) * p.s1;
	const float* sn = s + (second ? dir * p.s1 : 0);
	const float* t0 = p.tgt + i * p.t0 + (int64_t)(second ? jd : 0) * p.t1;
	const float* tm = t0 - (second ? p.t1 : 0), * tp = t0 + (second ? p.t1 : 0);
	const bool dot = q.mf == STEVI_MF_CC || q.mf == STEVI_MF_NCC || q.mf == STEVI_MF_ZCC || q.mf == STEVI_MF_ZNCC;
	const bool sq = q.mf == STEVI_MF_SSD || q.mf == STEVI_MF_ZSSD;
	const bool normalized = q.mf == STEVI_MF_NCC || q.mf == STEVI_MF_ZNCC;
	double a[3] = {0.0, 0.0, 0.0}, n[4] = {0.0, 0.0, 0.0, 0.0};
	for (int c = lane; c < F; c += GL) {
		const float v = (float)(0.5 * (double)s[c * p.s2] + 0.5 * (double)sn[c * p.s2]); // :1642
		const float t[3] = {tm[c * p.t2], t0[c * p.t2], tp[c * p.t2]};
		n[3] += (double)v * (double)v;
#pragma unroll
		for (int k = 0; k < 3; k++) {
			n[k] += (double)t[k] * (double)t[k];
			if (dot) a[k] += (double)v * (double)t[k];
			else { const double e = (double)v - (double)t[k]; a[k] += sq ? e * e : fabs(e); }
		}
	}
#pragma unroll
	for (int k = 0; k < 3; k++) { a[k] = gsum(a[k]); n[k] = gsum(n[k]); }
	n[3] = gsum(n[3]);
	if (!in || lane != 0) return;
	if (second) {
		float f[3];
		for (int k = 0; k < 3; k++) f[k] = (float)(normalized ? a[k] / (sqrt(n[3]) * sqrt(n[k])) : a[k]);
		const float par = __fdiv_rn(__fsub_rn(f[0], f[2]), __fmul_rn(2.f, __fadd_rn(__fsub_rn(f[2], __fmul_rn(2.f, f[1])), f[0])));
		const float delta2 = (float)((double)par - dir * 0.5); // :1676
		if (fabsf(delta2) < 1.f) delta = __fdiv_rn(__fadd_rn(delta, delta2), 2.f);
	}
	p.out[i * p.o0 + j * p.o1] = __fadd_rn((float)d, delta);
}

int32_t fill(stevi_ctx* c, const char* who, const View& fl, const View& fr, const View& idx, int idx_dims, int direction, const View& out, Bary& p) {
	if (fl.ndim != 3 || fr.ndim != 3 || fl.dtype != STEVI_F32 || fr.dtype != STEVI_F32)
		return fail(c, STEVI_E_DTYPE, "%s: feature volumes must be (H,W,F) f32", who);
	const View& src = direction == STEVI_RIGHT_TO_LEFT ? fr : fl;
	const View& tgt = direction == STEVI_RIGHT_TO_LEFT ? fl : fr;
	const int H = src.shape[0], W = src.shape[1], F = src.shape[2];
	if (tgt.shape[0] != H || tgt.shape[2] != F || tgt.shape[1] < W)
		return fail(c, STEVI_E_SHAPE, "%s: target feature volume (%d,%d,%d) does not cover the source's (%d,%d,%d)", who, tgt.shape[0], tgt.shape[1], tgt.shape[2], H, W, F);
	if (idx.dtype != STEVI_I32 || idx.ndim != idx_dims || idx.shape[0] != H || idx.shape[1] != W || (idx_dims == 3 && idx.shape[2] != 2))
		return fail(c, STEVI_E_SHAPE, "%s: index map must be (%d,%d%s) i32", who, H, W, idx_dims == 3 ? ",2" : "");
	if (out.dtype != STEVI_F32 || out.ndim != idx_dims || out.shape[0] != H || out.shape[1] != W || (idx_dims == 3 && out.shape[2] != 2))
		return fail(c, STEVI_E_SHAPE, "%s: out must be (%d,%d%s) f32", who, H, W, idx_dims == 3 ? ",2" : "");
	if (H > 65535) return fail(c, STEVI_E_UNSUPPORTED, "%s: more than 65535 rows", who);
	p = Bary{};
	p.src = src.as<float>(); p.s0 = src.st[0]; p.s1 = src.st[1]; p.s2 = src.st[2];
	p.tgt = tgt.as<float>(); p.t0 = tgt.st[0]; p.t1 = tgt.st[1]; p.t2 = tgt.st[2];
	p.idx = idx.as<int32_t>(); p.i0 = idx.st[0]; p.i1 = idx.st[1]; p.i2 = idx_dims == 3 ? idx.st[2] : 0;
	p.out = out.as<float>(); p.o0 = out.st[0]; p.o1 = out.st[1]; p.o2 = idx_dims == 3 ? out.st[2] : 0;
	p.H = H; p.W = W; p.F = F;
	p.sign = direction == STEVI_RIGHT_TO_LEFT ? 1 : -1;
	return STEVI_OK;
}

} // namespace

int32_t refine_barycentric_dev(stevi_ctx* c, int mf, const View& fl, const View& fr, const View& idx, int direction, const View& out) {
	Bary p;
	STEVI_TRY(fill(c, "refine_barycentric_disp", fl, fr, idx, 2, direction, out, p));
	p.family = family_of(mf);
	if (p.family == FAM_NONE) return fail(c, STEVI_E_UNSUPPORTED, "refine_barycentric_disp: matching function %d has no barycentric approximation (matching_costs.h:419-589)", mf);
	if (p.H == 0 || p.W == 0) return STEVI_OK;
	dim3 grid(cdiv(p.W, BLOCK / GL), p.H);
	if (p.family == FAM_L1) {
		const size_t smem = (size_t)(BLOCK / GL) * 2 * p.F * sizeof(float);
		if (smem > 48 * 1024) return fail(c, STEVI_E_UNSUPPORTED, "refine_barycentric_disp: more than %d features with SAD / ZSAD", 48 * 1024 / (BLOCK / GL) / 8);
		STEVI_LAUNCH(c, k_bary_1d_l1, grid, BLOCK, smem, p);
	} else STEVI_LAUNCH(c, k_bary_1d, grid, BLOCK, 0, p);
	return STEVI_OK;
}

int32_t refine_barycentric_symmetric_dev(stevi_ctx* c, int mf, int radius, const View& fl, const View& fr, const View& idx, int disp_width,
										 int direction, const View& out) {
	Bary p;
	STEVI_TRY(fill(c, "refine_barycentric_symmetric_disp", fl, fr, idx, 2, direction, out, p));
	p.family = family_of(mf);
	if (p.family == FAM_NONE) return fail(c, STEVI_E_UNSUPPORTED, "refine_barycentric_symmetric_disp: matching function %d has no barycentric approximation", mf);
	if (p.family == FAM_L1) return fail(c, STEVI_E_UNSUPPORTED, "refine_barycentric_symmetric_disp: SAD / ZSAD with more than two columns run the reference's randomised "
										"descent (l1optimization.h:98-252, Eigen's Random()): not reproducible, not provided");
	if (radius < 1 || radius > 3) return fail(c, STEVI_E_UNSUPPORTED, "refine_barycentric_symmetric_disp: refineRadius %d (1..3 are instantiated)", radius);
	p.disp_width = disp_width; p.radius = radius;
	if (p.H == 0 || p.W == 0) return STEVI_OK;
	dim3 grid(cdiv(p.W, BLOCK / GL), p.H);
	if (radius == 1) STEVI_LAUNCH(c, k_bary_sym_1d<1>, grid, BLOCK, 0, p);
	else if (radius == 2) STEVI_LAUNCH(c, k_bary_sym_1d<2>, grid, BLOCK, 0, p);
	else STEVI_LAUNCH(c, k_bary_sym_1d<3>, grid, BLOCK, 0, p);
	return STEVI_OK;
}

int32_t refine_barycentric_2d_dev(stevi_ctx* c, int mf, int contiguity, int symmetric, const View& fl, const View& fr, const View& idx,
								  const int32_t* search, int direction, const View& out) {
	Bary p;
	STEVI_TRY(fill(c, "refine_barycentric_2d_disp", fl, fr, idx, 3, direction, out, p));
	p.family = family_of(mf);
	if (p.family == FAM_NONE) return fail(c, STEVI_E_UNSUPPORTED, "refine_barycentric_2d_disp: matching function %d has no barycentric approximation", mf);
	if (contiguity < 0 || contiguity > 2) return fail(c, STEVI_E_ARG, "refine_barycentric_2d_disp: contiguity must be Queen (0), Rook (1) or Bishop (2)");
	if (p.family == FAM_L1) return fail(c, STEVI_E_UNSUPPORTED, "refine_barycentric_2d_disp: SAD / ZSAD run the reference's randomised descent (l1optimization.h:98-252): not provided");
	if (!search) return fail(c, STEVI_E_ARG, "refine_barycentric_2d_disp: search window missing");
	const View& tgt = direction == STEVI_RIGHT_TO_LEFT ? fl : fr;
	if (tgt.shape[1] != p.W) return fail(c, STEVI_E_SHAPE, "refine_barycentric_2d_disp: the feature volumes must have one size");
	p.lo0 = search[0]; p.up0 = search[1]; p.lo1 = search[2]; p.up1 = search[3];
	p.contiguity = contiguity; p.symmetric = symmetric ? 1 : 0;
	if (p.H == 0 || p.W == 0) return STEVI_OK;
	dim3 grid(cdiv(p.W, BLOCK / GL), p.H);
	const int K = symmetric ? (contiguity == STEVI_QUEEN ? 8 : 4) : (contiguity == STEVI_QUEEN ? 3 : contiguity == STEVI_ROOK ? 2 : 1);
	switch (K) {
	case 1: STEVI_LAUNCH(c, k_bary_2d<1>, grid, BLOCK, 0, p); break;
	case 2: STEVI_LAUNCH(c, k_bary_2d<2>, grid, BLOCK, 0, p); break;
	case 3: STEVI_LAUNCH(c, k_bary_2d<3>, grid, BLOCK, 0, p); break;
	case 4: STEVI_LAUNCH(c, k_bary_2d<4>, grid, BLOCK, 0, p); break;
	default: STEVI_LAUNCH(c, k_bary_2d<8>, grid, BLOCK, 0, p); break;
	}
	return STEVI_OK;
}

int32_t refine_cost_symmetric_dev(stevi_ctx* c, int mf, const View& fl, const View& fr, const View& idx, const View& cv, int direction, const View& out) {
	CostSym q;
	STEVI_TRY(fill(c, "refine_cost_symmetric_disp", fl, fr, idx, 2, direction, out, q.b));
	if (mf < STEVI_MF_CC || mf > STEVI_MF_ZSAD) return fail(c, STEVI_E_UNSUPPORTED, "refine_cost_symmetric_disp: matching function %d", mf);
	if (cv.ndim != 3 || cv.dtype != STEVI_F32 || cv.shape[0] != q.b.H || cv.shape[1] > q.b.W)
		return fail(c, STEVI_E_SHAPE, "refine_cost_symmetric_disp: cost volume must be (%d, <= %d, D) f32", q.b.H, q.b.W);
	q.cv = cv.as<float>(); q.c0 = cv.st[0]; q.c1 = cv.st[1]; q.c2 = cv.st[2];
	q.cvW = cv.shape[1]; q.cvD = cv.shape[2]; q.mf = mf;
	if (q.b.H == 0 || q.b.W == 0) return STEVI_OK;
	dim3 grid(cdiv(q.b.W, BLOCK / GL), q.b.H);
	STEVI_LAUNCH(c, k_cost_symmetric, grid, BLOCK, 0, q);
	return STEVI_OK;
}

} // namespace stevi
