/* TEST INFRASTRUCTURE — not product code.
 *
 * Least-squares solve by Householder QR with full pivoting, in float: the published algorithm of
 * Eigen::FullPivHouseholderQR<MatrixXf>::solve (Eigen 3.3, a system package of the reference that is absent from this
 * image: the reference's optimisation headers — optimization/l2optimization.h:37-42, sphericaloptimization.h:52-62 — call it).
 * Steps: at step k the entry of largest magnitude of the remaining block is brought to (k,k) by a row and a column
 * exchange, a Householder reflector clears the column below it; the factorisation stops when the largest remaining entry is
 * negligible against the first pivot (epsilon * min(m,n)); the rank counts the diagonal entries of R above
 * epsilon * min(m,n) * largest pivot; the solution is the basic one (free unknowns 0), not the minimum-norm one.
 * Same estimator and same precision class as Eigen's; the bits differ from Eigen's (blocked kernels, other summation orders).
 *
 * Used by oracle/shims/Eigen/QR (so that the unmodified reference headers of image_based_refinement.h can RUN in
 * oracle/_ref) and by oracle/stevi_oracle.c (the restatement).
 */
#ifndef STEVI_ORACLE_SMALL_QR_H
#define STEVI_ORACLE_SMALL_QR_H
#include <math.h>
#include <float.h>
#include <stdlib.h>
#include <string.h>

/* A: m x n, column-major (A[i + j*m]); b: m; x: n. */
static inline void stevi_qr_lstsq(int m, int n, const float* A, const float* b, float* x) {
	const int size = m < n ? m : n;
	float* qr = (float*)malloc(sizeof(float) * (size_t)m * n + sizeof(float) * m + sizeof(int) * n);
	float* c = qr + (size_t)m * n;
	int* perm = (int*)(c + m);
	memcpy(qr, A, sizeof(float) * (size_t)m * n);
	memcpy(c, b, sizeof(float) * m);
	for (int j = 0; j < n; j++) perm[j] = j;
	const float precision = FLT_EPSILON * (float)size;
	float biggest = 0.f, maxpivot = 0.f;
	int nonzero = size;
	for (int k = 0; k < size; k++) {
		int pr = k, pc = k;
		float big = -1.f;
		for (int j = k; j < n; j++) for (int i = k; i < m; i++) {
			const float v = fabsf(qr[i + (size_t)j * m]);
			if (v > big) { big = v; pr = i; pc = j; }
		}
		if (k == 0) biggest = big;
		if (big <= fabsf(biggest) * precision) { nonzero = k; break; }
		if (pr != k) {
			for (int j = 0; j < n; j++) { const float t = qr[k + (size_t)j * m]; qr[k + (size_t)j * m] = qr[pr + (size_t)j * m]; qr[pr + (size_t)j * m] = t; }
			const float t = c[k]; c[k] = c[pr]; c[pr] = t;
		}
		if (pc != k) {
			for (int i = 0; i < m; i++) { const float t = qr[i + (size_t)k * m]; qr[i + (size_t)k * m] = qr[i + (size_t)pc * m]; qr[i + (size_t)pc * m] = t; }
			const int t = perm[k]; perm[k] = perm[pc]; perm[pc] = t;
		}
		/* reflector H = I - tau v v^T, v = (1, essential), taking column k to (beta, 0, ..) */
		float* col = qr + (size_t)k * m;
		const float c0 = col[k];
		float tail = 0.f;
		for (int i = k + 1; i < m; i++) tail += col[i] * col[i];
		float tau, beta;
		if (tail <= FLT_MIN) { tau = 0.f; beta = c0; for (int i = k + 1; i < m; i++) col[i] = 0.f; }
		else {
			beta = sqrtf(c0 * c0 + tail);
			if (c0 >= 0.f) beta = -beta;
			for (int i = k + 1; i < m; i++) col[i] /= (c0 - beta);
			tau = (beta - c0) / beta;
		}
		col[k] = beta;
		if (fabsf(beta) > maxpivot) maxpivot = fabsf(beta);
		if (tau != 0.f) {
			for (int j = k + 1; j <= n; j++) { /* j == n: the right-hand side */
				float* y = j < n ? qr + (size_t)j * m : c;
				float s = y[k];
				for (int i = k + 1; i < m; i++) s += col[i] * y[i];
				s *= tau;
				y[k] -= s;
				for (int i = k + 1; i < m; i++) y[i] -= s * col[i];
			}
		}
	}
	int rank = 0;
	const float thr = fabsf(maxpivot) * precision;
	for (int k = 0; k < nonzero; k++) rank += fabsf(qr[k + (size_t)k * m]) > thr;
	for (int j = 0; j < n; j++) x[j] = 0.f;
	for (int k = rank - 1; k >= 0; k--) { /* back substitution on the leading rank x rank block of R */
		float s = c[k];
		for (int j = k + 1; j < rank; j++) s -= qr[k + (size_t)j * m] * c[j];
		c[k] = s / qr[k + (size_t)k * m];
	}
	for (int k = 0; k < rank; k++) x[perm[k]] = c[k];
	free(qr);
}
#endif
