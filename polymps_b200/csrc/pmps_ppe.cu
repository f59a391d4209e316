// pmps_ppe.cu — K1/K2: pressure Poisson equation of the incompressible scheme.
// Replaces MpsPressure::solvePressurePoissonPnd + solveConjugateGradient + setZeroOnNegativePressure (reference
// src/MpsPressure.cpp:169-286, :424-445, :471-476) and the Eigen arithmetic under them (ConjugateGradient.h:27-91,
// BasicPreconditioners.h:67-91).
//
// Assembly: the reference pushes triplets serially and lets Eigen build a column-major matrix; its CG is told
// Lower|Upper on that storage, so it multiplies by the TRANSPOSE of the assembled matrix (identical unless a polygon
// wall's one-sided mirror test breaks symmetry).  Here one thread per slot c writes row c of A^T directly as a
// compacted ELL row (column = slot of particle i, value = A_ic) by evaluating row i's predicate from c's candidate
// list, plus the diagonal and the right-hand side of its own row.  Everything is in slot order, so x-gathers in the
// SpMV hit nearby addresses.
// Solve: Eigen's Jacobi-preconditioned CG recurrence, warm start from pressurePPE, threshold eps^2 |b|^2, maxIt 2N.
// Dot products are two-stage (per-block partials, then one block) so the iteration is deterministic run to run.
#include "pmps_device.cuh"
#include <cfloat>
#include <cmath>

#define PPE_T 128
#define RED_T 256

struct PpeView {
	uint32_t* col; double* val; uint16_t* cnt;   // A^T, compacted ELL in the candidate-list layout
	double* diag; double* r; double* p; double* t;
	double* partial;   // [3][nBlocks] block partial sums
	double* scal;      // 0 rhs2, 1 res2, 2 absNew, 3 absOld, 4 pt, 5 alpha, 6 beta, 7 thr ; BiCGSTAB: 2 rho, 3 rho_old, 5 alpha, 6 beta, 11 w, 12 |r0|^2 (8, 9 hold the reduced sums)
	double* r0; double* v; double* s; double* y; double* z;   // BiCGSTAB work vectors (r, p, t above are shared with CG)
};

template <bool POLY>
__global__ void __launch_bounds__(PPE_T) k_ppe_assemble(const __grid_constant__ DevView S, const __grid_constant__ KParams K, const __grid_constant__ PpeView M, int rowsOfA) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= S.n) return;
	const double4 pi = S.pos[i]; const int ti = pk_type(pi);
	if (ti == K.p.ghost) { M.cnt[i] = 0; M.diag[i] = 1.0; S.sourceTerm[i] = 0.0; S.pressurePPE[i] = 0.0; return; }
	const int bci = S.particleBC[i];
	const bool rowActive = !(bci == K.p.other || bci == K.p.surface);
	const bool colOK = bci == K.p.inner && ti != K.p.dummyWall;
	const double mix = POLY ? S.mirrorParticlePos.x[i] : 0.0, miy = POLY ? S.mirrorParticlePos.y[i] : 0.0, miz = POLY ? S.mirrorParticlePos.z[i] : 0.0;
	const size_t base = (size_t)(i >> 5) * 32 * (size_t)S.listCap + (i & 31);
	const uint32_t* L = S.list + (size_t)i * (size_t)S.listCap; uint32_t* C = M.col + base; double* V = M.val + base;
	const int cnt = S.listCnt[i];
	double sum = 0.0, b = 0.0; int nz = 0;
	// particles moved into a hole by an out-of-domain swap of this step are found only through the bucket their array element is still
	// listed in (hid_visible, pmps_device.cuh): iFindsJ gates row i's own terms, jFindsI the transposed entry
	const int nHid = S.hid[0]; const int myHid = nHid > 0 ? hid_cell_of_slot(S, i) : -1;
	for (int k = 0; k < cnt; k++) {
		const int j = (int)L[k];
		double4 pj = S.pos[j];
		bool iFindsJ = true, jFindsI = true;
		if (nHid > 0) { const int hc = hid_cell_of_slot(S, j); iFindsJ = hc == -1 || hid_visible(K, hc, pi); jFindsI = myHid == -1 || hid_visible(K, myHid, pj); }
		if (K.p.domainTypeBC == 1) { if (S.wrapHid[j]) iFindsJ = false; if (S.wrapHid[i]) jFindsI = false; }   // wrapped in this step: found by nobody (pmps_domain.cu)
		const double pl = K.p.domainTypeBC == 1 ? per_pair_shift(S, K, i, j) : 0.0;   // periodic domain: j as i sees it
		if (K.p.domainTypeBC == 1) per_move(K, pj, pl);
		const double rx = pj.x - pi.x, ry = pj.y - pi.y, rz = pj.z - pi.z; const double d2 = sq3_exact(rx, ry, rz);
		if (!(d2 < K.p.reL2)) continue;
		const double wL = wgt(sqrt(d2), K.p.reL, K.p.weightType); const double a = wL * K.p.coeffPPE;
		const int tj = pk_type(pj);
		// row i (src/MpsPressure.cpp:213-229)
		bool mine = rowActive && iFindsJ;
		if (POLY && mine) { const double mx = pj.x - mix, my = pj.y - miy, mz = pj.z - miz; mine = d2 < sq3_exact(mx, my, mz); }
		if (mine) {
			if (tj == K.p.dummyWall) b -= K.p.coeffPPE * (S.RHO[j] * (rx * K.p.gravityX + ry * K.p.gravityY + rz * K.p.gravityZ) * wL);
			else sum -= a;
		}
		// BiCGSTAB multiplies by A itself (BiCGSTAB.h:81,88): row i of A = the entries row i's own predicate admits (:220-228)
		if (rowsOfA) {
			if (mine && tj != K.p.dummyWall && S.particleBC[j] == K.p.inner) { C[(size_t)nz * 32] = (uint32_t)j; V[(size_t)nz * 32] = a; nz++; }
		}
		// column i of row j: A_ji exists if row j is active, i is a real neighbour of j, i is not a dummy and BC[i] == inner
		else if (colOK) {
			const int bcj = S.particleBC[j];
			bool theirs = !(bcj == K.p.other || bcj == K.p.surface) && jFindsI;
			if (POLY && theirs) {
				double4 pim = pi; if (K.p.domainTypeBC == 1) per_move(K, pim, -pl);   // i as j sees it
				const double ex = pim.x - S.mirrorParticlePos.x[j], ey = pim.y - S.mirrorParticlePos.y[j], ez = pim.z - S.mirrorParticlePos.z[j];
				theirs = d2 < sq3_exact(ex, ey, ez);
			}
			if (theirs) { C[(size_t)nz * 32] = (uint32_t)j; V[(size_t)nz * 32] = a; nz++; }
		}
	}
	M.cnt[i] = (uint16_t)nz;
	if (rowActive) {
		const double rho = pk_ptype(pi) == 1 ? K.p.DNS_FL1 : K.p.DNS_FL2;
		sum -= K.p.alphaCompressibility * rho / (K.p.timeStep * K.p.timeStep);
		M.diag[i] = sum;
		if (K.p.mpsType == 3) b += -K.p.coeffPPESource * rho * (S.pndki[i] - K.p.pndSmallZero) + (1.0 - K.p.relaxPND) * rho * S.velDivergence[i] / K.p.timeStep;   // src/MpsPressure.cpp:373
		else b += -K.p.coeffPPESource * rho * (S.pndSmall[i] - K.p.pndSmallZero);
	} else { M.diag[i] = 1.0; b = 0.0; }
	S.sourceTerm[i] = b;
}

// ---- block reduction helpers -------------------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
	#pragma unroll
	for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
	return v;
}
template <int T> __device__ __forceinline__ double block_sum(double v, double* sh) {
	v = warp_sum(v);
	if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
	__syncthreads();
	double s = 0.0;
	if (threadIdx.x < 32) { s = threadIdx.x < T / 32 ? sh[threadIdx.x] : 0.0; s = warp_sum(s); }
	__syncthreads();
	return s;   // valid in thread 0
}

__device__ __forceinline__ double ell_row_dot(const DevView& S, const PpeView& M, int i, const double* __restrict__ x) {
	const size_t base = (size_t)(i >> 5) * 32 * (size_t)S.listCap + (i & 31);
	const uint32_t* C = M.col + base; const double* V = M.val + base; const int nz = M.cnt[i];
	double s = M.diag[i] * x[i];
	for (int k = 0; k < nz; k++) s += V[(size_t)k * 32] * x[C[(size_t)k * 32]];
	return s;
}

// r = b - A^T x0 ; partial sums of |b|^2 and |r|^2
__global__ void __launch_bounds__(PPE_T) k_ppe_residual0(const __grid_constant__ DevView S, const __grid_constant__ PpeView M, int nBlocks) {
	__shared__ double sh[PPE_T / 32];
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	double b2 = 0.0, r2 = 0.0;
	if (i < S.n) {
		const double b = S.sourceTerm[i]; const double r = b - ell_row_dot(S, M, i, S.pressurePPE);
		M.r[i] = r; if (owned_slot(S, i)) { b2 = b * b; r2 = r * r; }
	}
	const double sb = block_sum<PPE_T>(b2, sh); const double sr = block_sum<PPE_T>(r2, sh);
	if (threadIdx.x == 0) { M.partial[blockIdx.x] = sb; M.partial[nBlocks + blockIdx.x] = sr; }
}
// p = M^-1 r ; partial sums of r.p
__global__ void __launch_bounds__(PPE_T) k_ppe_init_p(const __grid_constant__ DevView S, const __grid_constant__ PpeView M, int nBlocks) {
	__shared__ double sh[PPE_T / 32];
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	double v = 0.0;
	if (i < S.n) { const double d = M.diag[i]; const double r = M.r[i]; const double p = (d != 0.0 ? 1.0 / d : 1.0) * r; M.p[i] = p; if (owned_slot(S, i)) v = r * p; }
	const double s = block_sum<PPE_T>(v, sh);
	if (threadIdx.x == 0) M.partial[blockIdx.x] = s;
}
// t = A^T p ; partial sums of p.t
__global__ void __launch_bounds__(PPE_T) k_ppe_spmv(const __grid_constant__ DevView S, const __grid_constant__ PpeView M, int nBlocks) {
	__shared__ double sh[PPE_T / 32];
	if (M.scal[14] != 0.0) return;   // (one-GPU batches: converged earlier in the batch)
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	double v = 0.0;
	if (i < S.n) { const double t = ell_row_dot(S, M, i, M.p); M.t[i] = t; if (owned_slot(S, i)) v = M.p[i] * t; }
	const double s = block_sum<PPE_T>(v, sh);
	if (threadIdx.x == 0) M.partial[blockIdx.x] = s;
}
// x += alpha p ; r -= alpha t ; partial sums of |r|^2 and r.(M^-1 r)
__global__ void __launch_bounds__(PPE_T) k_ppe_update_xr(const __grid_constant__ DevView S, const __grid_constant__ PpeView M, int nBlocks) {
	__shared__ double sh[PPE_T / 32];
	if (M.scal[14] != 0.0) return;
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	const double alpha = M.scal[5];
	double r2 = 0.0, rz = 0.0;
	if (i < S.n) {
		S.pressurePPE[i] += alpha * M.p[i];
		const double r = M.r[i] - alpha * M.t[i]; M.r[i] = r;
		const double d = M.diag[i]; const double z = (d != 0.0 ? 1.0 / d : 1.0) * r;
		if (owned_slot(S, i)) { r2 = r * r; rz = r * z; }
	}
	const double s1 = block_sum<PPE_T>(r2, sh); const double s2 = block_sum<PPE_T>(rz, sh);
	if (threadIdx.x == 0) { M.partial[blockIdx.x] = s1; M.partial[nBlocks + blockIdx.x] = s2; }
}
// p = M^-1 r + beta p
__global__ void __launch_bounds__(PPE_T) k_ppe_update_p(const __grid_constant__ DevView S, const __grid_constant__ PpeView M) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= S.n || M.scal[14] != 0.0) return;
	const double beta = M.scal[6]; const double d = M.diag[i];
	M.p[i] = (d != 0.0 ? 1.0 / d : 1.0) * M.r[i] + beta * M.p[i];
}
// second stage of the dot products: block partials -> sums[0..1] (device; on a slab rank the host then SUM-all-reduces them in place)
// mode 0: after residual0 (|b|^2, |r|^2) ; 1: after init_p (r.z) ; 2: after spmv (p.t) ; 3: after update_xr (|r|^2, r.z)
__global__ void __launch_bounds__(RED_T) k_ppe_reduce(const __grid_constant__ PpeView M, int nBlocks, int mode, double* sums) {
	__shared__ double sh[RED_T / 32];
	double a = 0.0, b = 0.0;
	for (int k = threadIdx.x; k < nBlocks; k += RED_T) { a += M.partial[k]; if (mode == 0 || mode == 3) b += M.partial[nBlocks + k]; }
	const double sa = block_sum<RED_T>(a, sh); const double sb = block_sum<RED_T>(b, sh);
	if (threadIdx.x == 0) { sums[0] = sa; sums[1] = sb; }
}
// the scalar recurrences of Eigen's CG (ConjugateGradient.h:50-89) from the (global) sums
__global__ void k_ppe_recur(const __grid_constant__ PpeView M, int mode, const double* sums, double* hostMirror) {
	const double sa = sums[0], sb = sums[1];
	double* s = M.scal;
	if (mode == 0) { s[0] = sa; s[1] = sb; double thr = DBL_EPSILON * DBL_EPSILON * sa; if (thr < DBL_MIN) thr = DBL_MIN; s[7] = thr; }
	else if (mode == 1) { s[2] = sa; }
	else if (mode == 2) { s[4] = sa; s[5] = s[2] / sa; }
	else { s[1] = sa; s[3] = s[2]; s[2] = sb; s[6] = sb / s[3]; }
	hostMirror[0] = s[0]; hostMirror[1] = s[1]; hostMirror[2] = s[7];
}
// ---- one GPU: second stage of the dot products + the scalar recurrence of Eigen's CG (ConjugateGradient.h:50-89) in ONE block of 1024
// threads, and the convergence test on the device: once the residual is below the threshold (scal[14] = 1) the remaining launches of
// a batch return at once, so the host reads the state back once per batch of iterations instead of synchronising in every one.
// 5 launches per iteration instead of 7, no per-iteration host round trip.  scal[15] counts the iterations.
#define RR_T 1024
__global__ void __launch_bounds__(RR_T) k_cg_reduce_recur(const __grid_constant__ PpeView M, int nBlocks, int mode) {
	__shared__ double sh[RR_T / 32];
	if (M.scal[14] != 0.0) return;
	double a = 0.0, b = 0.0;
	for (int k = threadIdx.x; k < nBlocks; k += RR_T) { a += M.partial[k]; if (mode == 3) b += M.partial[nBlocks + k]; }
	const double sa = block_sum<RR_T>(a, sh); const double sb = block_sum<RR_T>(b, sh);
	if (threadIdx.x != 0) return;
	double* s = M.scal;
	if (mode == 2) { s[4] = sa; s[5] = s[2] / sa; }   // alpha = absNew / p.Ap (:73)
	else {
		s[1] = sa; s[3] = s[2]; s[2] = sb; s[6] = sb / s[3];   // |r|^2 ; absOld ; absNew ; beta (:77-86)
		if (sa < s[7] || !(sa == sa)) s[14] = 1.0; else s[15] += 1.0;   // :78 "if(residualNorm2 < threshold) break" comes before "i++"
	}
}

// press = max(pressurePPE, 0)  (setZeroOnNegativePressure acts on press only, src/MpsPressure.cpp:471-476)
__global__ void k_ppe_finish(const __grid_constant__ DevView S, const __grid_constant__ KParams K, int zeroX) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= S.n) return;
	if (pk_type(S.pos[i]) == K.p.ghost) return;
	if (zeroX) S.pressurePPE[i] = 0.0;
	const double x = S.pressurePPE[i];
	S.press[i] = x < 0 ? 0.0 : x;
}

// ---- BiCGSTAB (Eigen BiCGSTAB.h:29-107 with the Jacobi preconditioner, BasicPreconditioners.h:67-91) ---------------------------
__device__ __forceinline__ double jacobi(const PpeView& M, int i) { const double d = M.diag[i]; return d != 0.0 ? 1.0 / d : 1.0; }
// r0 = r ; p = v = 0   (after residual0, and at a restart)
__global__ void __launch_bounds__(PPE_T) k_bicg_start(const __grid_constant__ DevView S, const __grid_constant__ PpeView M, int clearPV) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= S.n) return;
	M.r0[i] = M.r[i];
	if (clearPV) { M.p[i] = 0.0; M.v[i] = 0.0; }
}
// p = r + beta (p - w v) ; y = M^-1 p
__global__ void __launch_bounds__(PPE_T) k_bicg_p(const __grid_constant__ DevView S, const __grid_constant__ PpeView M) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= S.n) return;
	const double beta = M.scal[6], w = M.scal[11];
	const double p = M.r[i] + beta * (M.p[i] - w * M.v[i]);
	M.p[i] = p; M.y[i] = jacobi(M, i) * p;
}
// out = A in ; partial sums: which 0: r0.out   1: |out|^2 and out.s
__global__ void __launch_bounds__(PPE_T) k_bicg_spmv(const __grid_constant__ DevView S, const __grid_constant__ PpeView M, int nBlocks, int which) {
	__shared__ double sh[PPE_T / 32];
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	double a = 0.0, b = 0.0;
	if (i < S.n) {
		const double o = ell_row_dot(S, M, i, which == 0 ? M.y : M.z);
		if (which == 0) { M.v[i] = o; if (owned_slot(S, i)) a = M.r0[i] * o; }
		else { M.t[i] = o; if (owned_slot(S, i)) { a = o * o; b = o * M.s[i]; } }
	}
	const double sa = block_sum<PPE_T>(a, sh); const double sb = block_sum<PPE_T>(b, sh);
	if (threadIdx.x == 0) { M.partial[blockIdx.x] = sa; M.partial[nBlocks + blockIdx.x] = sb; }
}
// s = r - alpha v ; z = M^-1 s
__global__ void __launch_bounds__(PPE_T) k_bicg_s(const __grid_constant__ DevView S, const __grid_constant__ PpeView M) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= S.n) return;
	const double s = M.r[i] - M.scal[5] * M.v[i];
	M.s[i] = s; M.z[i] = jacobi(M, i) * s;
}
// x += alpha y + w z ; r = s - w t ; partial sums of |r|^2 and r0.r
__global__ void __launch_bounds__(PPE_T) k_bicg_xr(const __grid_constant__ DevView S, const __grid_constant__ PpeView M, int nBlocks) {
	__shared__ double sh[PPE_T / 32];
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	double a = 0.0, b = 0.0;
	if (i < S.n) {
		const double alpha = M.scal[5], w = M.scal[11];
		S.pressurePPE[i] += alpha * M.y[i] + w * M.z[i];
		const double r = M.s[i] - w * M.t[i]; M.r[i] = r;
		if (owned_slot(S, i)) { a = r * r; b = M.r0[i] * r; }
	}
	const double sa = block_sum<PPE_T>(a, sh); const double sb = block_sum<PPE_T>(b, sh);
	if (threadIdx.x == 0) { M.partial[blockIdx.x] = sa; M.partial[nBlocks + blockIdx.x] = sb; }
}
// r = b - A x (restart) ; partial sums of |r|^2
__global__ void __launch_bounds__(PPE_T) k_bicg_residual(const __grid_constant__ DevView S, const __grid_constant__ PpeView M, int nBlocks) {
	__shared__ double sh[PPE_T / 32];
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	double a = 0.0;
	if (i < S.n) { const double r = S.sourceTerm[i] - ell_row_dot(S, M, i, S.pressurePPE); M.r[i] = r; if (owned_slot(S, i)) a = r * r; }
	const double sa = block_sum<PPE_T>(a, sh); const double sb = block_sum<PPE_T>(0.0, sh);
	if (threadIdx.x == 0) { M.partial[blockIdx.x] = sa; M.partial[nBlocks + blockIdx.x] = sb; }
}
// scalar recurrences.  mode 10: after residual0 (|b|^2, |r|^2): rho = alpha = w = 1, |r0|^2 = |r|^2, then rho <- r0.r = |r|^2 for the
// first trip; 11: after spmv v (r0.v); 12: after spmv t (|t|^2, t.s); 13: after xr (|r|^2, r0.r); 14: restart (|r|^2)
__global__ void k_bicg_recur(const __grid_constant__ PpeView M, int mode, const double* sums, double* hostMirror) {
	const double sa = sums[0], sb = sums[1];
	double* s = M.scal;
	if (mode == 10) {
		s[0] = sa; s[1] = sb; s[12] = sb; s[7] = DBL_EPSILON * DBL_EPSILON * sa;
		s[3] = 1.0; s[5] = 1.0; s[11] = 1.0; s[2] = sb;               // rho_old = alpha = w = 1 ; rho = r0.r = |r|^2
		s[6] = (s[2] / s[3]) * (s[5] / s[11]);
	} else if (mode == 11) { s[5] = s[2] / sa; }
	else if (mode == 12) { s[11] = sa > 0.0 ? sb / sa : 0.0; }
	else if (mode == 13) { s[1] = sa; s[3] = s[2]; s[2] = sb; s[6] = (s[2] / s[3]) * (s[5] / s[11]); }
	else { s[1] = sa; s[12] = sa; s[2] = sa; s[6] = (s[2] / s[3]) * (s[5] / s[11]); }   // restart: r0 = r, rho = |r0|^2 = |r|^2 (rho_old stays)
	hostMirror[0] = s[0]; hostMirror[1] = s[1]; hostMirror[2] = s[7]; hostMirror[3] = s[2]; hostMirror[4] = s[12];
}

template <class T> static int palloc(pmps_handle* h, T** p, size_t count) {
	if (*p) return 0;
	void* q = nullptr;
	if (cudaMalloc(&q, sizeof(T) * (count ? count : 1)) != cudaSuccess) return 1;
	cudaMemsetAsync(q, 0, sizeof(T) * (count ? count : 1), h->stream);
	h->allocs.push_back(q); *p = (T*)q; return 0;
}

int pmps_ppe_solve(pmps_handle* h) {
	DevView& S = h->V; const KParams& K = h->K;
	if (K.p.solverType != 0 && K.p.solverType != 1) return pmps_fail(h, PMPS_E_ARG, "PPE: solver_type %d (0 = CG, 1 = BiCGSTAB)", K.p.solverType);
	const bool bicg = K.p.solverType == 1;
	{ const int rc = pmps_ensure_list(h); if (rc) return rc; }   // the assembly walks the Verlet candidate list
	const int n = S.n, nBlocks = cdiv(n, PPE_T);
	const size_t nWarps = ((size_t)S.n0 + 31) / 32, ell = nWarps * 32 * (size_t)S.listCap;
	int bad = 0;
	bad |= palloc(h, &h->ppe_col, ell); bad |= palloc(h, &h->ppe_val, ell); bad |= palloc(h, &h->ppe_cnt, nWarps * 32);
	bad |= palloc(h, &h->ppe_diag, S.n0); bad |= palloc(h, &h->ppe_r, S.n0); bad |= palloc(h, &h->ppe_p, S.n0); bad |= palloc(h, &h->ppe_t, S.n0);
	bad |= palloc(h, &h->ppe_partial, 2 * (size_t)cdiv(S.n0, PPE_T) + 2); bad |= palloc(h, &h->ppe_scal, 16);
	if (bicg) for (int k = 0; k < 5; k++) bad |= palloc(h, &h->ppe_bicg[k], S.n0);
	if (!h->ppe_host) { if (cudaMallocHost((void**)&h->ppe_host, 64) != cudaSuccess) bad |= 1; }
	if (bad) return pmps_fail(h, PMPS_E_CUDA, "PPE workspace allocation failed: %s", cudaGetErrorString(cudaGetLastError()));
	PpeView M; M.col = h->ppe_col; M.val = h->ppe_val; M.cnt = h->ppe_cnt; M.diag = h->ppe_diag; M.r = h->ppe_r; M.p = h->ppe_p; M.t = h->ppe_t;
	M.partial = h->ppe_partial; M.scal = h->ppe_scal;
	M.r0 = h->ppe_bicg[0]; M.v = h->ppe_bicg[1]; M.s = h->ppe_bicg[2]; M.y = h->ppe_bicg[3]; M.z = h->ppe_bicg[4];
	double* hm = h->ppe_host;
	S.ppeP = h->ppe_p;
	const bool slab = h->slab && h->slabSum;
	double* sums = slab ? h->slabWork : M.scal + 8;
	// one dot-product stage: partials -> sums, (slab: SUM over the ranks,) scalar recurrence
#define REDUCE(mode, nvals) do { LAUNCH(h, "ppe.reduce", k_ppe_reduce, 1, RED_T, M, nBlocks, (mode), sums); \
		if (slab) h->slabSum(h->slabUser, (nvals)); \
		LAUNCH(h, "ppe.recur", k_ppe_recur, 1, 1, M, (mode), sums, hm); } while (0)
	if (K.poly) { LAUNCH(h, "ppe.assemble", k_ppe_assemble<true>, nBlocks, PPE_T, S, K, M, bicg ? 1 : 0); }
	else { LAUNCH(h, "ppe.assemble", k_ppe_assemble<false>, nBlocks, PPE_T, S, K, M, bicg ? 1 : 0); }
	LAUNCH(h, "ppe.residual0", k_ppe_residual0, nBlocks, PPE_T, S, M, nBlocks);
	if (!bicg) REDUCE(0, 2);
	else { LAUNCH(h, "ppe.reduce", k_ppe_reduce, 1, RED_T, M, nBlocks, 0, sums); if (slab) h->slabSum(h->slabUser, 2); LAUNCH(h, "ppe.recur", k_bicg_recur, 1, 1, M, 10, sums, hm); }
	CUDA_TRY(h, cudaStreamSynchronize(h->stream));
	const double rhs2 = hm[0]; double res2 = hm[1]; const double thr = hm[2];
	int iters = 0; double err = 0.0; int zeroX = 0;
	if (rhs2 == 0.0) { zeroX = 1; if (bicg) { iters = 2 * h->numParticles; err = DBL_EPSILON; } }   // Eigen's BiCGSTAB returns before it touches iters / error
	else if (bicg) {
		// Eigen BiCGSTAB.h:29-107.  Per trip: p/y update, v = A y (+ r0.v), s/z update, t = A z (+ |t|^2, t.s), x/r update (+ |r|^2, r0.r);
		// one host read of {|r|^2, rho, |r0|^2} per trip decides about convergence and the restart
#define BREDUCE(mode, nvals) do { LAUNCH(h, "ppe.reduce", k_ppe_reduce, 1, RED_T, M, nBlocks, 0, sums); \
		if (slab) h->slabSum(h->slabUser, (nvals)); \
		LAUNCH(h, "ppe.recur", k_bicg_recur, 1, 1, M, (mode), sums, hm); } while (0)
		LAUNCH(h, "ppe.bicg_start", k_bicg_start, nBlocks, PPE_T, S, M, 1);
		const long long maxIt = 2ll * h->numParticles; int restarts = 0;
		const double eps2 = DBL_EPSILON * DBL_EPSILON;
		double rho = hm[3], r0n = hm[4];
		while (res2 > thr && iters < maxIt) {
			if (fabs(rho) < eps2 * r0n) {   // the residual became orthogonal to r0: restart from the true residual (:67-75)
				LAUNCH(h, "ppe.bicg_residual", k_bicg_residual, nBlocks, PPE_T, S, M, nBlocks);
				BREDUCE(14, 1);
				LAUNCH(h, "ppe.bicg_start", k_bicg_start, nBlocks, PPE_T, S, M, 0);
				if (restarts++ == 0) iters = 0;
			}
			LAUNCH(h, "ppe.bicg_p", k_bicg_p, nBlocks, PPE_T, S, M);
			if (slab) { S.ppeP = M.y; h->slabRefresh(h->slabUser, PMPS_HF_PPE_P); }
			LAUNCH(h, "ppe.spmv", k_bicg_spmv, nBlocks, PPE_T, S, M, nBlocks, 0);
			BREDUCE(11, 1);
			LAUNCH(h, "ppe.bicg_s", k_bicg_s, nBlocks, PPE_T, S, M);
			if (slab) { S.ppeP = M.z; h->slabRefresh(h->slabUser, PMPS_HF_PPE_P); }
			LAUNCH(h, "ppe.spmv", k_bicg_spmv, nBlocks, PPE_T, S, M, nBlocks, 1);
			BREDUCE(12, 2);
			LAUNCH(h, "ppe.bicg_xr", k_bicg_xr, nBlocks, PPE_T, S, M, nBlocks);
			BREDUCE(13, 2);
			CUDA_TRY(h, cudaStreamSynchronize(h->stream));
			res2 = hm[1]; rho = hm[3]; r0n = hm[4];
			iters++;
			if (!(res2 == res2)) break;
		}
#undef BREDUCE
		err = sqrt(res2 / rhs2);
	}
	else {
		if (res2 >= thr) {
			LAUNCH(h, "ppe.init_p", k_ppe_init_p, nBlocks, PPE_T, S, M, nBlocks);
			REDUCE(1, 1);
			const long long maxIt = 2ll * h->numParticles;
			if (!slab && !getenv("PMPS_CG_HOST_LOOP")) {
				// one GPU: reductions fused into the producing kernels, convergence tested on the device, one read-back per batch
				CUDA_TRY(h, cudaMemsetAsync(M.scal + 14, 0, 2 * sizeof(double), h->stream));   // done flag, iteration count
				const int batch = 6;
				while (true) {
					for (int k = 0; k < batch; k++) {
						LAUNCH(h, "ppe.spmv", k_ppe_spmv, nBlocks, PPE_T, S, M, nBlocks);
						LAUNCH(h, "ppe.reduce", k_cg_reduce_recur, 1, RR_T, M, nBlocks, 2);
						LAUNCH(h, "ppe.update_xr", k_ppe_update_xr, nBlocks, PPE_T, S, M, nBlocks);
						LAUNCH(h, "ppe.reduce", k_cg_reduce_recur, 1, RR_T, M, nBlocks, 3);
						LAUNCH(h, "ppe.update_p", k_ppe_update_p, nBlocks, PPE_T, S, M);
					}
					CUDA_TRY(h, cudaMemcpyAsync(hm, M.scal + 1, sizeof(double), cudaMemcpyDeviceToHost, h->stream));
					CUDA_TRY(h, cudaMemcpyAsync(hm + 1, M.scal + 14, 2 * sizeof(double), cudaMemcpyDeviceToHost, h->stream));
					CUDA_TRY(h, cudaStreamSynchronize(h->stream));
					res2 = hm[0]; iters = (int)hm[2];
					if (hm[1] != 0.0 || iters >= maxIt) break;
				}
			}
			else while (iters < maxIt) {
				if (slab) h->slabRefresh(h->slabUser, PMPS_HF_PPE_P);   // the halo copies' p before the matrix-vector product
				LAUNCH(h, "ppe.spmv", k_ppe_spmv, nBlocks, PPE_T, S, M, nBlocks);
				REDUCE(2, 1);
				LAUNCH(h, "ppe.update_xr", k_ppe_update_xr, nBlocks, PPE_T, S, M, nBlocks);
				REDUCE(3, 2);
				CUDA_TRY(h, cudaStreamSynchronize(h->stream));
				res2 = hm[1];
				if (res2 < thr || !(res2 == res2)) break;
				LAUNCH(h, "ppe.update_p", k_ppe_update_p, nBlocks, PPE_T, S, M);
				iters++;
			}
		}
		err = sqrt(res2 / rhs2);
	}
#undef REDUCE
	LAUNCH(h, "ppe.finish", k_ppe_finish, cdiv(n, 256), 256, S, K, zeroX);
	CUDA_TRY(h, cudaGetLastError());
	h->solverIter = iters; h->solverError = err;
	if (!(res2 == res2)) return pmps_fail(h, PMPS_E_STATE, "PPE: the conjugate-gradient residual became NaN after %d iterations", iters);
	return PMPS_OK;
}
