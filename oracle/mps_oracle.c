/* oracle/mps_oracle.c — CPU restatement of the PolyMPS per-timestep particle hot path.  TEST INFRASTRUCTURE ONLY
 * (see mps_oracle.h for the rules and the parity-pinning statement).
 *
 * This is NOT the reference's code: it is one generic neighbour walk (NEIGH_LOOP) plus one small body per stage,
 * written from the formulas, with each function citing the reference file:line it follows.  It keeps the
 * reference's evaluation ORDER (bucket order jz,jy,jx; list order = ascending particle index; left-to-right sums;
 * no FMA contraction: built with -ffp-contract=off) so that it reproduces the reference bit for bit, which is what
 * lets tests/ pin it against the reference dumps with exact equality instead of a tolerance.
 */
#include "mps_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>
#ifdef _OPENMP
#include <omp.h>
#endif

static int g_threads = 1;
void oracle_set_threads(int n) { g_threads = n < 1 ? 1 : n; }
#define PARFOR _Pragma("omp parallel for schedule(dynamic,64) num_threads(g_threads)")

/* ------------------------------------------------------------------------------------------------ helpers */
static inline double wgt(double r, double re, int type) {           /* src/MpsParticle.cpp:57-91 */
	switch (type) {
		case 1: return re / r + r / re - 2.0;
		case 2: return re / r - r / re;
		case 3: return (1.0 - r / re) * (1.0 - r / re) * (1.0 - r / re);
		case 4: return (1.0 - r / re) * (1.0 - r / re);
		default: return re / r - 1.0;
	}
}

static inline void bucket_of(const PmpsParams* P, const double* x, int* bx, int* by, int* bz) {   /* src/MpsBucket.cpp:420-425 */
	*bx = (int)((x[0] - P->domainMinX) * P->invBucketSide + P->epsilonZero);
	*by = (int)((x[1] - P->domainMinY) * P->invBucketSide + P->epsilonZero);
	*bz = (int)((x[2] - P->domainMinZ) * P->invBucketSide + P->epsilonZero);
}

/* The one neighbour walk every stage shares (canonical instance src/MpsViscosity.cpp:33-78).
 * Inside BODY: j, rx,ry,rz,d2 (r_ij = x_j - x_i), mx,my,mz,dm2 (x_j - mirror_i).  j == i is NOT filtered here. */
/* Periodic domain (domainTypeBC 1): the first / last bucket layer of the periodic axis carries the list of the last / first inner
 * layer (oracle_update_buckets) and its members are seen moved by -/+ periodicLength: r = (x_j + pl) - x_i, two roundings
 * (getPeriodicLengths + sqrDistBetweenParticles, src/MpsParticle.cpp:35-54). */
static inline double periodic_shift(const PmpsParams* P, int jb) {
	/* bucketPeriodicBC is looked up by the LINEAR bucket index (setBucketBC, src/MpsBoundaryCondition.cpp:20-60): a particle that has
	 * just crossed a periodic face sits in a border layer until the next out-of-domain check, its loop coordinate -1 / n along the axis
	 * then runs into the neighbouring row of the bucket array, and the shift is that of the bucket actually read */
	const int nX = P->numBucketsX, nXY = nX * P->numBucketsY;
	const int jz = jb / nXY, jy = (jb - jz * nXY) / nX, jx = jb - jz * nXY - jy * nX;
	const int c = P->periodicAxis == 0 ? jx : P->periodicAxis == 1 ? jy : jz;
	const int n = P->periodicAxis == 0 ? P->numBucketsX : P->periodicAxis == 1 ? P->numBucketsY : P->numBucketsZ;
	return c == 0 ? -P->periodicLength : c == n - 1 ? P->periodicLength : 0.0;
}
#define NEIGH_LOOP(P, A, B, xi, mi, ...)                                                                       \
	do {                                                                                                       \
		int bx_, by_, bz_; bucket_of(P, xi, &bx_, &by_, &bz_);                                                 \
		const int nX_ = (P)->numBucketsX, nXY_ = (P)->numBucketsX * (P)->numBucketsY;                          \
		const int z0_ = (P)->dim == 3 ? bz_ - 1 : 0, z1_ = (P)->dim == 3 ? bz_ + 1 : 0;                        \
		for (int jz_ = z0_; jz_ <= z1_; jz_++) for (int jy_ = by_ - 1; jy_ <= by_ + 1; jy_++)                  \
		for (int jx_ = bx_ - 1; jx_ <= bx_ + 1; jx_++) {                                                       \
			const int jb_ = jz_ * nXY_ + jy_ * nX_ + jx_;                                                       \
			const double pl_ = (P)->domainTypeBC == 1 ? periodic_shift(P, jb_) : 0.0;                           \
			for (int j = (B)->first[jb_]; j != -1; j = (B)->next[j]) {                                          \
				const double* xr_ = (A)->pos + 3 * (size_t)j; double xs_[3];                                   \
				const double* xj_ = xr_;                                                                       \
				if ((P)->domainTypeBC == 1) { xs_[0] = xr_[0]; xs_[1] = xr_[1]; xs_[2] = xr_[2]; xs_[(P)->periodicAxis] = xr_[(P)->periodicAxis] + pl_; xj_ = xs_; } \
				const double rx = xj_[0] - (xi)[0], ry = xj_[1] - (xi)[1], rz = xj_[2] - (xi)[2];              \
				const double d2 = rx * rx + ry * ry + rz * rz;                                                 \
				const double mx = xj_[0] - (mi)[0], my = xj_[1] - (mi)[1], mz = xj_[2] - (mi)[2];              \
				const double dm2 = mx * mx + my * my + mz * mz;                                                \
				(void)dm2; (void)mx; (void)my; (void)mz;                                                       \
				__VA_ARGS__                                                                                    \
			}                                                                                                  \
		}                                                                                                      \
	} while (0)

/* predicate P(re): "real neighbour"  */
#define PRED_P(P, re2) (d2 < (re2) && (d2 < dm2 || (P)->wallType == 0))
/* predicate M(re): mirror neighbourhood of a near-wall fluid particle */
#define PRED_M(re2) (d2 < (re2) && dm2 < (re2) && d2 < dm2)

/* unit normal i->wall = half the vector mirror->i, and the reflection R = I - 2 n n^T (e.g. src/MpsPndNeigh.cpp:114-136) */
static inline double wall_normal(const double* xi, const double* mi, double eps, int squared_test, double n[3]) {
	n[0] = 0.5 * (xi[0] - mi[0]); n[1] = 0.5 * (xi[1] - mi[1]); n[2] = 0.5 * (xi[2] - mi[2]);
	const double n2 = n[0] * n[0] + n[1] * n[1] + n[2] * n[2];
	const double len = sqrt(n2);
	/* two variants exist in the reference: test on |n| (calcWallNPCD, wall gradient, wall shifting) or on |n|^2
	 * (wall viscosity, wall diffusive term) */
	if ((squared_test ? n2 : len) > eps) { n[0] = n[0] / len; n[1] = n[1] / len; n[2] = n[2] / len; }
	else { n[0] = n[1] = n[2] = 0.0; }
	return len;
}
static inline void reflection(const double n[3], double R[9]) {
	R[0] = 1.0 - 2.0 * n[0] * n[0]; R[1] = 0.0 - 2.0 * n[0] * n[1]; R[2] = 0.0 - 2.0 * n[0] * n[2];
	R[3] = 0.0 - 2.0 * n[1] * n[0]; R[4] = 1.0 - 2.0 * n[1] * n[1]; R[5] = 0.0 - 2.0 * n[1] * n[2];
	R[6] = 0.0 - 2.0 * n[2] * n[0]; R[7] = 0.0 - 2.0 * n[2] * n[1]; R[8] = 1.0 - 2.0 * n[2] * n[2];
}
#define MATVEC(R, a, b, c, k) ((R)[3 * (k)] * (a) + (R)[3 * (k) + 1] * (b) + (R)[3 * (k) + 2] * (c))

/* ------------------------------------------------------------------------------------------------ buckets */
void oracle_update_buckets(const PmpsParams* P, PmpsHostArrays* A, OBuckets* B) {
	const int nB = P->numBucketsX * P->numBucketsY * P->numBucketsZ, nXY = P->numBucketsX * P->numBucketsY;
	for (int b = 0; b < nB; b++) { B->first[b] = -1; B->last[b] = -1; }
	for (int i = 0; i < A->numParticlesZero; i++) B->next[i] = -1;
	for (int i = 0; i < A->numParticles; i++) {
		if (A->particleType[i] == P->ghost) continue;
		const double* x = A->pos + 3 * (size_t)i; int ib;
		if (P->dim == 2) {   /* 2D insertion adds epsilonZero, 3D insertion does not (src/MpsBucket.cpp:312-313 vs :333-335) */
			int ix = (int)((x[0] - P->domainMinX) * P->invBucketSide + P->epsilonZero);
			int iy = (int)((x[1] - P->domainMinY) * P->invBucketSide + P->epsilonZero);
			ib = iy * P->numBucketsX + ix;
		} else {
			int ix = (int)((x[0] - P->domainMinX) * P->invBucketSide);
			int iy = (int)((x[1] - P->domainMinY) * P->invBucketSide);
			int iz = (int)((x[2] - P->domainMinZ) * P->invBucketSide);
			ib = iz * nXY + iy * P->numBucketsX + ix;
		}
		const int tail = B->last[ib];
		B->last[ib] = i;
		if (tail == -1) B->first[ib] = i; else B->next[tail] = i;
	}
	if (P->domainTypeBC == 1) {   /* copyDataBetweenBuckets, src/MpsBucket.cpp:356-417: the border layers alias the opposite inner layers */
		const int nX = P->numBucketsX, nY = P->numBucketsY, nZ = P->numBucketsZ, a = P->periodicAxis;
		const int nA = a == 0 ? nX : a == 1 ? nY : nZ, stride = a == 0 ? 1 : a == 1 ? nX : nXY;
		for (int iz = 0; iz < nZ; iz++) for (int iy = 0; iy < nY; iy++) for (int ix = 0; ix < nX; ix++) {
			const int c = a == 0 ? ix : a == 1 ? iy : iz, ib = iz * nXY + iy * nX + ix;
			if (c == 0) { B->first[ib] = B->first[ib + stride * (nA - 2)]; B->last[ib] = B->last[ib + stride * (nA - 2)]; }
			if (c == nA - 1) { B->first[ib] = B->first[ib - stride * (nA - 2)]; B->last[ib] = B->last[ib - stride * (nA - 2)]; }
		}
	}
}

int oracle_neighbors(const PmpsParams* P, const PmpsHostArrays* A, const OBuckets* B, int i, int kind, int32_t* out, int cap) {
	const double* xi = A->pos + 3 * (size_t)i; const double* mi = A->mirrorParticlePos + 3 * (size_t)i; int n = 0;
	NEIGH_LOOP(P, A, B, xi, mi, {
		int hit = 0;
		if (kind == 0) hit = PRED_P(P, P->reL2); else if (kind == 1) hit = PRED_P(P, P->reS2);
		else if (kind == 2) hit = PRED_M(P->reS2); else if (kind == 3) hit = PRED_M(P->reL2);
		else hit = (d2 < P->distCollisionLimit2 && (d2 < dm2 || P->wallType == 0));
		if (hit && j != i) { if (n < cap) out[n] = j; n++; }
	});
	return n;
}

/* ------------------------------------------------------------------------------------------------ elementwise */
void oracle_add_gravity(const PmpsParams* P, PmpsHostArrays* A) {
	for (int i = 0; i < A->numParticles; i++) {
		double* a = A->acc + 3 * (size_t)i; double* s = A->accStar + 3 * (size_t)i;
		a[0] += P->gravityX; a[1] += P->gravityY; a[2] += P->gravityZ;
		s[0] = a[0]; s[1] = a[1]; s[2] = a[2];
	}
}

void oracle_predict(const PmpsParams* P, PmpsHostArrays* A) {
	const double dt = P->timeStep, big = 10e8 * P->partDist;
	for (int i = 0; i < A->numParticles; i++) {
		double *x = A->pos + 3 * (size_t)i, *v = A->vel + 3 * (size_t)i, *a = A->acc + 3 * (size_t)i;
		if (A->particleType[i] == P->fluid) for (int k = 0; k < 3; k++) { v[k] += a[k] * dt; x[k] += v[k] * dt; }
		for (int k = 0; k < 3; k++) {
			a[k] = 0.0; A->npcdDeviation[3 * (size_t)i + k] = 0.0;
			if (A->dvelCollision) A->dvelCollision[3 * (size_t)i + k] = 0.0;
			if (P->wallType == 1) A->mirrorParticlePos[3 * (size_t)i + k] = big;
		}
		A->numNeighWallContribution[i] = 0; A->particleNearWall[i] = 0; A->distParticleWall2[i] = big;
	}
}

double oracle_correct(const PmpsParams* P, PmpsHostArrays* A) {
	const double dt = P->timeStep; double vmax = 0.0;
	for (int i = 0; i < A->numParticles; i++) {
		if (A->particleType[i] != P->fluid) continue;
		double *x = A->pos + 3 * (size_t)i, *v = A->vel + 3 * (size_t)i, *a = A->acc + 3 * (size_t)i;
		for (int k = 0; k < 3; k++) { v[k] += a[k] * dt; x[k] += a[k] * dt * dt; a[k] = 0.0; }
		const double v2 = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
		if (v2 > vmax * vmax) vmax = sqrt(v2);
	}
	return vmax;
}

/* Ghosting of fluid particles that left the domain: serial and order dependent (swap with the last live particle).
 * Periodic wrap is out of scope (domainTypeBC == 0 only).  Returns the number of particles removed. */
int oracle_check_out_domain(const PmpsParams* P, PmpsHostArrays* A) {
	const double lo[3] = { P->domainMinX + P->bucketSide, P->domainMinY + P->bucketSide, P->domainMinZ + P->bucketSide };
	const double hi[3] = { P->domainMaxX - P->bucketSide, P->domainMaxY - P->bucketSide, P->domainMaxZ - P->bucketSide };
	int live = A->numParticles, removed = 0;
	if (P->domainTypeBC == 1) {   /* periodic: every particle beyond a face of the physical domain re-enters through the opposite one (src/MpsParticle.cpp:236-269) */
		for (int i = 0; i < A->numParticles; i++) {
			double* x = A->pos + 3 * (size_t)i + P->periodicAxis;
			if (*x > P->physDomMax) *x -= P->periodicLength;
			if (*x < P->physDomMin) *x += P->periodicLength;
		}
	}
#define MV1(arr) do { if (A->arr) A->arr[i] = A->arr[last]; } while (0)
#define MV3(arr) do { if (A->arr) for (int k = 0; k < 3; k++) A->arr[3 * (size_t)i + k] = A->arr[3 * (size_t)last + k]; } while (0)
	for (int i = 0; i < A->numParticles; i++) {   /* the loop bound shrinks as particles are removed (:286,:392) */
		const double* x = A->pos + 3 * (size_t)i;
		if (A->particleType[i] == P->fluid && i < live &&
			(x[0] > hi[0] || x[0] < lo[0] || x[1] > hi[1] || x[1] < lo[1] || (P->dim == 3 && (x[2] > hi[2] || x[2] < lo[2])))) {
			const int last = live - 1;
			MV1(particleType); MV1(particleBC); MV1(numNeigh); MV1(press); MV1(pndi); MV1(pndki); MV1(pndski); MV1(pndSmall);
			MV1(npcdDeviation2); MV1(concentration); MV1(velDivergence); MV1(diffusiveTerm); MV1(nearMeshType);
			MV1(particleNearWall); MV1(numNeighWallContribution); MV1(pndWallContribution); MV1(deviationDotPolygonNormal);
			MV1(distParticleWall2); MV1(PTYPE); MV1(RHO); MV1(MEU); MV1(elementID);
			if (P->mpsType >= 2) { MV1(pressurePPE); MV1(sourceTerm); }
			MV3(acc); MV3(accStar); MV3(pos); MV3(vel); MV3(npcdDeviation); MV3(gradConcentration); MV3(correcMatrixRow1);
			MV3(correcMatrixRow2); MV3(correcMatrixRow3); MV3(dvelCollision); MV3(particleAtWallPos); MV3(mirrorParticlePos);
			MV3(polygonNormal); MV3(forceWall);
			if (P->fluidType == 1) {   /* src/MpsParticle.cpp:322-337 */
				MV1(Cv); MV1(II); MV1(MEU_Y); MV1(Inertia); MV1(p_rheo_new); MV1(p_smooth); MV1(VF);
				MV1(S12); MV1(S13); MV1(S23); MV1(S11); MV1(S22); MV1(S33);
			}
			A->particleType[last] = P->ghost; A->particleBC[last] = P->other; A->particleNearWall[last] = 0;
			A->nearMeshType[last] = 0; A->distParticleWall2[last] = 10e8 * P->partDist;
			for (int k = 0; k < 3; k++) A->vel[3 * (size_t)last + k] = 0.0;
			A->press[last] = 0.0;
			A->pos[3 * (size_t)last] = P->domainMaxX - P->partDist; A->pos[3 * (size_t)last + 1] = P->domainMaxY - P->partDist;
			A->pos[3 * (size_t)last + 2] = P->dim == 2 ? 0.0 : P->domainMaxZ - P->partDist;
			live--; removed++;
		}
		A->numParticles = live;
	}
#undef MV1
#undef MV3
	return removed;
}

void oracle_update_pnd(const PmpsParams* P, PmpsHostArrays* A) {
	(void)P;
	for (int i = 0; i < A->numParticles; i++) { A->pndi[i] = A->pndAuxVar1[i]; A->pndAuxVar1[i] = 0.0; }
}

/* MpsShifting::calcNormalParticles, src/MpsShifting.cpp:106-174: normal_i = -(1/n0S) sum_j r_ij w_S / |r_ij| over P(reS) */
void oracle_calc_normals(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B) {
	PARFOR
	for (int i = 0; i < A->numParticles; i++) {
		const double* xi = A->pos + 3 * (size_t)i; const double* mi = A->mirrorParticlePos + 3 * (size_t)i;
		double dx = 0.0, dy = 0.0, dz = 0.0;
		NEIGH_LOOP(P, A, B, xi, mi, {
			if (PRED_P(P, P->reS2) && j != i) {
				const double dst = sqrt(d2); const double w = wgt(dst, P->reS, P->weightType);
				dx += rx * w / dst; dy += ry * w / dst; dz += rz * w / dst;
			}
		});
		A->normal[3 * (size_t)i] = -dx / P->pndSmallZero; A->normal[3 * (size_t)i + 1] = -dy / P->pndSmallZero; A->normal[3 * (size_t)i + 2] = -dz / P->pndSmallZero;
	}
}
/* MpsShifting::calcWallNormalParticles, src/MpsShifting.cpp:319-429: the mirror neighbourhood's part, reflected back */
void oracle_calc_wall_normals(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B) {
	PARFOR
	for (int i = 0; i < A->numParticles; i++) {
		if (!(A->particleType[i] == P->fluid && A->particleNearWall[i])) continue;
		const double* xi = A->pos + 3 * (size_t)i; const double* mi = A->mirrorParticlePos + 3 * (size_t)i;
		double n[3], R[9]; wall_normal(xi, mi, P->epsilonZero, 0, n); reflection(n, R);
		double dx = 0.0, dy = 0.0, dz = 0.0;
		NEIGH_LOOP(P, A, B, xi, mi, {
			if (PRED_M(P->reS2) && j != i) {
				const double dst = sqrt(dm2); const double w = wgt(dst, P->reS, P->weightType);
				dx += mx * w / dst; dy += my * w / dst; dz += mz * w / dst;
			}
		});
		const double ex = xi[0] - mi[0], ey = xi[1] - mi[1], ez = xi[2] - mi[2]; const double dd = ex * ex + ey * ey + ez * ez;
		if (dd < P->reS2) { const double dst = sqrt(dd); const double w = wgt(dst, P->reS, P->weightType); dx += ex * w / dst; dy += ey * w / dst; dz += ez * w / dst; }
		const double rx_ = MATVEC(R, dx, dy, dz, 0), ry_ = MATVEC(R, dx, dy, dz, 1), rz_ = MATVEC(R, dx, dy, dz, 2);
		A->normal[3 * (size_t)i] += -rx_ / P->pndSmallZero; A->normal[3 * (size_t)i + 1] += -ry_ / P->pndSmallZero; A->normal[3 * (size_t)i + 2] += -rz_ / P->pndSmallZero;
	}
}

/* MpsBoundaryCondition::updateParticleBC, src/MpsBoundaryCondition.cpp:66-399 (PND_NEIGH, PND_NPCD, PND_ARC) */
void oracle_update_bc(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B) {
	PARFOR
	for (int i = 0; i < A->numParticles; i++) {
		if (A->particleType[i] == P->dummyWall) { A->particleBC[i] = P->other; continue; }
		int bc = A->pndSmall[i] < P->betaPnd ? P->surface : P->inner;
		if (P->freeSurfType == 0) bc = (A->pndSmall[i] < P->betaPnd && A->numNeigh[i] < P->betaNeigh) ? P->surface : P->inner;
		else if (P->freeSurfType == 1) { if (bc == P->surface && A->numNeigh[i] > 4 && A->npcdDeviation2[i] < P->delta2) bc = P->inner; }
		else if (P->freeSurfType == 2) {   /* :205-390 */
			double nx = A->normal[3 * (size_t)i], ny = A->normal[3 * (size_t)i + 1], nz = A->normal[3 * (size_t)i + 2];
			const double norm2 = nx * nx + ny * ny + nz * nz;
			if (A->pndi[i] >= P->betaPnd && A->numNeigh[i] >= P->betaNeigh && norm2 <= P->normThreshold2) bc = P->inner;
			else {
				const double* xi = A->pos + 3 * (size_t)i; const double* mi = A->mirrorParticlePos + 3 * (size_t)i;
				if (norm2 > P->epsilonZero) { const double nm = sqrt(norm2); nx /= nm; ny /= nm; nz /= nm; } else { nx = ny = nz = 0.0; }
				double nw[3]; nw[0] = 0.5 * (xi[0] - mi[0]); nw[1] = 0.5 * (xi[1] - mi[1]); nw[2] = 0.5 * (xi[2] - mi[2]);
				const double nw2 = nw[0] * nw[0] + nw[1] * nw[1] + nw[2] * nw[2];
				if (nw2 > P->epsilonZero) { const double m_ = sqrt(nw2); nw[0] /= m_; nw[1] /= m_; nw[2] /= m_; } else { nw[0] = nw[1] = nw[2] = 0.0; }
				double R[9]; reflection(nw, R);
				double qx = MATVEC(R, nx, ny, nz, 0), qy = MATVEC(R, nx, ny, nz, 1), qz = MATVEC(R, nx, ny, nz, 2);   /* normal of the mirror particle */
				const double q2 = qx * qx + qy * qy + qz * qz;
				if (q2 > P->epsilonZero) { const double qm = sqrt(q2); qx /= qm; qy /= qm; qz /= qm; } else { qx = qy = qz = 0.0; }
				const double lo = P->partDist;
				const double tx = xi[0] + lo * nx, ty = xi[1] + lo * ny, tz = xi[2] + lo * nz;          /* point one particle distance along the normal */
				const double ux = mi[0] + lo * qx, uy = mi[1] + lo * qy, uz = mi[2] + lo * qz;          /* the same from the mirror particle */
				int found = 0;
				NEIGH_LOOP(P, A, B, xi, mi, {
					if (!found) {   /* the reference leaves both loops at the first hit (goto endloop) */
						const double* xj = A->pos + 3 * (size_t)j;
						if (PRED_P(P, P->reS2) && j != i) {
							const double ax = xj[0] - tx, ay = xj[1] - ty, az = xj[2] - tz; const double dn2 = ax * ax + ay * ay + az * az;
							const double rn = rx * nx + ry * ny + rz * nz; const double ang = acos(rn / sqrt(d2));
							if ((d2 >= P->dstThreshold2 && dn2 < P->hThreshold2) || (d2 < P->dstThreshold2 && ang < P->thetaArc)) { bc = P->inner; found = 1; }
							else bc = P->surface;
						}
						if (!found && P->wallType == 1 && PRED_M(P->reS2) && j != i) {
							const double ax = xj[0] - ux, ay = xj[1] - uy, az = xj[2] - uz; const double dn2 = ax * ax + ay * ay + az * az;
							const double rn = mx * qx + my * qy + mz * qz; const double ang = acos(rn / sqrt(dm2));
							if ((dm2 >= P->dstThreshold2 && dn2 < P->hThreshold2) || (dm2 < P->dstThreshold2 && ang < P->thetaArc)) { bc = P->inner; found = 1; }
							else bc = P->surface;
						}
					}
				});
			}
		}
		A->particleBC[i] = bc;
	}
}

void oracle_calc_press_eos(const PmpsParams* P, PmpsHostArrays* A) {
	for (int i = 0; i < A->numParticles; i++) {
		const double mi = A->PTYPE[i] == 1 ? P->DNS_FL1 : P->DNS_FL2; double p = 0.0;
		if (A->particleBC[i] == P->inner) {
			if (P->mpsType == 0) p = (A->pndi[i] - P->pndSmallZero) * P->coeffPressEMPS * mi;
			else p = (mi * P->coeffPressWCMPS / P->gamma) * (pow(A->pndi[i] / P->pndSmallZero, P->gamma) - 1);
		}
		if (p < 0.0) p = 0.0;
		A->press[i] = p;
	}
}

/* ------------------------------------------------------------------------------------------------ viscosity */
void oracle_calc_viscosity(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B) {
	PARFOR
	for (int i = 0; i < A->numParticles; i++) {
		const double* xi = A->pos + 3 * (size_t)i; const double* mi = A->mirrorParticlePos + 3 * (size_t)i;
		const double* vi = A->vel + 3 * (size_t)i; const double mu_i = A->MEU[i];
		double sx = 0.0, sy = 0.0, sz = 0.0;
		NEIGH_LOOP(P, A, B, xi, mi, {
			if (PRED_P(P, P->reL2) && j != i) {
				const double w = wgt(sqrt(d2), P->reL, P->weightType);
				double nu = (mu_i + A->MEU[j]) > P->epsilonZero ? 2.0 * mu_i * A->MEU[j] / (mu_i + A->MEU[j]) : 0.0;
				if (A->particleType[j] == P->wall) nu = 2.0 * mu_i;
				nu = nu / A->RHO[i];
				const double* vj = A->vel + 3 * (size_t)j;
				sx += (vj[0] - vi[0]) * w * nu; sy += (vj[1] - vi[1]) * w * nu; sz += (vj[2] - vi[2]) * w * nu;
			}
		});
		A->acc[3 * (size_t)i] = P->coeffViscMultiphase * sx; A->acc[3 * (size_t)i + 1] = P->coeffViscMultiphase * sy;
		A->acc[3 * (size_t)i + 2] = P->coeffViscMultiphase * sz;
	}
}

/* mirror velocity of a near-wall particle: no-slip -(v - 2(v_w - (n.v_w) n)), free-slip R v
 * (src/MpsViscosity.cpp:1729-1752 no-slip; :1565-1590 free-slip) */
static inline void mirror_velocity(const PmpsParams* P, const PmpsHostArrays* A, int i, const double n[3], const double R[9], double vm[3]) {
	const double* v = A->vel + 3 * (size_t)i; double vw[3] = { 0.0, 0.0, 0.0 };
	if (A->nearMeshType[i] == 2) { vw[0] = P->uniformVelWall[0]; vw[1] = P->uniformVelWall[1]; vw[2] = P->uniformVelWall[2]; }
	const double dn = n[0] * vw[0] + n[1] * vw[1] + n[2] * vw[2];
	if (P->slipCondition == 0) {
		double t[3]; for (int k = 0; k < 3; k++) t[k] = v[k] - 2.0 * (vw[k] - dn * n[k]);
		vm[0] = (-1.0 * t[0] + 0.0 * t[1] + 0.0 * t[2]); vm[1] = (0.0 * t[0] + -1.0 * t[1] + 0.0 * t[2]); vm[2] = (0.0 * t[0] + 0.0 * t[1] + -1.0 * t[2]);
	} else {   /* free-slip: plain reflection, wall velocity ignored */
		for (int k = 0; k < 3; k++) vm[k] = MATVEC(R, v[0], v[1], v[2], k);
	}
}

void oracle_wall_viscosity(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B) {
	const double vol = pow(P->partDist, (double)P->dim);
	PARFOR
	for (int i = 0; i < A->numParticles; i++) {
		if (!(A->particleType[i] == P->fluid && A->particleNearWall[i])) continue;
		const double* xi = A->pos + 3 * (size_t)i; const double* mi = A->mirrorParticlePos + 3 * (size_t)i;
		const double* vi = A->vel + 3 * (size_t)i; const double mu_i = A->MEU[i];
		double n[3], R[9], T[9], vm[3];
		wall_normal(xi, mi, P->epsilonZero, 1, n); reflection(n, R); mirror_velocity(P, A, i, n, R, vm);
		if (P->slipCondition == 0) { for (int k = 0; k < 9; k++) T[k] = 0.0; T[0] = T[4] = T[8] = -1.0; }   /* R_inv = -I */
		else memcpy(T, R, sizeof T);
		const double rho_div = A->PTYPE[i] == 1 ? P->DNS_FL1 : P->DNS_FL2;
		double sx = 0.0, sy = 0.0, sz = 0.0;
		NEIGH_LOOP(P, A, B, xi, mi, {
			if (PRED_M(P->reL2) && j != i) {
				const double w = wgt(sqrt(dm2), P->reL, P->weightType);
				double nu = (mu_i + A->MEU[j]) > P->epsilonZero ? 2.0 * mu_i * A->MEU[j] / (mu_i + A->MEU[j]) : 0.0;
				nu = nu / rho_div;
				const double* vj = A->vel + 3 * (size_t)j;
				sx += (vj[0] - vm[0]) * w * nu; sy += (vj[1] - vm[1]) * w * nu; sz += (vj[2] - vm[2]) * w * nu;
			}
		});
		{   /* i itself is a neighbour of its mirror image */
			const double ex = xi[0] - mi[0], ey = xi[1] - mi[1], ez = xi[2] - mi[2]; const double dd = ex * ex + ey * ey + ez * ez;
			if (dd < P->reL2) {
				const double w = wgt(sqrt(dd), P->reL, P->weightType);
				double nu = mu_i > P->epsilonZero ? 2.0 * mu_i * mu_i / (mu_i + mu_i) : 0.0;
				nu = nu / rho_div;
				sx += (vi[0] - vm[0]) * w * nu; sy += (vi[1] - vm[1]) * w * nu; sz += (vi[2] - vm[2]) * w * nu;
			}
		}
		for (int k = 0; k < 3; k++) {
			const double t = MATVEC(T, sx, sy, sz, k);
			A->acc[3 * (size_t)i + k] += t * P->coeffViscMultiphase;
			A->forceWall[3 * (size_t)i + k] += -t * P->coeffViscMultiphase * vol * A->RHO[i];
		}
	}
}

/* ------------------------------------------------------------------------------------------------ collisions (PC) */
/* dynamic particle collision (collision type 1), src/MpsParticleCollision.cpp:124-292: pairs closer than partDist; approaching pairs
 * exchange momentum scaled by kappa, separating ones are pushed apart by a background pressure; positions move with the velocity jump */
static void collisions_dpc(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B) {
	const double Wij5 = 0.5 * 0.5 * 0.5 * 0.5 * 3.0, invLo = 1.0 / P->partDist;
	const double gravityMod = sqrt(P->gravityX * P->gravityX + P->gravityY * P->gravityY + P->gravityZ * P->gravityZ);
	double* dv = (double*)calloc(3 * (size_t)A->numParticlesZero, sizeof(double));
	PARFOR
	for (int i = 0; i < A->numParticles; i++) {
		if (A->particleType[i] != P->fluid) continue;
		const double* xi = A->pos + 3 * (size_t)i; const double* mi = A->mirrorParticlePos + 3 * (size_t)i;
		const double* vi = A->vel + 3 * (size_t)i; const double m_i = A->PTYPE[i] == 1 ? P->DNS_FL1 : P->DNS_FL2;
		double dx = 0.0, dy = 0.0, dz = 0.0;
		NEIGH_LOOP(P, A, B, xi, mi, {
			if (d2 < P->partDist * P->partDist && (d2 < dm2 || P->wallType == 0) && j != i) {
				const int jf = A->particleType[j] == P->fluid;
				const double m_j = jf ? (A->PTYPE[j] == 1 ? P->DNS_FL1 : P->DNS_FL2) : P->densityWall;
				const double dst = sqrt(d2);
				double Wij = 0.0;
				if (dst > P->epsilonZero) { const double w1 = 1.0 - dst * invLo; const double w2 = 4.0 * dst * invLo + 1.0; Wij = w1 * w1 * w1 * w1 * w2; }
				const double chi = sqrt(Wij / Wij5);
				const double kappa = dst < 0.5 * P->partDist ? 1.0 : chi;
				const double* vj = A->vel + 3 * (size_t)j;
				double f = (vi[0] - vj[0]) * rx + (vi[1] - vj[1]) * ry + (vi[2] - vj[2]) * rz;
				if (f > 0.0) f *= kappa * 2.0 * m_j / (m_i + m_j) / d2;
				else {
					const double pmax = 2.0 / 3.0 * A->RHO[i] * gravityMod * 0.2, pmin = A->RHO[i] * gravityMod * P->partDist;
					const double a = P->lambdaCollision * fabs(A->press[i] + A->press[j]), b = P->lambdaCollision * pmax;
					const double lo_ = a < b ? a : b; const double ptil = lo_ > pmin ? lo_ : pmin;
					const double pb = ptil * chi;
					f = P->timeStep / m_i * chi * pb / d2;
				}
				if (jf) { dx -= rx * f; dy -= ry * f; dz -= rz * f; }
				else { dx -= 2.0 * rx * f; dy -= 2.0 * ry * f; dz -= 2.0 * rz * f; }
			}
		});
		dv[3 * (size_t)i] = dx; dv[3 * (size_t)i + 1] = dy; dv[3 * (size_t)i + 2] = dz;
	}
	for (int i = 0; i < A->numParticles; i++)
		if (A->particleType[i] == P->fluid) for (int k = 0; k < 3; k++) { A->vel[3 * (size_t)i + k] += dv[3 * (size_t)i + k]; A->pos[3 * (size_t)i + k] += dv[3 * (size_t)i + k] * P->timeStep; }
	free(dv);
}

void oracle_collisions(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B) {
	if (P->collisionType == 1) { collisions_dpc(P, A, B); return; }
	double* dv = (double*)calloc(3 * (size_t)A->numParticlesZero, sizeof(double));
	PARFOR
	for (int i = 0; i < A->numParticles; i++) {
		if (A->particleType[i] != P->fluid) continue;
		const double* xi = A->pos + 3 * (size_t)i; const double* mi = A->mirrorParticlePos + 3 * (size_t)i;
		const double* vi = A->vel + 3 * (size_t)i; const double m_i = A->PTYPE[i] == 1 ? P->DNS_FL1 : P->DNS_FL2;
		double dx = 0.0, dy = 0.0, dz = 0.0;
		NEIGH_LOOP(P, A, B, xi, mi, {
			if (d2 < P->distCollisionLimit2 && (d2 < dm2 || P->wallType == 0) && j != i) {
				const double* vj = A->vel + 3 * (size_t)j;
				double f = (vi[0] - vj[0]) * rx + (vi[1] - vj[1]) * ry + (vi[2] - vj[2]) * rz;
				if (f > 0.0) {
					const int jf = A->particleType[j] == P->fluid;
					const double m_j = jf ? (A->PTYPE[j] == 1 ? P->DNS_FL1 : P->DNS_FL2) : P->densityWall;
					f *= P->restitutionCollision * m_j / (m_i + m_j) / d2;
					if (jf) { dx -= rx * f; dy -= ry * f; dz -= rz * f; }
					else { dx -= 2.0 * rx * f; dy -= 2.0 * ry * f; dz -= 2.0 * rz * f; }
				}
			}
		});
		dv[3 * (size_t)i] = dx; dv[3 * (size_t)i + 1] = dy; dv[3 * (size_t)i + 2] = dz;
	}
	for (int i = 0; i < A->numParticles; i++)
		if (A->particleType[i] == P->fluid) for (int k = 0; k < 3; k++) A->vel[3 * (size_t)i + k] += dv[3 * (size_t)i + k];
	free(dv);
}

/* ------------------------------------------------------------------------------------------------ PND family */
void oracle_wall_npcd(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B) {
	PARFOR
	for (int i = 0; i < A->numParticles; i++) {
		if (!(A->particleType[i] == P->fluid && A->particleNearWall[i])) continue;
		const double* xi = A->pos + 3 * (size_t)i; const double* mi = A->mirrorParticlePos + 3 * (size_t)i;
		double n[3], R[9]; wall_normal(xi, mi, P->epsilonZero, 0, n); reflection(n, R);
		double sx = 0.0, sy = 0.0, sz = 0.0;
		NEIGH_LOOP(P, A, B, xi, mi, {
			if (PRED_M(P->reS2) && j != i) { const double w = wgt(sqrt(dm2), P->reS, P->weightType); sx += mx * w; sy += my * w; sz += mz * w; }
		});
		const double ex = xi[0] - mi[0], ey = xi[1] - mi[1], ez = xi[2] - mi[2]; const double dd = ex * ex + ey * ey + ez * ez;
		if (dd < P->reS2) { const double w = wgt(sqrt(dd), P->reS, P->weightType); sx += ex * w; sy += ey * w; sz += ez * w; }
		for (int k = 0; k < 3; k++) A->npcdDeviation[3 * (size_t)i + k] += MATVEC(R, sx, sy, sz, k);
	}
}

/* shared tail of calcPndnNeighNPCD / setInitialPndNumberOfNeighNPCD (src/MpsPndNeigh.cpp:287-305, :81-100) */
static inline void npcd_finish(PmpsHostArrays* A, const PmpsParams* P, int i, double wsum) {
	double* d = A->npcdDeviation + 3 * (size_t)i;
	if (wsum > P->epsilonZero) { d[0] /= A->pndSmall[i]; d[1] /= A->pndSmall[i]; d[2] /= A->pndSmall[i]; }
	A->npcdDeviation2[i] = d[0] * d[0] + d[1] * d[1] + d[2] * d[2];
	const double* pn = A->polygonNormal + 3 * (size_t)i;
	A->deviationDotPolygonNormal[i] = (d[0] * pn[0] + d[1] * pn[1] + d[2] * pn[2] < 0.0) ? 1 : -1;
}

void oracle_pnd_neigh_npcd(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B) {
	PARFOR
	for (int i = 0; i < A->numParticles; i++) {
		const double* xi = A->pos + 3 * (size_t)i; const double* mi = A->mirrorParticlePos + 3 * (size_t)i;
		double* dev = A->npcdDeviation + 3 * (size_t)i;   /* accumulates ON TOP of the wall part written by wall_npcd */
		double ni = 0.0, wsum = 0.0; int cnt = A->numNeighWallContribution[i];
		NEIGH_LOOP(P, A, B, xi, mi, {
			if (PRED_P(P, P->reL2) && j != i) {
				cnt += 1;
				if (d2 < P->reS2) { const double w = wgt(sqrt(d2), P->reS, P->weightType); ni += w; dev[0] += rx * w; dev[1] += ry * w; dev[2] += rz * w; wsum += w; }
			}
		});
		A->numNeigh[i] = cnt;
		if (P->pndType == 0 || P->pndType == 1) { A->pndki[i] = A->pndi[i]; A->pndi[i] = ni + A->pndWallContribution[i]; }
		A->pndSmall[i] = ni + A->pndWallContribution[i];
		npcd_finish(A, P, i, wsum);
	}
}

void oracle_initial_pnd(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B) {
	PARFOR
	for (int i = 0; i < A->numParticles; i++) {
		const double* xi = A->pos + 3 * (size_t)i; const double* mi = A->mirrorParticlePos + 3 * (size_t)i;
		double* dev = A->npcdDeviation + 3 * (size_t)i; dev[0] = dev[1] = dev[2] = 0.0;
		double wsum = 0.0, acc = A->pndi[i]; int cnt = 0;
		NEIGH_LOOP(P, A, B, xi, mi, {
			if (PRED_P(P, P->reL2) && j != i) {
				cnt += 1;
				if (d2 < P->reS2) { const double w = wgt(sqrt(d2), P->reS, P->weightType); acc += w; dev[0] += rx * w; dev[1] += ry * w; dev[2] += rz * w; wsum += w; }
			}
		});
		acc += A->pndWallContribution[i];
		if (A->particleType[i] == P->wall) acc = P->pndSmallZero;
		A->pndi[i] = acc; A->pndSmall[i] = acc; A->pndki[i] = acc;
		A->numNeigh[i] = cnt + A->numNeighWallContribution[i];
		npcd_finish(A, P, i, wsum);
	}
}

void oracle_pnd_diffusive(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B) {
	PARFOR
	for (int i = 0; i < A->numParticles; i++) {
		const double ni = A->pndi[i];
		if (ni < P->epsilonZero) continue;
		const double* xi = A->pos + 3 * (size_t)i; const double* mi = A->mirrorParticlePos + 3 * (size_t)i;
		const double* vi = A->vel + 3 * (size_t)i; const double Pi = A->press[i];
		const double *c1 = A->correcMatrixRow1 + 3 * (size_t)i, *c2 = A->correcMatrixRow2 + 3 * (size_t)i, *c3 = A->correcMatrixRow3 + 3 * (size_t)i;
		double Di = 0.0, div = 0.0;
		NEIGH_LOOP(P, A, B, xi, mi, {
			if (PRED_P(P, P->reL2) && j != i) {
				const double r = sqrt(d2); const double wL = wgt(r, P->reL, P->weightType);
				if (A->particleType[i] == P->fluid && A->particleType[j] == P->fluid) {
					const double pgh = -2.0 * (A->RHO[i] * A->RHO[j] / (A->RHO[i] + A->RHO[j])) * (P->gravityX * rx + P->gravityY * ry + P->gravityZ * rz);
					Di += P->timeStep / A->RHO[i] * P->coeffViscMultiphase * (A->press[j] - Pi - pgh) * wL;
				}
				if (d2 < P->reS2) {
					const double* vj = A->vel + 3 * (size_t)j;
					const double ux = vj[0] - vi[0], uy = vj[1] - vi[1], uz = vj[2] - vi[2];
					const double wS = wgt(r, P->reS, P->weightType);
					if (!P->gradientCorrection) {
						if (ni > P->epsilonZero) div += ((double)P->dim / P->pndSmallZero) * (A->pndi[j] / ni) * (ux * rx + uy * ry + uz * rz) * wS / d2;
					} else {
						const double qx = (rx * c1[0] + ry * c1[1] + rz * c1[2]), qy = (rx * c2[0] + ry * c2[1] + rz * c2[2]), qz = (rx * c3[0] + ry * c3[1] + rz * c3[2]);
						div += ((double)P->dim / P->pndSmallZero) * (ux * qx + uy * qy + uz * qz) * wS / d2;
					}
				}
			}
		});
		A->pndAuxVar1[i] = A->pndi[i] * (1.0 + P->timeStep * (-div + Di * 1.0));
		A->velDivergence[i] = div; A->diffusiveTerm[i] = Di;
	}
}

static void wall_pnd_diffusive_impl(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B, int divergence);
void oracle_wall_pnd_diffusive(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B) { wall_pnd_diffusive_impl(P, A, B, 0); }
/* calcWallNoSlipVelDivergence / calcWallSlipVelDivergence, src/MpsPressure.cpp:554-788: the same sum, added to velDivergence */
void oracle_wall_vel_divergence(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B) { wall_pnd_diffusive_impl(P, A, B, 1); }
static void wall_pnd_diffusive_impl(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B, int divergence) {
	PARFOR
	for (int i = 0; i < A->numParticles; i++) {
		const double ni = A->pndi[i];
		if (!(A->particleType[i] == P->fluid && A->particleNearWall[i] && ni > P->epsilonZero)) continue;
		const double* xi = A->pos + 3 * (size_t)i; const double* mi = A->mirrorParticlePos + 3 * (size_t)i;
		const double* vi = A->vel + 3 * (size_t)i;
		double n[3], R[9], vm[3]; wall_normal(xi, mi, P->epsilonZero, 1, n); reflection(n, R); mirror_velocity(P, A, i, n, R, vm);
		double div = 0.0;
		NEIGH_LOOP(P, A, B, xi, mi, {
			if (PRED_M(P->reS2) && j != i) {
				const double wS = wgt(sqrt(dm2), P->reS, P->weightType);
				const double* vj = A->vel + 3 * (size_t)j;
				double ux = -(vj[0] - vm[0]), uy = -(vj[1] - vm[1]), uz = -(vj[2] - vm[2]);
				double qx = MATVEC(R, mx, my, mz, 0), qy = MATVEC(R, mx, my, mz, 1), qz = MATVEC(R, mx, my, mz, 2);
				if (P->slipCondition == 1) { ux = vj[0] - vm[0]; uy = vj[1] - vm[1]; uz = vj[2] - vm[2]; qx = mx; qy = my; qz = mz; }   /* :556-560 */
				div += ((double)P->dim / P->pndSmallZero) * (A->pndi[j] / ni) * (ux * qx + uy * qy + uz * qz) * wS / dm2;
			}
		});
		const double ex = xi[0] - mi[0], ey = xi[1] - mi[1], ez = xi[2] - mi[2]; const double dd = ex * ex + ey * ey + ez * ez;
		if (dd < P->reS2) {
			const double wS = wgt(sqrt(dd), P->reS, P->weightType);
			double ux = -(vi[0] - vm[0]), uy = -(vi[1] - vm[1]), uz = -(vi[2] - vm[2]);
			double qx = MATVEC(R, ex, ey, ez, 0), qy = MATVEC(R, ex, ey, ez, 1), qz = MATVEC(R, ex, ey, ez, 2);
			if (P->slipCondition == 1) { ux = vi[0] - vm[0]; uy = vi[1] - vm[1]; uz = vi[2] - vm[2]; qx = ex; qy = ey; qz = ez; }   /* :580-583 */
			div += ((double)P->dim / P->pndSmallZero) * (A->pndi[i] / ni) * (ux * qx + uy * qy + uz * qz) * wS / dd;
		}
		if (divergence) A->velDivergence[i] += div;
		else A->pndAuxVar1[i] += -A->pndi[i] * P->timeStep * div;
	}
}

void oracle_vel_divergence(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B) {   /* src/MpsPressure.cpp:479-551 */
	PARFOR
	for (int i = 0; i < A->numParticles; i++) {
		A->velDivergence[i] = 0.0;
		if (A->particleType[i] != P->fluid) continue;
		const double* xi = A->pos + 3 * (size_t)i; const double* mi = A->mirrorParticlePos + 3 * (size_t)i; const double* vi = A->vel + 3 * (size_t)i;
		const double ni = A->pndi[i]; double div = 0.0;
		const double *c1 = A->correcMatrixRow1 + 3 * (size_t)i, *c2 = A->correcMatrixRow2 + 3 * (size_t)i, *c3 = A->correcMatrixRow3 + 3 * (size_t)i;
		NEIGH_LOOP(P, A, B, xi, mi, {
			if (PRED_P(P, P->reS2) && j != i) {
				const double* vj = A->vel + 3 * (size_t)j;
				const double ux = vj[0] - vi[0], uy = vj[1] - vi[1], uz = vj[2] - vi[2];
				const double wS = wgt(sqrt(d2), P->reS, P->weightType);
				if (!P->gradientCorrection) { if (ni > P->epsilonZero) div += ((double)P->dim / P->pndSmallZero) * (A->pndi[j] / ni) * (ux * rx + uy * ry + uz * rz) * wS / d2; }
				else {
					const double qx = (rx * c1[0] + ry * c1[1] + rz * c1[2]), qy = (rx * c2[0] + ry * c2[1] + rz * c2[2]), qz = (rx * c3[0] + ry * c3[1] + rz * c3[2]);
					div += ((double)P->dim / P->pndSmallZero) * (ux * qx + uy * qy + uz * qz) * wS / d2;
				}
			}
		});
		A->velDivergence[i] = div;
	}
}

/* weighted mean of neighbours' pndi (two-pass: all sums first, then the writes), src/MpsPndNeigh.cpp:777-852 */
void oracle_mean_pnd_wall_dummy_surface(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B) {
	PARFOR
	for (int i = 0; i < A->numParticles; i++) {
		A->pndAuxVar1[i] = 0.0; A->pndAuxVar2[i] = 0.0;
		if (!(A->particleType[i] == P->wall || A->particleBC[i] == P->surface)) continue;
		const double* xi = A->pos + 3 * (size_t)i; const double* mi = A->mirrorParticlePos + 3 * (size_t)i;
		double up = 0.0, dn = 0.0;
		NEIGH_LOOP(P, A, B, xi, mi, {
			if (PRED_P(P, P->reS2) && j != i) { const double w = wgt(sqrt(d2), P->reS, P->weightType); up += A->pndi[j] * w; dn += w; }
		});
		A->pndAuxVar1[i] = up; A->pndAuxVar2[i] = dn;
	}
	for (int i = 0; i < A->numParticles; i++) {
		if (A->particleType[i] == P->wall || A->particleBC[i] == P->surface)
			A->pndi[i] = A->pndAuxVar2[i] < P->epsilonZero ? A->pndAuxVar1[i] : A->pndAuxVar1[i] / (A->pndAuxVar2[i]);
		A->pndAuxVar1[i] = 0.0; A->pndAuxVar2[i] = 0.0;
	}
}

void oracle_mean_wall_pnd(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B) {   /* src/MpsPndNeigh.cpp:923-994 */
	PARFOR
	for (int i = 0; i < A->numParticles; i++) {
		A->pndAuxVar1[i] = 0.0; A->pndAuxVar2[i] = 0.0;
		if (!A->particleNearWall[i]) continue;
		const double* xi = A->pos + 3 * (size_t)i; const double* mi = A->mirrorParticlePos + 3 * (size_t)i;
		double up = 0.0, dn = 0.0;
		NEIGH_LOOP(P, A, B, xi, mi, {
			if (PRED_M(P->reS2) && j != i) { const double w = wgt(sqrt(dm2), P->reS, P->weightType); up += A->pndi[j] * w; dn += w; }
		});
		const double ex = xi[0] - mi[0], ey = xi[1] - mi[1], ez = xi[2] - mi[2]; const double dd = ex * ex + ey * ey + ez * ez;
		if (dd < P->reS2) { const double w = wgt(sqrt(dd), P->reS, P->weightType); up += A->pndi[i] * w; dn += w; }
		A->pndAuxVar1[i] = up; A->pndAuxVar2[i] = dn;
	}
}

void oracle_mean_pnd(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B) {        /* src/MpsPndNeigh.cpp:855-920 */
	PARFOR
	for (int i = 0; i < A->numParticles; i++) {
		const double* xi = A->pos + 3 * (size_t)i; const double* mi = A->mirrorParticlePos + 3 * (size_t)i;
		double up = 0.0, dn = 0.0;
		NEIGH_LOOP(P, A, B, xi, mi, {
			if (PRED_P(P, P->reS2) && j != i) { const double w = wgt(sqrt(d2), P->reS, P->weightType); up += A->pndi[j] * w; dn += w; }
		});
		A->pndAuxVar1[i] += up; A->pndAuxVar2[i] += dn;
	}
	for (int i = 0; i < A->numParticles; i++) {
		A->pndi[i] = A->numNeigh[i] < 1 ? A->pndAuxVar1[i] : A->pndAuxVar1[i] / A->pndAuxVar2[i];
		A->pndAuxVar1[i] = 0.0; A->pndAuxVar2[i] = 0.0;
	}
}

void oracle_mean_neigh_fluid_pnd(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B) {   /* src/MpsPndNeigh.cpp:997-1049 */
	PARFOR
	for (int i = 0; i < A->numParticles; i++) {
		const double* xi = A->pos + 3 * (size_t)i; const double* mi = A->mirrorParticlePos + 3 * (size_t)i;
		double up = 0.0, dn = 0.0;
		NEIGH_LOOP(P, A, B, xi, mi, {
			if (PRED_P(P, P->reS2) && j != i && A->particleType[j] == P->fluid) { const double w = wgt(sqrt(d2), P->reS, P->weightType); up += A->pndki[j] * w; dn += w; }
		});
		A->pndski[i] = dn > P->epsilonZero ? up / dn : up;
	}
}

/* ------------------------------------------------------------------------------------------------ pressure gradient */
static inline double grad_phi(const PmpsParams* P, double w, double Pj, double Pi, double Pmin, double ni, double nj, double dd) {
	switch (P->gradientType) {                       /* src/MpsPressure.cpp:1288-1300 */
		case 0: return w * ((Pj - Pmin) / dd);
		case 1: return w * ((Pj + Pi) / dd);
		case 2: return w * ((Pj + Pi - 2.0 * Pmin) / dd);
		default: return (ni > P->epsilonZero && nj > P->epsilonZero) ? w * ((ni * Pj / nj + nj * Pi / ni) / dd) : w;
	}
}

static void press_gradient_impl(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B, int prediction);
void oracle_press_gradient(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B) { press_gradient_impl(P, A, B, 0); }
/* relax_fact < 1 (src/MpsPressure.cpp:932-1028): no correction matrix, acc += (1 - relaxPress) ... */
void oracle_prediction_press_gradient(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B) { press_gradient_impl(P, A, B, 1); }
static void press_gradient_impl(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B, int prediction) {
	PARFOR
	for (int i = 0; i < A->numParticles; i++) {
		const double* xi = A->pos + 3 * (size_t)i; const double* mi = A->mirrorParticlePos + 3 * (size_t)i;
		const double Pi = A->press[i], ni = A->pndi[i]; double Pmin = Pi;
		const double *c1 = A->correcMatrixRow1 + 3 * (size_t)i, *c2 = A->correcMatrixRow2 + 3 * (size_t)i, *c3 = A->correcMatrixRow3 + 3 * (size_t)i;
		if (P->gradientType == 0 || P->gradientType == 2)
			NEIGH_LOOP(P, A, B, xi, mi, { if (PRED_P(P, P->reS2) && j != i) { if (Pmin > A->press[j]) Pmin = A->press[j]; } });
		double sx = 0.0, sy = 0.0, sz = 0.0;
		NEIGH_LOOP(P, A, B, xi, mi, {
			if (PRED_P(P, P->reS2) && j != i) {
				const double w = grad_phi(P, wgt(sqrt(d2), P->reS, P->weightType), A->press[j], Pi, Pmin, ni, A->pndi[j], d2);
				if (prediction || !P->gradientCorrection) { sx += rx * w; sy += ry * w; sz += rz * w; }
				else {
					sx += (rx * c1[0] + ry * c1[1] + rz * c1[2]) * w; sy += (rx * c2[0] + ry * c2[1] + rz * c2[2]) * w;
					sz += (rx * c3[0] + ry * c3[1] + rz * c3[2]) * w;
				}
			}
		});
		if (prediction) {
			A->acc[3 * (size_t)i] += (1.0 - P->relaxPress) * sx * P->coeffPressGrad / A->RHO[i];
			A->acc[3 * (size_t)i + 1] += (1.0 - P->relaxPress) * sy * P->coeffPressGrad / A->RHO[i];
			A->acc[3 * (size_t)i + 2] += (1.0 - P->relaxPress) * sz * P->coeffPressGrad / A->RHO[i];
			continue;
		}
		A->acc[3 * (size_t)i] = P->relaxPress * sx * P->coeffPressGrad / A->RHO[i];
		A->acc[3 * (size_t)i + 1] = P->relaxPress * sy * P->coeffPressGrad / A->RHO[i];
		A->acc[3 * (size_t)i + 2] = P->relaxPress * sz * P->coeffPressGrad / A->RHO[i];
	}
}

/* wall repulsion, src/MpsPressure.cpp:1554-1611: returns the scalar that multiplies -n */
static inline double repulsion(const PmpsParams* P, double r, double rho, double vmax2) {
	const double re = P->reRepulsiveForce;
	switch (P->repulsiveForceType) {
		case 0: return rho / (P->timeStep * P->timeStep) * (re - r);
		case 1: return P->repForceCoefMitsume * wgt(r, re, P->weightType);
		case 2: { const double R1 = (re / r) * (re / r); const double R2 = R1 * R1; return (P->repForceCoefLennardJones * vmax2 / r) * (R2 - R1) * rho; }
		default: { const double W1 = (1.0 + 3.0 * 0.5 * r / (re)); const double W2 = (1.0 - r / (re)) * (1.0 - r / (re)) * (1.0 - r / (re));
			return (P->repForceCoefMonaghanKajtar * vmax2 / (r - 0.0 * P->partDist)) * (1.0 / 8.0) * (W1) * (W2)*rho; }
	}
}

static void wall_press_gradient_impl(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B, double velMax, int prediction);
void oracle_wall_press_gradient(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B, double velMax) { wall_press_gradient_impl(P, A, B, velMax, 0); }
void oracle_prediction_wall_press_gradient(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B, double velMax) { wall_press_gradient_impl(P, A, B, velMax, 1); }   /* src/MpsPressure.cpp:1031-1214 */
static void wall_press_gradient_impl(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B, double velMax, int prediction) {
	const double vol = pow(P->partDist, (double)P->dim);
	const double vcap = velMax < P->expectMaxVelocity ? velMax : P->expectMaxVelocity; const double vmax2 = vcap * vcap;
	PARFOR
	for (int i = 0; i < A->numParticles; i++) {
		if (!(A->particleType[i] == P->fluid && A->particleNearWall[i])) continue;
		const double* xi = A->pos + 3 * (size_t)i; const double* mi = A->mirrorParticlePos + 3 * (size_t)i;
		const double Pi = A->press[i], ni = A->pndi[i]; double Pmin = Pi;
		double n[3], R[9]; const double r = wall_normal(xi, mi, P->epsilonZero, 0, n); reflection(n, R);
		if (P->gradientType == 0 || P->gradientType == 2)
			NEIGH_LOOP(P, A, B, xi, mi, { if (d2 < P->reS2 && d2 < dm2 && j != i) { if (Pmin > A->press[j]) Pmin = A->press[j]; } });
		double sx = 0.0, sy = 0.0, sz = 0.0;
		NEIGH_LOOP(P, A, B, xi, mi, {
			if (PRED_M(P->reS2) && j != i) {
				const double w = grad_phi(P, wgt(sqrt(dm2), P->reS, P->weightType), A->press[j], Pi, Pmin, ni, A->pndi[j], dm2);
				sx += mx * w; sy += my * w; sz += mz * w;
			}
		});
		const double ex = xi[0] - mi[0], ey = xi[1] - mi[1], ez = xi[2] - mi[2]; const double dd = ex * ex + ey * ey + ez * ez;
		if (dd < P->reS2) {
			const double w = grad_phi(P, wgt(sqrt(dd), P->reS, P->weightType), Pi, Pi, Pmin, ni, A->pndi[i], dd);
			sx += ex * w; sy += ey * w; sz += ez * w;
		}
		double f[3] = { 0.0, 0.0, 0.0 };
		if (r < P->reRepulsiveForce && r > P->epsilonZero) { const double q = repulsion(P, r, A->RHO[i], vmax2); f[0] = -q * n[0]; f[1] = -q * n[1]; f[2] = -q * n[2]; }
		for (int k = 0; k < 3; k++) {
			const double g = ((prediction ? 1.0 - P->relaxPress : P->relaxPress) * MATVEC(R, sx, sy, sz, k) * P->coeffPressGrad - f[k]);
			A->acc[3 * (size_t)i + k] += g / A->RHO[i];
			if (!prediction) A->forceWall[3 * (size_t)i + k] += -g * vol;
		}
	}
}

/* ------------------------------------------------------------------------------------------------ shifting (type 2) */
/* The reference updates pos[i] inside the same loop in which later particles read pos[j] (src/MpsShifting.cpp:478-536):
 * with one thread this is a Gauss-Seidel sweep in index order (shiftingJacobi = 0 reproduces it bit for bit); with
 * several threads it is a data race.  shiftingJacobi = 1 reads a snapshot of the positions instead — the only
 * order-independent definition, and the one the CUDA path implements. */
static inline double del_wgt(double r, double re, int type) {   /* src/MpsParticle.cpp:94-108 */
	switch (type) {
		case 1: return -re / (r * r) + 1.0 / re;
		case 2: return -re / (r * r) - 1.0 / re;
		case 3: return -3.0 / re * (1.0 - r / re) * (1.0 - r / re);
		case 4: return -2.0 / re * (1.0 - r / re);
		default: return -re / (r * r);
	}
}
/* shifting type 1 (velocity adjustment), src/MpsShifting.cpp:23-101.  The reference updates vel in place inside the particle
 * loop (a Gauss-Seidel sweep with one thread); shiftingJacobi reads a snapshot instead, as for type 2. */
static void shifting_velocity(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B) {
	const int N = A->numParticles; double* snap = NULL; const double* vsrc = A->vel;
	if (P->shiftingJacobi) { snap = (double*)malloc(sizeof(double) * 3 * (size_t)N); memcpy(snap, A->vel, sizeof(double) * 3 * (size_t)N); vsrc = snap; }
	for (int i = 0; i < N; i++) {
		if (A->particleType[i] != P->fluid) continue;
		const double* xi = A->pos + 3 * (size_t)i; const double* mi = A->mirrorParticlePos + 3 * (size_t)i;
		const double vx = vsrc[3 * (size_t)i], vy = vsrc[3 * (size_t)i + 1], vz = vsrc[3 * (size_t)i + 2];
		double dux = 0.0, duy = 0.0, duz = 0.0;
		NEIGH_LOOP(P, A, B, xi, mi, {
			if (PRED_P(P, P->reS2) && j != i) {
				const double dw = del_wgt(sqrt(d2), P->reS, P->weightType);
				dux += dw * (vsrc[3 * (size_t)j] - vx); duy += dw * (vsrc[3 * (size_t)j + 1] - vy); duz += dw * (vsrc[3 * (size_t)j + 2] - vz);
			}
		});
		if (A->particleBC[i] == P->inner) {
			A->vel[3 * (size_t)i] -= P->coeffShifting1 * dux; A->vel[3 * (size_t)i + 1] -= P->coeffShifting1 * duy; A->vel[3 * (size_t)i + 2] -= P->coeffShifting1 * duz;
		}
	}
	free(snap);
}

void oracle_shifting(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B) {
	if (P->shiftingType == 1) { shifting_velocity(P, A, B); return; }
	PARFOR
	for (int i = 0; i < A->numParticles; i++) {
		const double* xi = A->pos + 3 * (size_t)i; const double* mi = A->mirrorParticlePos + 3 * (size_t)i;
		double c = 0.0;
		NEIGH_LOOP(P, A, B, xi, mi, { if (PRED_P(P, P->reS2) && j != i) c += wgt(sqrt(d2), P->reS, P->weightType); });
		c += A->pndWallContribution[i];
		A->concentration[i] = c / P->pndSmallZero;
	}
	const int jac = P->shiftingJacobi; PmpsHostArrays S = *A; double* snap = NULL;
	if (jac) { snap = (double*)malloc(3 * sizeof(double) * (size_t)A->numParticlesZero); memcpy(snap, A->pos, 3 * sizeof(double) * (size_t)A->numParticlesZero); S.pos = snap; }
	const int nthr = jac ? g_threads : 1;
	#pragma omp parallel for schedule(dynamic,64) num_threads(nthr)
	for (int i = 0; i < A->numParticles; i++) {
		if (A->particleType[i] != P->fluid) continue;
		const double* xi = S.pos + 3 * (size_t)i; const double* mi = A->mirrorParticlePos + 3 * (size_t)i;
		const double ci = A->concentration[i]; double gx = 0.0, gy = 0.0, gz = 0.0;
		NEIGH_LOOP(P, &S, B, xi, mi, {
			if (PRED_P(P, P->reS2) && j != i) {
				const double w = wgt(sqrt(d2), P->reS, P->weightType);
				gx += (ci + A->concentration[j]) * rx * w / d2; gy += (ci + A->concentration[j]) * ry * w / d2; gz += (ci + A->concentration[j]) * rz * w / d2;
			}
		});
		double* g = A->gradConcentration + 3 * (size_t)i;
		g[0] = gx * -P->coeffPressGrad; g[1] = gy * -P->coeffPressGrad; g[2] = gz * -P->coeffPressGrad;
		if (A->particleBC[i] == P->inner) for (int k = 0; k < 3; k++) A->pos[3 * (size_t)i + k] -= P->coeffShifting2 * g[k];
	}
	free(snap);
}

void oracle_wall_shifting(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B) {
	const int jac = P->shiftingJacobi; PmpsHostArrays S = *A; double* snap = NULL;
	if (jac) { snap = (double*)malloc(3 * sizeof(double) * (size_t)A->numParticlesZero); memcpy(snap, A->pos, 3 * sizeof(double) * (size_t)A->numParticlesZero); S.pos = snap; }
	const int nthr = jac ? g_threads : 1;
	#pragma omp parallel for schedule(dynamic,64) num_threads(nthr)
	for (int i = 0; i < A->numParticles; i++) {
		if (!(A->particleType[i] == P->fluid && A->particleNearWall[i])) continue;
		const double* xi = S.pos + 3 * (size_t)i; const double* mi = A->mirrorParticlePos + 3 * (size_t)i;
		const double ci = A->concentration[i];
		double n[3], R[9]; wall_normal(xi, mi, P->epsilonZero, 0, n); reflection(n, R);
		double sx = 0.0, sy = 0.0, sz = 0.0;
		NEIGH_LOOP(P, &S, B, xi, mi, {
			if (PRED_M(P->reS2) && j != i) {
				const double w = wgt(sqrt(dm2), P->reS, P->weightType);
				sx += (ci + A->concentration[j]) * mx * w / dm2; sy += (ci + A->concentration[j]) * my * w / dm2; sz += (ci + A->concentration[j]) * mz * w / dm2;
			}
		});
		const double ex = xi[0] - mi[0], ey = xi[1] - mi[1], ez = xi[2] - mi[2]; const double dd = ex * ex + ey * ey + ez * ez;
		if (dd < P->reS2) {
			const double w = wgt(sqrt(dd), P->reS, P->weightType);
			sx += (ci + ci) * ex * w / dd; sy += (ci + ci) * ey * w / dd; sz += (ci + ci) * ez * w / dd;
		}
		for (int k = 0; k < 3; k++) {
			const double gw = -MATVEC(R, sx, sy, sz, k) * P->coeffPressGrad;
			A->gradConcentration[3 * (size_t)i + k] += gw;
			if (A->particleBC[i] == P->inner) A->pos[3 * (size_t)i + k] -= P->coeffShifting2 * gw;
		}
	}
	free(snap);
}

/* ------------------------------------------------------------------------------------------------ polygon wall */
/* Closest point on a triangle, region tests in the order of Ericson ch.5 as used by
 * libigl point_simplex_squared_distance.cpp:44-110.  dot3 = (a0b0 + a1b1) + a2b2. */
static inline double dot3(const double* a, const double* b) { return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2]; }
static void closest_on_triangle(const double* p, const double* a, const double* b, const double* c, double* q) {
	double ab[3], ac[3], ap[3], bp[3], cp[3];
	for (int k = 0; k < 3; k++) { ab[k] = b[k] - a[k]; ac[k] = c[k] - a[k]; ap[k] = p[k] - a[k]; }
	const double d1 = dot3(ab, ap), d2 = dot3(ac, ap);
	if (d1 <= 0.0 && d2 <= 0.0) { memcpy(q, a, 24); return; }
	for (int k = 0; k < 3; k++) bp[k] = p[k] - b[k];
	const double d3 = dot3(ab, bp), d4 = dot3(ac, bp);
	if (d3 >= 0.0 && d4 <= d3) { memcpy(q, b, 24); return; }
	const double vc = d1 * d4 - d3 * d2;
	if (a[0] != b[0] || a[1] != b[1] || a[2] != b[2]) {
		if (vc <= 0.0 && d1 >= 0.0 && d3 <= 0.0) { const double v = d1 / (d1 - d3); for (int k = 0; k < 3; k++) q[k] = a[k] + v * ab[k]; return; }
	}
	for (int k = 0; k < 3; k++) cp[k] = p[k] - c[k];
	const double d5 = dot3(ab, cp), d6 = dot3(ac, cp);
	if (d6 >= 0.0 && d5 <= d6) { memcpy(q, c, 24); return; }
	const double vb = d5 * d2 - d1 * d6;
	if (vb <= 0.0 && d2 >= 0.0 && d6 <= 0.0) { const double w = d2 / (d2 - d6); for (int k = 0; k < 3; k++) q[k] = a[k] + w * ac[k]; return; }
	const double va = d3 * d6 - d5 * d4;
	if (va <= 0.0 && (d4 - d3) >= 0.0 && (d5 - d6) >= 0.0) {
		const double w = (d4 - d3) / ((d4 - d3) + (d5 - d6)); for (int k = 0; k < 3; k++) q[k] = b[k] + w * (c[k] - b[k]); return;
	}
	const double denom = 1.0 / (va + vb + vc); const double v = vb * denom, w = vc * denom;
	for (int k = 0; k < 3; k++) q[k] = (a[k] + ab[k] * v) + ac[k] * w;
}

/* The libigl AABB hierarchy (AABB.cpp:29-196) in heap order (children of node t at 2t+1, 2t+2), rebuilt here because
 * the ORDER in which it visits equidistant triangles decides which wall a corner particle is mirrored across
 * (first strictly smaller distance wins, AABB.cpp:806-830). */
int oracle_tree_nodes(int nF) { int d = 0; while ((1 << d) < nF) d++; return (1 << (d + 1)) - 1 + 2; }

typedef struct { const double* key; } SortCtx;
static const double* g_sort_key;
static int cmp_key(const void* a, const void* b) {   /* ties keep input order: libstdc++ std::sort on <=16 items is an insertion sort */
	const int ia = *(const int*)a, ib = *(const int*)b;
	if (g_sort_key[ia] < g_sort_key[ib]) return -1; if (g_sort_key[ia] > g_sort_key[ib]) return 1; return ia - ib;
}
static void build_node(OMesh* m, int t, const int* I, int n, const int* SI /* nF x 3 rank table */) {
	double lo[3] = { DBL_MAX, DBL_MAX, DBL_MAX }, hi[3] = { -DBL_MAX, -DBL_MAX, -DBL_MAX };
	for (int e = 0; e < n; e++) for (int c = 0; c < 3; c++) for (int k = 0; k < 3; k++) {
		const double v = m->V[3 * m->F[3 * I[e] + c] + k]; if (v < lo[k]) lo[k] = v; if (v > hi[k]) hi[k] = v;
	}
	for (int k = 0; k < 3; k++) { m->bbMin[3 * t + k] = lo[k]; m->bbMax[3 * t + k] = hi[k]; }
	if (n == 1) { m->prim[t] = I[0]; return; }
	m->prim[t] = -1;
	int ax = 0; for (int k = 1; k < 3; k++) if (hi[k] - lo[k] > hi[ax] - lo[ax]) ax = k;   /* maxCoeff: first maximum */
	int* r = (int*)malloc(sizeof(int) * n); int* s = (int*)malloc(sizeof(int) * n);
	for (int e = 0; e < n; e++) r[e] = s[e] = SI[3 * I[e] + ax];
	for (int a = 1; a < n; a++) { int v = s[a], b = a; while (b > 0 && s[b - 1] > v) { s[b] = s[b - 1]; b--; } s[b] = v; }
	const int med = s[(n - 1) / 2];
	int* L = (int*)malloc(sizeof(int) * n); int* Rr = (int*)malloc(sizeof(int) * n); int nl = 0, nr = 0;
	for (int e = 0; e < n; e++) { if (r[e] <= med) L[nl++] = I[e]; else Rr[nr++] = I[e]; }
	if (nl > 0) build_node(m, 2 * t + 1, L, nl, SI);
	if (nr > 0) build_node(m, 2 * t + 2, Rr, nr, SI);
	free(r); free(s); free(L); free(Rr);
}
void oracle_build_tree(OMesh* m) {
	const int nF = m->nF; m->nNodes = oracle_tree_nodes(nF);
	for (int t = 0; t < m->nNodes; t++) m->prim[t] = -2;
	double* bc = (double*)malloc(sizeof(double) * nF); int* idx = (int*)malloc(sizeof(int) * nF); int* SI = (int*)malloc(sizeof(int) * 3 * nF);
	for (int k = 0; k < 3; k++) {
		for (int f = 0; f < nF; f++) {   /* barycenter: ((v0+v1)+v2)/3, igl/barycenter.cpp:19-31 */
			double s = 0.0; for (int c = 0; c < 3; c++) s += m->V[3 * m->F[3 * f + c] + k];
			bc[f] = s / 3.0; idx[f] = f;
		}
		g_sort_key = bc; qsort(idx, nF, sizeof(int), cmp_key);
		for (int rnk = 0; rnk < nF; rnk++) SI[3 * idx[rnk] + k] = rnk;
	}
	for (int f = 0; f < nF; f++) idx[f] = f;
	build_node(m, 0, idx, nF, SI);
	free(bc); free(idx); free(SI);
}

static inline double box_ext2(const double* lo, const double* hi, const double* p) {   /* Eigen AlignedBox.h:408-427 */
	double d = 0.0;
	for (int k = 0; k < 3; k++) { if (lo[k] > p[k]) { const double a = lo[k] - p[k]; d += a * a; } else if (p[k] > hi[k]) { const double a = p[k] - hi[k]; d += a * a; } }
	return d;
}
static inline int box_has(const double* lo, const double* hi, const double* p) {
	return lo[0] <= p[0] && p[0] <= hi[0] && lo[1] <= p[1] && p[1] <= hi[1] && lo[2] <= p[2] && p[2] <= hi[2];
}
/* AABB::squared_distance (AABB.cpp:363-450): returns the best squared distance found under node t given bound `best` */
static double tree_query(const OMesh* m, int t, const double* p, double best, int* face, double* c) {
	double sq = best;
	if (m->prim[t] >= 0) {
		const int f = m->prim[t]; double q[3];
		closest_on_triangle(p, m->V + 3 * m->F[3 * f], m->V + 3 * m->F[3 * f + 1], m->V + 3 * m->F[3 * f + 2], q);
		const double e0 = p[0] - q[0], e1 = p[1] - q[1], e2 = p[2] - q[2]; const double dd = (e0 * e0 + e1 * e1) + e2 * e2;
		if (dd < sq) { sq = dd; *face = f; memcpy(c, q, 24); }
		return sq;
	}
	const int L = 2 * t + 1, Rn = 2 * t + 2; int seenL = 0, seenR = 0;
#define LOOK(child, seen) do { int f_ = -1; double c_[3]; memcpy(c_, c, 24); const double s_ = tree_query(m, child, p, sq, &f_, c_); \
		if (s_ < sq) { sq = s_; *face = f_; memcpy(c, c_, 24); } seen = 1; } while (0)
	if (box_has(m->bbMin + 3 * L, m->bbMax + 3 * L, p)) LOOK(L, seenL);
	if (box_has(m->bbMin + 3 * Rn, m->bbMax + 3 * Rn, p)) LOOK(Rn, seenR);
	const double dl = box_ext2(m->bbMin + 3 * L, m->bbMax + 3 * L, p), dr = box_ext2(m->bbMin + 3 * Rn, m->bbMax + 3 * Rn, p);
	if (dl < dr) { if (!seenL && dl < sq) LOOK(L, seenL); if (!seenR && dr < sq) LOOK(Rn, seenR); }
	else { if (!seenR && dr < sq) LOOK(Rn, seenR); if (!seenL && dl < sq) LOOK(L, seenL); }
#undef LOOK
	return sq;
}
void oracle_closest_point(const OMesh* m, const double p[3], double* sqrD, int* face, double c[3]) {
	*face = -1; c[0] = c[1] = c[2] = 0.0;
	*sqrD = tree_query(m, 0, p, INFINITY, face, c);
}

/* table lerp with extrapolation, src/PolygonMesh.cpp:533-581 */
static inline double table_lerp(const double* xs, const double* ys, const int32_t* yi, int n, double x) {
	int k = 0;
	if (x >= xs[n - 2]) k = n - 2; else while (x > xs[k + 1]) k++;
	const double xL = xs[k], xR = xs[k + 1]; const double yL = ys ? ys[k] : (double)yi[k], yR = ys ? ys[k + 1] : (double)yi[k + 1];
	const double dydx = (yR - yL) / (xR - xL);
	return yL + dydx * (x - xL);
}

void oracle_wall_closest_point(const PmpsParams* P, PmpsHostArrays* A, const OMesh* meshes) {
	for (int me = 0; me < P->numOfMeshs; me++) {   /* src/PolygonMesh.cpp:585-651; mesh order rigid, deformable, forced */
		const OMesh* m = meshes + me;
		PARFOR
		for (int i = 0; i < A->numParticles; i++) {
			const double* x = A->pos + 3 * (size_t)i; double sq, c[3]; int f;
			oracle_closest_point(m, x, &sq, &f, c);
			if (sq < A->distParticleWall2[i]) {
				A->distParticleWall2[i] = sq;
				for (int k = 0; k < 3; k++) {
					A->particleAtWallPos[3 * (size_t)i + k] = c[k];
					A->mirrorParticlePos[3 * (size_t)i + k] = x[k] + 2.0 * (c[k] - x[k]);
					A->polygonNormal[3 * (size_t)i + k] = m->N[3 * f + k];
				}
				/* mesh slot -> type id: slot 1 is the deformable mesh when femOn, slot 2 the forced one (src/PolygonMesh.cpp:630-640) */
				A->nearMeshType[i] = me;
				if (me == 1) A->elementID[i] = f;
			}
		}
	}
	const OMesh* m0 = meshes; const double reS = sqrt(P->reS2), reL = sqrt(P->reL2);   /* src/PolygonMesh.cpp:718-795 */
	for (int i = 0; i < A->numParticles; i++) {
		A->pndWallContribution[i] = 0.0; A->numNeighWallContribution[i] = 0;
		const double x2 = A->distParticleWall2[i];
		if (x2 < P->reL2) {
			A->particleNearWall[i] = 1; const double x = sqrt(x2);
			A->pndWallContribution[i] = table_lerp(m0->xDataPND, m0->Zij, NULL, m0->nTable, x / reS);
			A->numNeighWallContribution[i] = (int)table_lerp(m0->xDataNeigh, NULL, m0->nNeighWall, m0->nTable, x / reL);
		}
	}
}

/* ------------------------------------------------------------------------------------------------ gradient correction */
static void invert3(double M[9], int dim, double eps) {   /* Gauss-Jordan without pivoting, src/MpsVectorMatrix.cpp:96-166 */
	double a[3][3], inv[3][3];
	for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) { a[r][c] = M[3 * r + c]; inv[r][c] = r == c ? 1.0 : 0.0; }
	if (dim == 2) { a[0][2] = a[1][2] = a[2][0] = a[2][1] = 0.0; a[2][2] = 1.0; }
	const double det = (a[0][0] * a[1][1] * a[2][2] + a[0][1] * a[1][2] * a[2][0] + a[0][2] * a[1][0] * a[2][1])
	                 - (a[0][2] * a[1][1] * a[2][0] + a[0][1] * a[1][0] * a[2][2] + a[0][0] * a[1][2] * a[2][1]);
	if (det <= eps) { for (int k = 0; k < 9; k++) M[k] = (k % 4 == 0) ? 1.0 : 0.0; return; }
	for (int k = 0; k < dim; k++) for (int r = 0; r < dim; r++) {
		if (r == k) continue;
		const double m = a[r][k] / a[k][k];
		for (int c = 0; c < dim; c++) { a[r][c] -= m * a[k][c]; inv[r][c] -= m * inv[k][c]; }
	}
	for (int r = 0; r < dim; r++) for (int c = 0; c < dim; c++) inv[r][c] /= a[r][r];
	for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) M[3 * r + c] = inv[r][c];
}

void oracle_correction_matrix(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B) {
	PARFOR
	for (int i = 0; i < A->numParticles; i++) {
		const double* xi = A->pos + 3 * (size_t)i; const double* mi = A->mirrorParticlePos + 3 * (size_t)i;
		double M[9] = { 0, 0, 0, 0, 0, 0, 0, 0, 0 };
		NEIGH_LOOP(P, A, B, xi, mi, {
			if (PRED_P(P, P->reS2) && j != i) {
				const double w = wgt(sqrt(d2), P->reS, P->weightType); const double inv = 1.0 / d2; const double r[3] = { rx, ry, rz };
				for (int a = 0; a < 3; a++) for (int b = 0; b < 3; b++) M[3 * a + b] += w * r[a] * r[b] * inv;
			}
		});
		for (int k = 0; k < 9; k++) M[k] *= -P->coeffPressGrad;
		invert3(M, P->dim, P->epsilonZero);
		if (A->numNeigh[i] < 4) for (int k = 0; k < 9; k++) M[k] = (k % 4 == 0) ? 1.0 : 0.0;
		for (int k = 0; k < 3; k++) { A->correcMatrixRow1[3 * (size_t)i + k] = M[k]; A->correcMatrixRow2[3 * (size_t)i + k] = M[3 + k]; A->correcMatrixRow3[3 * (size_t)i + k] = M[6 + k]; }
	}
}

/* ------------------------------------------------------------------------------------------------ PPE (incompressible) */
/* Assembly follows src/MpsPressure.cpp:169-286.  Eigen is told Lower|Upper on a column-major matrix, so its CG
 * multiplies by the TRANSPOSE of the assembled matrix (ConjugateGradient.h:200-207); A is symmetric except where a
 * polygon-wall mirror test is one-sided, and we keep that detail: y = A^T x.  The Krylov recurrence is Eigen's
 * (ConjugateGradient.h:27-91) with the Jacobi preconditioner of BasicPreconditioners.h:67-91, tolerance = DBL_EPSILON,
 * maxIt = 2N, warm start from pressurePPE.  Eigen's dot products are SIMD-blocked, so this part is pinned to the
 * reference by tolerance on the converged pressure, not bit for bit. */
typedef struct { int n; int* rowStart; int* col; double* val; double* diag; } Csr;

int oracle_solve_ppe(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B, double* errOut) {
	const int N = A->numParticles; double* b = A->sourceTerm; double* x = A->pressurePPE;
	int* cnt = (int*)calloc((size_t)N + 1, sizeof(int));
	/* rows of A (row-wise triplets, ascending j inside a row is NOT guaranteed by the walk, so sort later via transpose) */
	int cap = N * 64 + 16, nnz = 0; int* ti = (int*)malloc(sizeof(int) * cap); int* tj = (int*)malloc(sizeof(int) * cap); double* tv = (double*)malloc(sizeof(double) * cap);
	double* diag = (double*)malloc(sizeof(double) * N);
	for (int i = 0; i < N; i++) {
		b[i] = 0.0;
		if (A->particleBC[i] == P->other || A->particleBC[i] == P->surface) { diag[i] = 1.0; continue; }
		const double* xi = A->pos + 3 * (size_t)i; const double* mi = A->mirrorParticlePos + 3 * (size_t)i;
		double sum = 0.0;
		NEIGH_LOOP(P, A, B, xi, mi, {
			if (PRED_P(P, P->reL2) && j != i) {
				const double wL = wgt(sqrt(d2), P->reL, P->weightType); const double a = wL * P->coeffPPE;
				if (A->particleType[j] == P->dummyWall) {
					const double pgh = A->RHO[j] * (rx * P->gravityX + ry * P->gravityY + rz * P->gravityZ) * wL;
					b[i] -= P->coeffPPE * pgh;
				} else {
					sum -= a;
					if (A->particleBC[j] == P->inner) {
						if (nnz == cap) { cap *= 2; ti = (int*)realloc(ti, sizeof(int) * cap); tj = (int*)realloc(tj, sizeof(int) * cap); tv = (double*)realloc(tv, sizeof(double) * cap); }
						ti[nnz] = i; tj[nnz] = j; tv[nnz] = a; nnz++;
					}
				}
			}
		});
		const double rho = A->PTYPE[i] == 1 ? P->DNS_FL1 : P->DNS_FL2;
		sum -= P->alphaCompressibility * rho / (P->timeStep * P->timeStep);
		diag[i] = sum;
		if (P->mpsType == 3) b[i] += -P->coeffPPESource * rho * (A->pndki[i] - P->pndSmallZero) + (1.0 - P->relaxPND) * rho * A->velDivergence[i] / P->timeStep;   /* src/MpsPressure.cpp:373 */
		else b[i] += -P->coeffPPESource * rho * (A->pndSmall[i] - P->pndSmallZero);
	}
	/* column-compressed storage of A == row-compressed storage of A^T: bucket the off-diagonals by column j */
	for (int k = 0; k < nnz; k++) cnt[tj[k] + 1]++;
	for (int i = 0; i < N; i++) cnt[i + 1] += cnt[i];
	int* fill = (int*)malloc(sizeof(int) * ((size_t)N + 1)); memcpy(fill, cnt, sizeof(int) * ((size_t)N + 1));
	int* ci = (int*)malloc(sizeof(int) * (nnz + 1)); double* cv = (double*)malloc(sizeof(double) * (nnz + 1));
	for (int k = 0; k < nnz; k++) { const int s = fill[tj[k]]++; ci[s] = ti[k]; cv[s] = tv[k]; }   /* rows ascending since i ascends */
#define SPMV_T(y, v) do { for (int c_ = 0; c_ < N; c_++) { double t_ = 0.0; int dd_ = 0;                                  \
		for (int s_ = cnt[c_]; s_ < cnt[c_ + 1]; s_++) { if (!dd_ && ci[s_] > c_) { t_ += diag[c_] * (v)[c_]; dd_ = 1; } t_ += cv[s_] * (v)[ci[s_]]; } \
		if (!dd_) t_ += diag[c_] * (v)[c_]; (y)[c_] = t_; } } while (0)
	double *r = (double*)malloc(sizeof(double) * N), *p = (double*)malloc(sizeof(double) * N), *z = (double*)malloc(sizeof(double) * N), *t = (double*)malloc(sizeof(double) * N);
	int iters = 0; double err = 0.0;
	double rhs2 = 0.0, res2 = 0.0;
	if (P->solverType == 1) {
		/* BiCGSTAB multiplies by A itself (Eigen BiCGSTAB.h:29-107; Jacobi preconditioner, tolerance DBL_EPSILON, maxIt 2N, warm start).
		 * The triplets are in row order (i ascends in the assembly loop): row-compressed A straight from them. */
		int* rs = (int*)calloc((size_t)N + 1, sizeof(int));
		for (int k = 0; k < nnz; k++) rs[ti[k] + 1]++;
		for (int i = 0; i < N; i++) rs[i + 1] += rs[i];
#define SPMV_A(y, v) do { for (int r_ = 0; r_ < N; r_++) { double t_ = diag[r_] * (v)[r_]; for (int s_ = rs[r_]; s_ < rs[r_ + 1]; s_++) t_ += tv[s_] * (v)[tj[s_]]; (y)[r_] = t_; } } while (0)
#define DOT(a, c, out) do { double d_ = 0.0; for (int i_ = 0; i_ < N; i_++) d_ += (a)[i_] * (c)[i_]; (out) = d_; } while (0)
#define DINV(i) (diag[i] != 0.0 ? 1.0 / diag[i] : 1.0)
		double *r0 = (double*)malloc(sizeof(double) * N), *v = (double*)calloc(N, sizeof(double)), *sv = (double*)malloc(sizeof(double) * N), *y = (double*)malloc(sizeof(double) * N);
		for (int i = 0; i < N; i++) p[i] = 0.0;
		SPMV_A(t, x);
		for (int i = 0; i < N; i++) { r[i] = b[i] - t[i]; r0[i] = r[i]; }
		double r0n; DOT(r0, r0, r0n); DOT(b, b, rhs2);
		if (rhs2 == 0.0) { for (int i = 0; i < N; i++) x[i] = 0.0; iters = 2 * N; err = DBL_EPSILON; }
		else {
			double rho = 1.0, alpha = 1.0, w = 1.0; const double tol2 = DBL_EPSILON * DBL_EPSILON * rhs2, eps2 = DBL_EPSILON * DBL_EPSILON;
			const int maxIt = 2 * N; int restarts = 0;
			DOT(r, r, res2);
			while (res2 > tol2 && iters < maxIt) {
				const double rho_old = rho;
				DOT(r0, r, rho);
				if (fabs(rho) < eps2 * r0n) {
					SPMV_A(t, x);
					for (int i = 0; i < N; i++) { r[i] = b[i] - t[i]; r0[i] = r[i]; }
					DOT(r, r, rho); r0n = rho;
					if (restarts++ == 0) iters = 0;
				}
				const double beta = (rho / rho_old) * (alpha / w);
				for (int i = 0; i < N; i++) { p[i] = r[i] + beta * (p[i] - w * v[i]); y[i] = DINV(i) * p[i]; }
				SPMV_A(v, y);
				double r0v; DOT(r0, v, r0v); alpha = rho / r0v;
				for (int i = 0; i < N; i++) { sv[i] = r[i] - alpha * v[i]; z[i] = DINV(i) * sv[i]; }
				SPMV_A(t, z);
				double tt, ts; DOT(t, t, tt); DOT(t, sv, ts);
				w = tt > 0.0 ? ts / tt : 0.0;
				for (int i = 0; i < N; i++) { x[i] += alpha * y[i] + w * z[i]; r[i] = sv[i] - w * t[i]; }
				DOT(r, r, res2);
				++iters;
			}
			err = sqrt(res2 / rhs2);
		}
#undef SPMV_A
#undef DOT
#undef DINV
		free(rs); free(r0); free(v); free(sv); free(y);
	} else {
	SPMV_T(t, x);
	for (int i = 0; i < N; i++) { r[i] = b[i] - t[i]; rhs2 += b[i] * b[i]; }
	if (rhs2 == 0.0) { for (int i = 0; i < N; i++) x[i] = 0.0; iters = 0; err = 0.0; }
	else {
		const double tol = DBL_EPSILON; double thr = tol * tol * rhs2; if (thr < DBL_MIN) thr = DBL_MIN;
		for (int i = 0; i < N; i++) res2 += r[i] * r[i];
		if (res2 >= thr) {
			double absNew = 0.0;
			for (int i = 0; i < N; i++) { p[i] = (diag[i] != 0.0 ? 1.0 / diag[i] : 1.0) * r[i]; absNew += r[i] * p[i]; }
			const int maxIt = 2 * N;
			while (iters < maxIt) {
				SPMV_T(t, p);
				double pt = 0.0; for (int i = 0; i < N; i++) pt += p[i] * t[i];
				const double alpha = absNew / pt;
				res2 = 0.0;
				for (int i = 0; i < N; i++) { x[i] += alpha * p[i]; r[i] -= alpha * t[i]; res2 += r[i] * r[i]; }
				if (res2 < thr) break;
				const double absOld = absNew; absNew = 0.0;
				for (int i = 0; i < N; i++) { z[i] = (diag[i] != 0.0 ? 1.0 / diag[i] : 1.0) * r[i]; absNew += r[i] * z[i]; }
				const double beta = absNew / absOld;
				for (int i = 0; i < N; i++) p[i] = z[i] + beta * p[i];
				iters++;
			}
		}
		err = sqrt(res2 / rhs2);
	}
	}
#undef SPMV_T
	for (int i = 0; i < N; i++) { A->press[i] = x[i]; if (A->press[i] < 0) A->press[i] = 0.0; }
	free(cnt); free(ti); free(tj); free(tv); free(diag); free(fill); free(ci); free(cv); free(r); free(p); free(z); free(t);
	if (errOut) *errOut = err;
	return iters;
}

/* ------------------------------------------------------------------------------------------------ particle walls */
void oracle_extrapolate_press(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B) {
	PARFOR
	for (int i = 0; i < A->numParticles; i++) {
		if (!(A->particleType[i] == P->dummyWall || (P->mpsType == 1 && A->particleType[i] == P->wall))) continue;
		const double* xi = A->pos + 3 * (size_t)i; const double* mi = xi; double ni = 0.0, pr = 0.0;
		NEIGH_LOOP(P, A, B, xi, mi, {
			if (d2 < P->reS2 && j != i && A->particleType[j] == P->fluid) {
				const double w = wgt(sqrt(d2), P->reS, P->weightType); ni += w;
				pr += (A->press[j] - A->RHO[j] * (P->gravityX * rx + P->gravityY * ry + P->gravityZ * rz)) * w;
			}
		});
		if (pr < 0.0) continue;
		A->press[i] = ni > 0 ? pr / ni : pr;
	}
}

void oracle_update_wall_dummy_vel(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B) {
	PARFOR
	for (int i = 0; i < A->numParticles; i++) {
		if (A->particleType[i] != P->dummyWall) continue;
		const double* xi = A->pos + 3 * (size_t)i; const double* mi = xi; double ni = 0.0, ux = 0.0, uy = 0.0, uz = 0.0;
		NEIGH_LOOP(P, A, B, xi, mi, {
			if (d2 < P->reL2 && j != i && A->particleType[j] == P->fluid) {
				const double w = wgt(sqrt(d2), P->reL, P->weightType); const double* vj = A->vel + 3 * (size_t)j;
				ni += w; ux += vj[0] * w; uy += vj[1] * w; uz += vj[2] * w;
			}
		});
		double* v = A->vel + 3 * (size_t)i;
		if (ni > P->epsilonZero) { v[0] = 2.0 * 0.0 - ux / ni; v[1] = 2.0 * 0.0 - uy / ni; v[2] = 2.0 * 0.0 - uz / ni; }
		else { v[0] = 2.0 * 0.0 - ux; v[1] = 2.0 * 0.0 - uy; v[2] = 2.0 * 0.0 - uz; }
	}
}

/* ------------------------------------------------------------------------------------------------ drivers */
static void pressure_stage(PmpsParams* P, PmpsHostArrays* A, OBuckets* B, int* it, double* err) {
	if (P->mpsType == 0 || P->mpsType == 1) oracle_calc_press_eos(P, A);
	else {
		if (P->mpsType == 3) { oracle_vel_divergence(P, A, B); if (P->wallType == 1) oracle_wall_vel_divergence(P, A, B); }   /* src/MpsPressure.cpp:44-56 */
		const int k = oracle_solve_ppe(P, A, B, err); if (it) *it = k;
	}
}

void oracle_init_step0(PmpsParams* P, PmpsHostArrays* A, OBuckets* B, const OMesh* meshes) {   /* src/main.cpp:121-180 */
	oracle_update_buckets(P, A, B);
	oracle_check_out_domain(P, A);
	if (P->wallType == 1) { oracle_wall_closest_point(P, A, meshes); oracle_wall_npcd(P, A, B); }
	oracle_initial_pnd(P, A, B);
	if (P->wallType == 1) oracle_mean_wall_pnd(P, A, B);
	oracle_mean_pnd(P, A, B);
	oracle_mean_neigh_fluid_pnd(P, A, B);
	if (P->freeSurfType == 2) { oracle_calc_normals(P, A, B); if (P->wallType == 1) oracle_calc_wall_normals(P, A, B); }   /* src/main.cpp:170-176 */
	oracle_update_bc(P, A, B);
	pressure_stage(P, A, B, NULL, NULL);
}

/* ------------------------------------------------------------------------------------------------ non-Newtonian rheology */
void oracle_calc_volume_fraction(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B) {   /* src/MpsViscosity.cpp:104-190 */
	PARFOR
	for (int i = 0; i < A->numParticles; i++) {
		if (A->particleType[i] != P->fluid) continue;
		const double* xi = A->pos + 3 * (size_t)i; double sum1 = 0.0, sum2 = 0.0;
		NEIGH_LOOP(P, A, B, xi, xi, {
			if (d2 < P->reS2) {
				if (j != i && A->particleType[j] == P->fluid) {
					double w = 1.0;
					if (P->fractionMethod != 1) { const double dst = sqrt(d2); w = wgt(dst, P->reS, P->weightType); }
					sum1 += w;
					if (A->PTYPE[j] >= 2) sum2 += w;
				}
			}
		});
		A->Cv[i] = sum1 < P->epsilonZero ? 0.0 : sum2 / sum1;
	}
}

/* src/MpsViscosity.cpp:194-700.  The column maps are filled in index order (one thread), as the reference does with OMP_NUM_THREADS=1;
 * returns CFLvisc (:690). */
double oracle_calc_viscosity_interaction(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B) {
	const double gravityMod = sqrt(P->gravityX * P->gravityX + P->gravityY * P->gravityY + P->gravityZ * P->gravityZ);
	const double Xmin = P->domainMinX, Xmax = P->domainMaxX, Ymin = P->domainMinY;
	double Ymax = Ymin, Zmin = 0.0;
	if (P->dim == 3) { Ymax = P->domainMaxY; Zmin = P->domainMinZ; }
	const double aaL0 = 2.0 * P->partDist;
	const int kx_max = (int)((Xmax - Xmin) / aaL0) + 1, ky_max = (int)((Ymax - Ymin) / aaL0) + 1;
	const int stride = ky_max + 1; const size_t cells = (size_t)(kx_max + 1) * stride;
	double* BL = (double*)malloc(sizeof(double) * cells); double* WL = (double*)malloc(sizeof(double) * cells); double* PS = (double*)malloc(sizeof(double) * cells);
	for (size_t c = 0; c < cells; c++) { BL[c] = P->dim == 2 ? Ymin : Zmin; WL[c] = BL[c]; PS[c] = 0.0; }
	for (int i = 0; i < A->numParticles; i++) {   /* :273-301 */
		if (A->particleType[i] != P->fluid) continue;
		const double* x = A->pos + 3 * (size_t)i;
		const int kx = (int)((x[0] - Xmin) / aaL0) + 1;
		const int ky = P->dim == 2 ? 1 : (int)((x[1] - Ymin) / aaL0) + 1;
		const double h = P->dim == 2 ? x[1] : x[2]; const size_t c = (size_t)kx * stride + ky;
		if (h > BL[c] && A->Cv[i] > 0.5) { BL[c] = h; PS[c] = A->press[i]; }
		if (h > WL[c] && A->PTYPE[i] == 1) WL[c] = h;
	}
	PARFOR
	for (int i = 0; i < A->numParticles; i++) {   /* strain rate, :304-437 */
		if (A->particleType[i] != P->fluid) continue;
		double sum1 = 0.0, sum2 = 0.0, sum3 = 0.0, sum4 = 0.0, sum5 = 0.0, sum6 = 0.0, sum7 = 0.0, sum8 = 0.0, sum9 = 0.0, sum10 = 0.0, sumWs = 0.0;
		const double* xi = A->pos + 3 * (size_t)i; const double* mi = A->mirrorParticlePos + 3 * (size_t)i; const double* vi = A->vel + 3 * (size_t)i;
		double MC[9];
		for (int k = 0; k < 3; k++) { MC[k] = A->correcMatrixRow1[3 * (size_t)i + k]; MC[3 + k] = A->correcMatrixRow2[3 * (size_t)i + k]; MC[6 + k] = A->correcMatrixRow3[3 * (size_t)i + k]; }
		NEIGH_LOOP(P, A, B, xi, mi, {
			if (PRED_P(P, P->reS2)) {
				if (j != i) {
					const double dst = sqrt(d2); const double wS = wgt(dst, P->reS, P->weightType);
					const double ux = A->vel[3 * (size_t)j] - vi[0], uy = A->vel[3 * (size_t)j + 1] - vi[1], uz = A->vel[3 * (size_t)j + 2] - vi[2];
					const double inv = 1.0 / d2;
					double r0 = rx, r1 = ry, r2 = rz;
					if (P->gradientCorrection) {
						r0 = (rx * MC[0] + ry * MC[1] + rz * MC[2]); r1 = (rx * MC[3] + ry * MC[4] + rz * MC[5]); r2 = (rx * MC[6] + ry * MC[7] + rz * MC[8]);
					}
					sum1 += ux * r0 * wS * inv; sum2 += ux * r1 * wS * inv; sum3 += ux * r2 * wS * inv;
					sum4 += uy * r0 * wS * inv; sum5 += uy * r1 * wS * inv; sum6 += uy * r2 * wS * inv;
					sum7 += uz * r0 * wS * inv; sum8 += uz * r1 * wS * inv; sum9 += uz * r2 * wS * inv;
					sum10 += A->press[j] * wS; sumWs += wS;
				}
			}
		});
		const double a = -P->coeffPressGrad;
		const double Uxx = a * sum1, Uxy = a * sum2, Uxz = a * sum3, Uyx = a * sum4, Uyy = a * sum5, Uyz = a * sum6, Uzx = a * sum7, Uzy = a * sum8, Uzz = a * sum9;
		if (sumWs > P->epsilonZero) A->p_smooth[i] = sum10 / sumWs; else A->p_smooth[i] = A->press[i];
		if (A->p_smooth[i] < P->epsilonZero) A->p_smooth[i] = 0.0;
		A->S11[i] = 0.5 * (Uxx + Uxx); A->S12[i] = 0.5 * (Uxy + Uyx); A->S13[i] = 0.5 * (Uxz + Uzx);
		A->S22[i] = 0.5 * (Uyy + Uyy); A->S23[i] = 0.5 * (Uyz + Uzy); A->S33[i] = 0.5 * (Uzz + Uzz);
		A->II[i] = 0.5 * (A->S11[i] * A->S11[i] + 2.0 * A->S12[i] * A->S12[i] + 2.0 * A->S13[i] * A->S13[i] + A->S22[i] * A->S22[i] + 2.0 * A->S23[i] * A->S23[i] + A->S33[i] * A->S33[i]);
		if (A->II[i] < P->epsilonZero || A->II[i] * 0 != 0) A->II[i] = 0.0;
	}
	double mi_max = 0.0;
	for (int i = 0; i < A->numParticles; i++) {   /* :463-640 */
		if (A->particleType[i] != P->fluid) continue;
		const double* x = A->pos + 3 * (size_t)i; const double* v = A->vel + 3 * (size_t)i;
		const double vel2i = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
		if (A->PTYPE[i] == 1) A->MEU[i] = P->KNM_VS1 * P->DNS_FL1;
		else if (A->PTYPE[i] == 2) {
			double phi = (A->Cv[i] - 0.25) * P->PHI_1 / (1.0 - 0.25), phi2 = (A->Cv[i] - 0.25) * P->PHI_2 / (1.0 - 0.25);
			if (A->Cv[i] <= 0.25) { phi = P->epsilonZero; phi2 = P->epsilonZero; }
			A->p_rheo_new[i] = A->p_smooth[i];
			const int kx = (int)((x[0] - Xmin) / aaL0) + 1; int ky; double normal_stress; size_t c;
			if (P->dim == 2) {
				ky = 1; c = (size_t)kx * stride + ky;
				normal_stress = (BL[c] - x[1] + P->partDist * 0.5) * (P->DNS_FL2 - P->DNS_FL1) * gravityMod - vel2i * (P->DNS_FL2 - P->DNS_FL1) * 0.5;
				if (WL[c] < BL[c]) normal_stress = A->p_smooth[i];
				else normal_stress = A->p_smooth[i] - PS[c];
			} else {
				ky = (int)((x[1] - Ymin) / aaL0) + 1; c = (size_t)kx * stride + ky;
				normal_stress = (BL[c] - x[2] + P->partDist * 0.5) * (P->DNS_FL2 - P->DNS_FL1) * gravityMod - vel2i * (P->DNS_FL2 - P->DNS_FL1) * 0.5;
				if (A->p_smooth[i] - (WL[c] - x[2]) * P->DNS_FL1 * gravityMod < P->epsilonZero) A->p_smooth[i] = (WL[c] - x[2]) * P->DNS_FL1 * gravityMod;
				if (P->timeCurrent <= 1.0) normal_stress = (1.0 - P->timeCurrent) * (A->p_smooth[i] - (WL[c] - x[2]) * P->DNS_FL1 * gravityMod) + P->timeCurrent * normal_stress;
			}
			if (normal_stress < 1.0 || A->Cv[i] < 0.5) normal_stress = 1.0;
			A->p_rheo_new[i] = normal_stress;
			const double modE = sqrt(A->II[i]);
			if (WL[c] < BL[c]) A->Inertia[i] = modE * P->DG / sqrt(normal_stress / P->DNS_SDT);
			else A->Inertia[i] = modE * P->DG / sqrt(normal_stress / (P->DNS_FL1 * P->Cd));
			A->VF[i] = P->VF_max - (P->VF_max - P->VF_min) * A->Inertia[i];
			if (A->VF[i] < P->VF_min) A->VF[i] = P->VF_min;
			A->RHO[i] = P->DNS_SDT * A->VF[i] + (1.0 - A->VF[i]) * P->DNS_FL1;
			phi *= (A->VF[i] / P->VF_max);
			const double alpha_1 = 2.0 * sqrt(3.0) * sin(phi) / (3.0 - sin(phi)), beta_1 = 2.0 * sqrt(3.0) * cos(phi) / (3.0 - sin(phi));
			double yield_stress = P->cohes * beta_1 + normal_stress * alpha_1;
			if (yield_stress < P->epsilonZero) yield_stress = 0.0;
			A->MEU_Y[i] = yield_stress / (2.0 * sqrt(modE * modE + P->epsilonZero));
			double meu_0;
			if (WL[c] < BL[c]) meu_0 = 0.5 * (tan(phi2) - tan(phi)) * normal_stress * P->DG / (P->I0 * sqrt(normal_stress / P->DNS_FL2) + modE * P->DG) * modE / (sqrt(modE * modE + P->epsilonZero));
			else meu_0 = 0.5 * (tan(phi2) - tan(phi)) * normal_stress * P->DG / (P->I0 * sqrt(normal_stress / (P->DNS_FL1 * P->Cd)) + modE * P->DG) * modE / (sqrt(modE * modE + P->epsilonZero));
			if (modE <= P->epsilonZero || (meu_0 * 0) != 0) meu_0 = P->MEU0;
			const double visc_max = (meu_0 + yield_stress * P->mm * 0.5);
			A->MEU[i] = A->MEU_Y[i] + meu_0;
			if (A->II[i] <= P->epsilonZero) A->MEU[i] = visc_max;
			if (A->MEU[i] > mi_max) mi_max = A->MEU[i];
		}
		if (A->PTYPE[i] >= 2) {
			if (A->Cv[i] > 0.5) A->RHO[i] = P->DNS_FL2;
			else A->RHO[i] = A->Cv[i] * P->DNS_FL2 + (1.0 - A->Cv[i]) * P->DNS_FL1;
		}
	}
	free(BL); free(WL); free(PS);
	return P->timeStep * mi_max / ((P->DNS_SDT * P->VF_max + (1.0 - P->VF_max) * P->DNS_FL1) * P->partDist * P->partDist);
}

/* calcWallSlipViscosityInteractionVal / calcWallNoSlipViscosityInteractionVal, src/MpsViscosity.cpp:703-1111 / :1112-1526 (polygon walls, 3D):
 * for the fluid particles near a wall the velocity gradient over the MIRROR neighbourhood M(reS) is added to the strain-rate invariant
 * and the smoothed pressure, and their viscosity is evaluated again — with the older regularised form of the model these two functions
 * still carry (exponential yield term, power-law index N).  The bed / water level maps are rebuilt from the same state as in
 * calcViscosityInteractionVal, which always runs right before (src/main.cpp:387-399). */
void oracle_calc_wall_viscosity_interaction(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B) {
	const int slip = P->slipCondition == 1;
	const double gravityMod = sqrt(P->gravityX * P->gravityX + P->gravityY * P->gravityY + P->gravityZ * P->gravityZ);
	const double Xmin = P->domainMinX, Xmax = P->domainMaxX, Ymin = P->domainMinY;
	double Ymax = Ymin, Zmin = 0.0;
	if (P->dim == 3) { Ymax = P->domainMaxY; Zmin = P->domainMinZ; }
	const double aaL0 = 2.0 * P->partDist;
	const int kx_max = (int)((Xmax - Xmin) / aaL0) + 1, ky_max = (int)((Ymax - Ymin) / aaL0) + 1;
	const int stride = ky_max + 1; const size_t cells = (size_t)(kx_max + 1) * stride;
	double* BL = (double*)malloc(sizeof(double) * cells); double* WL = (double*)malloc(sizeof(double) * cells);
	for (size_t c = 0; c < cells; c++) { BL[c] = Zmin; WL[c] = Zmin; }
	for (int i = 0; i < A->numParticles; i++) {   /* :1165-1176 */
		if (A->particleType[i] != P->fluid) continue;
		const double* x = A->pos + 3 * (size_t)i;
		const size_t c = (size_t)((int)((x[0] - Xmin) / aaL0) + 1) * stride + ((int)((x[1] - Ymin) / aaL0) + 1);
		if (x[2] > BL[c] && A->Cv[i] > 0.5) BL[c] = x[2];
		if (x[2] > WL[c] && A->PTYPE[i] == 1) WL[c] = x[2];
	}
	PARFOR
	for (int i = 0; i < A->numParticles; i++) {   /* :1179-1345 */
		if (!(A->particleType[i] == P->fluid && A->particleNearWall[i])) continue;
		double sum1 = 0.0, sum2 = 0.0, sum3 = 0.0, sum4 = 0.0, sum5 = 0.0, sum6 = 0.0, sum7 = 0.0, sum8 = 0.0, sum9 = 0.0, sum10 = 0.0;
		const double* xi = A->pos + 3 * (size_t)i; const double* mi = A->mirrorParticlePos + 3 * (size_t)i; const double* vi = A->vel + 3 * (size_t)i;
		double n[3], R[9]; wall_normal(xi, mi, P->epsilonZero, 1, n); reflection(n, R);
		double vm[3];
		if (slip) { for (int k = 0; k < 3; k++) vm[k] = MATVEC(R, vi[0], vi[1], vi[2], k); }   /* v_i' = R v_i, :800-803 */
		else {   /* v_i' = -(v_i - 2 (v_wall - (n.v_wall) n)), :1214-1240 */
			double vw[3] = { 0.0, 0.0, 0.0 };
			if (A->nearMeshType[i] == 2) { vw[0] = P->uniformVelWall[0]; vw[1] = P->uniformVelWall[1]; vw[2] = P->uniformVelWall[2]; }
			const double dotnv = n[0] * vw[0] + n[1] * vw[1] + n[2] * vw[2];
			double vt[3]; for (int k = 0; k < 3; k++) vt[k] = vi[k] - 2.0 * (vw[k] - dotnv * n[k]);
			vm[0] = (-1.0 * vt[0] + 0.0 * vt[1] + 0.0 * vt[2]); vm[1] = (0.0 * vt[0] + -1.0 * vt[1] + 0.0 * vt[2]); vm[2] = (0.0 * vt[0] + 0.0 * vt[1] + -1.0 * vt[2]);
		}
		NEIGH_LOOP(P, A, B, xi, mi, {
			if (PRED_M(P->reS2)) {
				if (j != i) {
					const double dst = sqrt(dm2); const double wS = wgt(dst, P->reS, P->weightType);
					const double ux = A->vel[3 * (size_t)j] - vm[0], uy = A->vel[3 * (size_t)j + 1] - vm[1], uz = A->vel[3 * (size_t)j + 2] - vm[2];
					sum1 += ux * mx * wS / dm2; sum2 += ux * my * wS / dm2; sum3 += ux * mz * wS / dm2;
					sum4 += uy * mx * wS / dm2; sum5 += uy * my * wS / dm2; sum6 += uy * mz * wS / dm2;
					sum7 += uz * mx * wS / dm2; sum8 += uz * my * wS / dm2; sum9 += uz * mz * wS / dm2;
					sum10 += A->press[j] * wS;
				}
			}
		});
		double Uxx, Uxy, Uxz, Uyx, Uyy, Uyz, Uzx, Uzy, Uzz; const double cg = P->coeffPressGrad;
		if (slip) {   /* R (grad u) R, :868-890 */
			const double aUxx = -cg * (R[0] * sum1 + R[1] * sum4 + R[2] * sum7), aUxy = -cg * (R[0] * sum2 + R[1] * sum5 + R[2] * sum8), aUxz = -cg * (R[0] * sum3 + R[1] * sum6 + R[2] * sum9);
			const double aUyx = -cg * (R[3] * sum1 + R[4] * sum4 + R[5] * sum7), aUyy = -cg * (R[3] * sum2 + R[4] * sum5 + R[5] * sum8), aUyz = -cg * (R[3] * sum3 + R[4] * sum6 + R[5] * sum9);
			const double aUzx = -cg * (R[6] * sum1 + R[7] * sum4 + R[8] * sum7), aUzy = -cg * (R[6] * sum2 + R[7] * sum5 + R[8] * sum8), aUzz = -cg * (R[6] * sum3 + R[7] * sum6 + R[8] * sum9);
			Uxx = aUxx * R[0] + aUxy * R[3] + aUxz * R[6]; Uxy = aUxx * R[1] + aUxy * R[4] + aUxz * R[7]; Uxz = aUxx * R[2] + aUxy * R[5] + aUxz * R[8];
			Uyx = aUyx * R[0] + aUyy * R[3] + aUyz * R[6]; Uyy = aUyx * R[1] + aUyy * R[4] + aUyz * R[7]; Uyz = aUyx * R[2] + aUyy * R[5] + aUyz * R[8];
			Uzx = aUzx * R[0] + aUzy * R[3] + aUzz * R[6]; Uzy = aUzx * R[1] + aUzy * R[4] + aUzz * R[7]; Uzz = aUzx * R[2] + aUzy * R[5] + aUzz * R[8];
		} else {      /* -(grad u) R, :1289-1300 */
			Uxx = cg * (sum1 * R[0] + sum2 * R[3] + sum3 * R[6]); Uxy = cg * (sum1 * R[1] + sum2 * R[4] + sum3 * R[7]); Uxz = cg * (sum1 * R[2] + sum2 * R[5] + sum3 * R[8]);
			Uyx = cg * (sum4 * R[0] + sum5 * R[3] + sum6 * R[6]); Uyy = cg * (sum4 * R[1] + sum5 * R[4] + sum6 * R[7]); Uyz = cg * (sum4 * R[2] + sum5 * R[5] + sum6 * R[8]);
			Uzx = cg * (sum7 * R[0] + sum8 * R[3] + sum9 * R[6]); Uzy = cg * (sum7 * R[1] + sum8 * R[4] + sum9 * R[7]); Uzz = cg * (sum7 * R[2] + sum8 * R[5] + sum9 * R[8]);
		}
		A->p_smooth[i] += sum10 / P->pndGradientZero;
		if (A->p_smooth[i] < P->epsilonZero) A->p_smooth[i] = 0.0;
		A->S11[i] = 0.5 * (Uxx + Uxx); A->S12[i] = 0.5 * (Uxy + Uyx); A->S13[i] = 0.5 * (Uxz + Uzx);
		A->S22[i] = 0.5 * (Uyy + Uyy); A->S23[i] = 0.5 * (Uyz + Uzy); A->S33[i] = 0.5 * (Uzz + Uzz);
		A->II[i] += 0.5 * (A->S11[i] * A->S11[i] + 2.0 * A->S12[i] * A->S12[i] + 2.0 * A->S13[i] * A->S13[i] + A->S22[i] * A->S22[i] + 2.0 * A->S23[i] * A->S23[i] + A->S33[i] * A->S33[i]);
		if (A->II[i] < P->epsilonZero || A->II[i] * 0 != 0) A->II[i] = 0.0;
	}
	for (int i = 0; i < A->numParticles; i++) {   /* :1360-1515 */
		if (!(A->particleType[i] == P->fluid && A->particleNearWall[i])) continue;
		const double* x = A->pos + 3 * (size_t)i; const double* v = A->vel + 3 * (size_t)i;
		const double vel2i = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
		if (A->PTYPE[i] == 1) A->MEU[i] = P->KNM_VS1 * P->DNS_FL1;
		else if (A->PTYPE[i] == 2) {
			double phi = (A->Cv[i] - 0.25) * P->PHI_1 / (1.0 - 0.25), phi2 = (A->Cv[i] - 0.25) * P->PHI_2 / (1.0 - 0.25);
			if (A->Cv[i] <= 0.25) { phi = P->epsilonZero; phi2 = P->epsilonZero; }
			const size_t c = (size_t)((int)((x[0] - Xmin) / aaL0) + 1) * stride + ((int)((x[1] - Ymin) / aaL0) + 1);
			double normal_stress = (BL[c] - x[2] + P->partDist * 0.5) * (P->DNS_FL2 - P->DNS_FL1) * gravityMod - vel2i * (P->DNS_FL2 - P->DNS_FL1) * 0.5;
			if (A->p_smooth[i] < (WL[c] - x[2]) * P->DNS_FL1 * gravityMod) A->p_smooth[i] = (WL[c] - x[2]) * P->DNS_FL1 * gravityMod;
			if (P->timeCurrent <= 1.0) normal_stress = (1.0 - P->timeCurrent) * (A->p_smooth[i] - (WL[c] - x[2]) * P->DNS_FL1 * gravityMod) + P->timeCurrent * normal_stress;
			if (normal_stress < 1.0 || A->Cv[i] < 0.5) normal_stress = 1.0;
			A->p_rheo_new[i] = normal_stress;
			A->Inertia[i] = sqrt(A->II[i]) * P->DG / sqrt(normal_stress / (P->DNS_FL1 * P->Cd));
			A->VF[i] = P->VF_max - (P->VF_max - P->VF_min) * A->Inertia[i];
			if (A->VF[i] < P->VF_min) A->VF[i] = P->VF_min;
			A->RHO[i] = P->DNS_SDT * A->VF[i] + (1.0 - A->VF[i]) * P->DNS_FL1;
			phi = phi * A->VF[i] / P->VF_max;
			double yield_stress = P->cohes * cos(phi) + normal_stress * sin(phi);
			if (yield_stress < (slip ? P->epsilonZero : 0.0)) yield_stress = 0.0;   /* :1083 against :1497 */
			double visc_max = (yield_stress * P->mm * 0.5 + P->MEU0);
			if (A->II[i] > P->epsilonZero) A->MEU_Y[i] = yield_stress * (1.0 - exp(-P->mm * sqrt(A->II[i]))) * 0.5 / sqrt(A->II[i]);
			else A->MEU_Y[i] = visc_max;
			double meu_0 = 0.5 * (tan(phi2) - tan(phi)) * normal_stress * P->DG / (P->I0 * sqrt(normal_stress / (P->DNS_FL1 * P->Cd)) + sqrt(A->II[i]) * P->DG);
			if (A->II[i] <= P->epsilonZero || (meu_0 * 0) != 0) meu_0 = P->MEU0;
			visc_max = (yield_stress * P->mm * 0.5 + meu_0);
			A->MEU[i] = A->MEU_Y[i] + P->MEU0 * pow(4.0 * A->II[i], (P->powerLawN - 1.0) * 0.5);
			if (A->II[i] == 0 || A->MEU[i] > visc_max) A->MEU[i] = visc_max;
		}
		if (A->PTYPE[i] >= 2) {
			if (A->Cv[i] > 0.5) A->RHO[i] = P->DNS_FL2;
			else A->RHO[i] = A->Cv[i] * P->DNS_FL2 + (1.0 - A->Cv[i]) * P->DNS_FL1;
		}
	}
	free(BL); free(WL);
}

void oracle_step(PmpsParams* P, PmpsHostArrays* A, OBuckets* B, const OMesh* meshes, int* solverIter, double* solverErr) {
	const int poly = P->wallType == 1;   /* order: src/main.cpp:256-456 */
	oracle_update_buckets(P, A, B);
	if (P->fluidType == 1) { oracle_calc_volume_fraction(P, A, B); oracle_calc_viscosity_interaction(P, A, B); }   /* src/main.cpp:259-262 */
	oracle_calc_viscosity(P, A, B);
	oracle_add_gravity(P, A);
	if (P->relaxPress < 1.0) { oracle_prediction_press_gradient(P, A, B); if (poly) oracle_prediction_wall_press_gradient(P, A, B, P->velMax); }   /* src/main.cpp:268-276 */
	oracle_predict(P, A);
	oracle_check_out_domain(P, A);
	oracle_collisions(P, A, B);
	if (poly) { oracle_wall_closest_point(P, A, meshes); oracle_wall_npcd(P, A, B); }
	if (P->gradientCorrection) oracle_correction_matrix(P, A, B);
	oracle_pnd_neigh_npcd(P, A, B);
	if (P->pndType == 2) {
		oracle_pnd_diffusive(P, A, B);
		if (poly) oracle_wall_pnd_diffusive(P, A, B);
		oracle_update_pnd(P, A);
		oracle_mean_pnd_wall_dummy_surface(P, A, B);
	} else if (P->pndType == 1) {
		if (poly) oracle_mean_wall_pnd(P, A, B);
		oracle_mean_pnd(P, A, B);
	}
	if (P->freeSurfType == 2) { oracle_calc_normals(P, A, B); if (poly) oracle_calc_wall_normals(P, A, B); }   /* src/main.cpp:351-357 */
	oracle_update_bc(P, A, B);
	pressure_stage(P, A, B, solverIter, solverErr);
	if (!poly) oracle_extrapolate_press(P, A, B);
	if (P->gradientCorrection) oracle_correction_matrix(P, A, B);
	oracle_press_gradient(P, A, B);
	if (poly) oracle_wall_press_gradient(P, A, B, P->velMax);
	if (P->fluidType == 1) { oracle_calc_volume_fraction(P, A, B); oracle_calc_viscosity_interaction(P, A, B); if (poly) oracle_calc_wall_viscosity_interaction(P, A, B); }   /* src/main.cpp:387-399 */
	if (poly) oracle_wall_viscosity(P, A, B);
	P->velMax = oracle_correct(P, A);
	oracle_check_out_domain(P, A);
	if (P->shiftingType == 1) oracle_shifting(P, A, B);
	if (P->shiftingType == 2) { oracle_shifting(P, A, B); if (poly) oracle_wall_shifting(P, A, B); }
	if (!poly) oracle_update_wall_dummy_vel(P, A, B);
	oracle_check_out_domain(P, A);
	P->timeCurrent += P->timeStep;   /* src/main.cpp:460 */
}
