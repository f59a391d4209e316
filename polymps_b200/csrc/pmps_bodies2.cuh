// pmps_bodies2.cuh — stage bodies, continued: particle number density, pressure gradient, correction matrix.
#pragma once

// ------------------------------------------------------------------------------------------------ PND family
// MpsPndNeigh::calcWallNPCD, src/MpsPndNeigh.cpp:105-195
struct BWallNpcd {
	typedef void wall_frame_tag;
	struct Ctx { double sx, sy, sz; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		if (!s.nearWallFluid) return false;
		c.sx = c.sy = c.sz = 0.0; return true;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const Nb& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_M(q, K.p.reS2)) return;
		const double w = wgt_r(fast_rsqrt(q.dm2), q.dm2, K.p.reS, K.invReS, K.p.weightType);
		c.sx += q.mx * w; c.sy += q.my * w; c.sz += q.mz * w;
	}
	__device__ __forceinline__ static void reduce(Ctx& c, unsigned m) { c.sx = gsum(c.sx, m); c.sy = gsum(c.sy, m); c.sz = gsum(c.sz, m); }
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		const double ex = s.pi.x - s.mix, ey = s.pi.y - s.miy, ez = s.pi.z - s.miz; const double dd = sq3_exact(ex, ey, ez);
		if (dd < K.p.reS2) { const double w = wgt(sqrt(dd), K.p.reS, K.p.weightType); c.sx += ex * w; c.sy += ey * w; c.sz += ez * w; }
		WallFrame f; wall_frame_of(f, s, K.p.epsilonZero, false);
		S.npcdDeviation.x[s.i] += MATVEC(f.R, c.sx, c.sy, c.sz, 0);
		S.npcdDeviation.y[s.i] += MATVEC(f.R, c.sx, c.sy, c.sz, 1);
		S.npcdDeviation.z[s.i] += MATVEC(f.R, c.sx, c.sy, c.sz, 2);
	}
};

// shared tail of calcPndnNeighNPCD / setInitialPndNumberOfNeighNPCD (src/MpsPndNeigh.cpp:287-305, :81-100)
__device__ __forceinline__ void npcd_finish(const DevView& S, const KParams& K, int i, double dx, double dy, double dz, double wsum, double pndSmall) {
	if (wsum > K.p.epsilonZero) { dx /= pndSmall; dy /= pndSmall; dz /= pndSmall; }
	S.npcdDeviation.x[i] = dx; S.npcdDeviation.y[i] = dy; S.npcdDeviation.z[i] = dz;
	S.npcdDeviation2[i] = dx * dx + dy * dy + dz * dz;
	S.deviationDotPolygonNormal[i] = (dx * S.polygonNormal.x[i] + dy * S.polygonNormal.y[i] + dz * S.polygonNormal.z[i] < 0.0) ? 1 : -1;
}

// MpsPndNeigh::calcPndnNeighNPCD, src/MpsPndNeigh.cpp:199-331.  The wall part of the deviation vector (calcWallNPCD,
// which runs before it — as a separate stage or as the body in front of this one in the fused pass) is added in end().
struct BPndNeighNpcd {
	struct Ctx { double ni, wsum, dx, dy, dz; int cnt; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		c.ni = c.wsum = 0.0; c.cnt = 0; c.dx = c.dy = c.dz = 0.0; return true;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const Nb& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_P<POLY>(q, K.p.reL2)) return;
		c.cnt += 1;
		if (q.d2 < K.p.reS2) {
			const double w = wgt_r(q.rinv, q.d2, K.p.reS, K.invReS, K.p.weightType);
			c.ni += w; c.dx += q.rx * w; c.dy += q.ry * w; c.dz += q.rz * w; c.wsum += w;
		}
	}
	__device__ __forceinline__ static void reduce(Ctx& c, unsigned m) { c.ni = gsum(c.ni, m); c.wsum = gsum(c.wsum, m); c.dx = gsum(c.dx, m); c.dy = gsum(c.dy, m); c.dz = gsum(c.dz, m); c.cnt = gsum(c.cnt, m); }
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		S.numNeigh[s.i] = c.cnt + S.numNeighWallContribution[s.i];
		const double ps = c.ni + S.pndWallContribution[s.i];
		if (K.p.pndType == 0 || K.p.pndType == 1) { S.pndki[s.i] = S.pndi[s.i]; S.pndi[s.i] = ps; }
		S.pndSmall[s.i] = ps;
		npcd_finish(S, K, s.i, S.npcdDeviation.x[s.i] + c.dx, S.npcdDeviation.y[s.i] + c.dy, S.npcdDeviation.z[s.i] + c.dz, c.wsum, ps);
	}
};

// MpsPndNeigh::setInitialPndNumberOfNeighNPCD, src/MpsPndNeigh.cpp:21-101 (0-th step)
struct BInitialPnd {
	struct Ctx { double acc, wsum, dx, dy, dz; int cnt; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		c.acc = 0.0; c.wsum = 0.0; c.cnt = 0; c.dx = c.dy = c.dz = 0.0; return true;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const Nb& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_P<POLY>(q, K.p.reL2)) return;
		c.cnt += 1;
		if (q.d2 < K.p.reS2) {
			const double w = wgt_r(q.rinv, q.d2, K.p.reS, K.invReS, K.p.weightType);
			c.acc += w; c.dx += q.rx * w; c.dy += q.ry * w; c.dz += q.rz * w; c.wsum += w;
		}
	}
	__device__ __forceinline__ static void reduce(Ctx& c, unsigned m) { c.acc = gsum(c.acc, m); c.wsum = gsum(c.wsum, m); c.dx = gsum(c.dx, m); c.dy = gsum(c.dy, m); c.dz = gsum(c.dz, m); c.cnt = gsum(c.cnt, m); }
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		double a = S.pndi[s.i] + c.acc + S.pndWallContribution[s.i];
		if (s.type == K.p.wall) a = K.p.pndSmallZero;
		S.pndi[s.i] = a; S.pndSmall[s.i] = a; S.pndki[s.i] = a;
		S.numNeigh[s.i] = c.cnt + S.numNeighWallContribution[s.i];
		npcd_finish(S, K, s.i, c.dx, c.dy, c.dz, c.wsum, a);
	}
};

// MpsPndNeigh::calcPndDiffusiveTerm, src/MpsPndNeigh.cpp:334-489
struct BPndDiffusive {
	struct Ctx { double ni, invNi, vx, vy, vz, Pi, rho, kDi, Di, div; double c1[3], c2[3], c3[3]; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		c.ni = OWN(pndi, S.pndi[s.i]);
		if (c.ni < K.p.epsilonZero) return false;
		c.invNi = 1.0 / c.ni;
		if (TV_OWN_STAGED && s.nb) { c.vx = s.nb->vx[s.ms]; c.vy = s.nb->vy[s.ms]; c.vz = s.nb->vz[s.ms]; }
		else { const double4 v = S.vel[s.i]; c.vx = v.x; c.vy = v.y; c.vz = v.z; }
		c.Pi = OWN(press, S.press[s.i]); c.rho = OWN(RHO, S.RHO[s.i]);
		c.kDi = K.p.timeStep / c.rho * K.p.coeffViscMultiphase;
		if (K.p.gradientCorrection) {
			c.c1[0] = S.correcMatrixRow1.x[s.i]; c.c1[1] = S.correcMatrixRow1.y[s.i]; c.c1[2] = S.correcMatrixRow1.z[s.i];
			c.c2[0] = S.correcMatrixRow2.x[s.i]; c.c2[1] = S.correcMatrixRow2.y[s.i]; c.c2[2] = S.correcMatrixRow2.z[s.i];
			c.c3[0] = S.correcMatrixRow3.x[s.i]; c.c3[1] = S.correcMatrixRow3.y[s.i]; c.c3[2] = S.correcMatrixRow3.z[s.i];
		}
		c.Di = c.div = 0.0; return true;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const Nb& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_P<POLY>(q, K.p.reL2)) return;
		if (s.type == K.p.fluid && q.tj == K.p.fluid) {
			const double wL = wgt_r(q.rinv, q.d2, K.p.reL, K.invReL, K.p.weightType);
			const double rhoj = S.RHO[q.j];
			const double pgh = -2.0 * (c.rho * rhoj * fast_rcp(c.rho + rhoj)) * (K.p.gravityX * q.rx + K.p.gravityY * q.ry + K.p.gravityZ * q.rz);
			c.Di += c.kDi * (S.press[q.j] - c.Pi - pgh) * wL;
		}
		if (q.d2 < K.p.reS2) {
			const double vjx = S.vx[q.j], vjy = S.vy[q.j], vjz = S.vz[q.j];
			const double ux = vjx - c.vx, uy = vjy - c.vy, uz = vjz - c.vz;
			const double wS = wgt_r(q.rinv, q.d2, K.p.reS, K.invReS, K.p.weightType) * (q.rinv * q.rinv);   // w / d2
			if (!K.p.gradientCorrection) c.div += S.pndi[q.j] * (ux * q.rx + uy * q.ry + uz * q.rz) * wS;   // * (dim/n0S)/ni in end()
			else {
				const double qx = (q.rx * c.c1[0] + q.ry * c.c1[1] + q.rz * c.c1[2]), qy = (q.rx * c.c2[0] + q.ry * c.c2[1] + q.rz * c.c2[2]), qz = (q.rx * c.c3[0] + q.ry * c.c3[1] + q.rz * c.c3[2]);
				c.div += (ux * qx + uy * qy + uz * qz) * wS;
			}
		}
	}
	__device__ __forceinline__ static void reduce(Ctx& c, unsigned m) { c.Di = gsum(c.Di, m); c.div = gsum(c.div, m); }
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		double div = c.div * ((double)K.p.dim / K.p.pndSmallZero);
		if (!K.p.gradientCorrection) div *= c.invNi;
		S.pndAuxVar1[s.i] = c.ni * (1.0 + K.p.timeStep * (-div + c.Di * 1.0));
		S.velDivergence[s.i] = div; S.diffusiveTerm[s.i] = c.Di;
	}
};

// MpsPndNeigh::calcWallNoSlipPndDiffusiveTerm / calcWallSlipPndDiffusiveTerm, src/MpsPndNeigh.cpp:492-744
// DIV = true: calcWallNoSlipVelDivergence / calcWallSlipVelDivergence (src/MpsPressure.cpp:554-788): the same sum, added to velDivergence
template <bool DIV> struct BWallPndDiffusiveT {
	typedef void wall_frame_tag;
	struct Ctx { double ni, v[3], vm[3], div; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		c.ni = S.pndi[s.i];
		if (!(s.nearWallFluid && c.ni > K.p.epsilonZero)) return false;
		const double4 v = S.vel[s.i]; c.v[0] = v.x; c.v[1] = v.y; c.v[2] = v.z;
		WallFrame f; wall_frame_of(f, s, K.p.epsilonZero, true);
		mirror_velocity(K, S.nearMeshType[s.i], f, c.v, c.vm);
		c.div = 0.0; return true;
	}
	__device__ __forceinline__ static double term(const Ctx& c, const KParams& K, const Self& s, double vjx, double vjy, double vjz, double mx, double my, double mz, double pndj, double dd) {
		const double wS = wgt(sqrt(dd), K.p.reS, K.p.weightType);
		WallFrame f; wall_frame_of(f, s, K.p.epsilonZero, true);
		double ux = -(vjx - c.vm[0]), uy = -(vjy - c.vm[1]), uz = -(vjz - c.vm[2]);
		double qx = MATVEC(f.R, mx, my, mz, 0), qy = MATVEC(f.R, mx, my, mz, 1), qz = MATVEC(f.R, mx, my, mz, 2);
		if (K.p.slipCondition == 1) { ux = vjx - c.vm[0]; uy = vjy - c.vm[1]; uz = vjz - c.vm[2]; qx = mx; qy = my; qz = mz; }
		return ((double)K.p.dim / K.p.pndSmallZero) * (pndj / c.ni) * (ux * qx + uy * qy + uz * qz) * wS / dd;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const Nb& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_M(q, K.p.reS2)) return;
		const double vjx = S.vx[q.j], vjy = S.vy[q.j], vjz = S.vz[q.j];
		c.div += term(c, K, s, vjx, vjy, vjz, q.mx, q.my, q.mz, S.pndi[q.j], q.dm2);
	}
	__device__ __forceinline__ static void reduce(Ctx& c, unsigned m) { c.div = gsum(c.div, m); }
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		const double ex = s.pi.x - s.mix, ey = s.pi.y - s.miy, ez = s.pi.z - s.miz; const double dd = sq3_exact(ex, ey, ez);
		if (dd < K.p.reS2) c.div += term(c, K, s, c.v[0], c.v[1], c.v[2], ex, ey, ez, c.ni, dd);
		if (DIV) S.velDivergence[s.i] += c.div;
		else S.pndAuxVar1[s.i] += -c.ni * K.p.timeStep * c.div;
	}
};
typedef BWallPndDiffusiveT<false> BWallPndDiffusive;
typedef BWallPndDiffusiveT<true> BWallVelDivergence;

// MpsPressure::calcVelDivergence, src/MpsPressure.cpp:479-551 (IMPLICIT_PND_DIVU): divergence of the velocity over P(reS), fluid particles
struct BVelDivergence {
	struct Ctx { double ni, vx, vy, vz, div; double m[9]; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		if (s.type != K.p.fluid) { if (s.lead) S.velDivergence[s.i] = 0.0; return false; }
		c.ni = S.pndi[s.i];
		const double4 v = S.vel[s.i]; c.vx = v.x; c.vy = v.y; c.vz = v.z; c.div = 0.0;
		if (K.p.gradientCorrection) {
			c.m[0] = S.correcMatrixRow1.x[s.i]; c.m[1] = S.correcMatrixRow1.y[s.i]; c.m[2] = S.correcMatrixRow1.z[s.i];
			c.m[3] = S.correcMatrixRow2.x[s.i]; c.m[4] = S.correcMatrixRow2.y[s.i]; c.m[5] = S.correcMatrixRow2.z[s.i];
			c.m[6] = S.correcMatrixRow3.x[s.i]; c.m[7] = S.correcMatrixRow3.y[s.i]; c.m[8] = S.correcMatrixRow3.z[s.i];
		}
		return true;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const Nb& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_P<POLY>(q, K.p.reS2)) return;
		const double ux = S.vx[q.j] - c.vx, uy = S.vy[q.j] - c.vy, uz = S.vz[q.j] - c.vz;
		const double wS = wgt_r(q.rinv, q.d2, K.p.reS, K.invReS, K.p.weightType) * (q.rinv * q.rinv);
		if (!K.p.gradientCorrection) { if (c.ni > K.p.epsilonZero) c.div += S.pndi[q.j] * (ux * q.rx + uy * q.ry + uz * q.rz) * wS; }
		else {
			const double qx = q.rx * c.m[0] + q.ry * c.m[1] + q.rz * c.m[2], qy = q.rx * c.m[3] + q.ry * c.m[4] + q.rz * c.m[5], qz = q.rx * c.m[6] + q.ry * c.m[7] + q.rz * c.m[8];
			c.div += (ux * qx + uy * qy + uz * qz) * wS;
		}
	}
	__device__ __forceinline__ static void reduce(Ctx& c, unsigned m) { c.div = gsum(c.div, m); }
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		double div = c.div * ((double)K.p.dim / K.p.pndSmallZero);
		if (!K.p.gradientCorrection) div = c.ni > K.p.epsilonZero ? div / c.ni : 0.0;
		S.velDivergence[s.i] = div;
	}
};

// sum_j pndi_j w / sum_j w over P(reS): first half of MpsPndNeigh::meanPndParticlesWallDummySurface (src/MpsPndNeigh.cpp:777-852)
// and of meanPnd (:855-920).  SEL 0: wall or surface particles, overwrite; SEL 1: every particle, accumulate (meanPnd)
template <int SEL> struct BMeanPnd {
	typedef void sparse_tag;
	struct Ctx { double up, dn; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		c.up = c.dn = 0.0;
		if (SEL == 0) {
			S.pndAuxVar1[s.i] = 0.0; S.pndAuxVar2[s.i] = 0.0;
			return s.type == K.p.wall || S.particleBC[s.i] == K.p.surface;
		}
		return true;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const Nb& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_P<POLY>(q, K.p.reS2)) return;
		const double w = wgt_r(q.rinv, q.d2, K.p.reS, K.invReS, K.p.weightType); c.up += S.pndi[q.j] * w; c.dn += w;
	}
	__device__ __forceinline__ static void reduce(Ctx& c, unsigned m) { c.up = gsum(c.up, m); c.dn = gsum(c.dn, m); }
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		if (SEL == 0) { S.pndAuxVar1[s.i] = c.up; S.pndAuxVar2[s.i] = c.dn; }
		else { S.pndAuxVar1[s.i] += c.up; S.pndAuxVar2[s.i] += c.dn; }
	}
};

// MpsPndNeigh::meanWallPnd, src/MpsPndNeigh.cpp:923-994
struct BMeanWallPnd {
	struct Ctx { double up, dn; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		c.up = c.dn = 0.0; S.pndAuxVar1[s.i] = 0.0; S.pndAuxVar2[s.i] = 0.0;
		return S.particleNearWall[s.i] != 0;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const Nb& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_M(q, K.p.reS2)) return;
		const double w = wgt(sqrt(q.dm2), K.p.reS, K.p.weightType); c.up += S.pndi[q.j] * w; c.dn += w;
	}
	__device__ __forceinline__ static void reduce(Ctx& c, unsigned m) { c.up = gsum(c.up, m); c.dn = gsum(c.dn, m); }
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		const double ex = s.pi.x - s.mix, ey = s.pi.y - s.miy, ez = s.pi.z - s.miz; const double dd = sq3_exact(ex, ey, ez);
		if (dd < K.p.reS2) { const double w = wgt(sqrt(dd), K.p.reS, K.p.weightType); c.up += S.pndi[s.i] * w; c.dn += w; }
		S.pndAuxVar1[s.i] = c.up; S.pndAuxVar2[s.i] = c.dn;
	}
};

// MpsPndNeigh::meanNeighFluidPnd, src/MpsPndNeigh.cpp:997-1049
struct BMeanNeighFluidPnd {
	struct Ctx { double up, dn; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) { c.up = c.dn = 0.0; return true; }
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const Nb& S, const KParams& K, const Self& s, const Pair& q) {
		if (!(pred_P<POLY>(q, K.p.reS2) && q.tj == K.p.fluid)) return;
		const double w = wgt_r(q.rinv, q.d2, K.p.reS, K.invReS, K.p.weightType); c.up += S.pndki[q.j] * w; c.dn += w;
	}
	__device__ __forceinline__ static void reduce(Ctx& c, unsigned m) { c.up = gsum(c.up, m); c.dn = gsum(c.dn, m); }
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		S.pndski[s.i] = c.dn > K.p.epsilonZero ? c.up / c.dn : c.up;
	}
};

// ------------------------------------------------------------------------------------------------ gradient correction
// MpsVectorMatrix::correctionMatrix + inverseMatrix, src/MpsVectorMatrix.cpp:20-166
__device__ __forceinline__ void invert3(double M[9], int dim, double eps) {   // Gauss-Jordan without pivoting
	double a[3][3], inv[3][3];
	for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) { a[r][c] = M[3 * r + c]; inv[r][c] = r == c ? 1.0 : 0.0; }
	if (dim == 2) { a[0][2] = a[1][2] = a[2][0] = a[2][1] = 0.0; a[2][2] = 1.0; }
	const double det = (a[0][0] * a[1][1] * a[2][2] + a[0][1] * a[1][2] * a[2][0] + a[0][2] * a[1][0] * a[2][1])
	                 - (a[0][2] * a[1][1] * a[2][0] + a[0][1] * a[1][0] * a[2][2] + a[0][0] * a[1][2] * a[2][1]);
	if (det <= eps) { for (int k = 0; k < 9; k++) M[k] = (k % 4 == 0) ? 1.0 : 0.0; return; }
	for (int k = 0; k < 3; k++) {
		if (k >= dim) break;
		for (int r = 0; r < 3; r++) {
			if (r >= dim || r == k) continue;
			const double m = a[r][k] / a[k][k];
			for (int c = 0; c < 3; c++) { if (c >= dim) break; a[r][c] -= m * a[k][c]; inv[r][c] -= m * inv[k][c]; }
		}
	}
	for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) if (r < dim && c < dim) inv[r][c] /= a[r][r];
	for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) M[3 * r + c] = inv[r][c];
}

struct BCorrectionMatrix {
	struct Ctx { double M[9]; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) { for (int k = 0; k < 9; k++) c.M[k] = 0.0; return true; }
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const Nb& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_P<POLY>(q, K.p.reS2)) return;
		const double w = wgt_r(q.rinv, q.d2, K.p.reS, K.invReS, K.p.weightType) * (q.rinv * q.rinv); const double r[3] = { q.rx, q.ry, q.rz };
		for (int a = 0; a < 3; a++) for (int b = 0; b < 3; b++) c.M[3 * a + b] += w * r[a] * r[b];
	}
	__device__ __forceinline__ static void reduce(Ctx& c, unsigned m) { for (int k = 0; k < 9; k++) c.M[k] = gsum(c.M[k], m); }
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		for (int k = 0; k < 9; k++) c.M[k] *= -K.p.coeffPressGrad;
		invert3(c.M, K.p.dim, K.p.epsilonZero);
		if (S.numNeigh[s.i] < 4) for (int k = 0; k < 9; k++) c.M[k] = (k % 4 == 0) ? 1.0 : 0.0;
		S.correcMatrixRow1.x[s.i] = c.M[0]; S.correcMatrixRow1.y[s.i] = c.M[1]; S.correcMatrixRow1.z[s.i] = c.M[2];
		S.correcMatrixRow2.x[s.i] = c.M[3]; S.correcMatrixRow2.y[s.i] = c.M[4]; S.correcMatrixRow2.z[s.i] = c.M[5];
		S.correcMatrixRow3.x[s.i] = c.M[6]; S.correcMatrixRow3.y[s.i] = c.M[7]; S.correcMatrixRow3.z[s.i] = c.M[8];
	}
};

#include "pmps_bodies3.cuh"
