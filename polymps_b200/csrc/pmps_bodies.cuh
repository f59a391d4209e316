// pmps_bodies.cuh — one stage body per reference neighbour loop.  A body is the {begin, pair, end} triple that the
// list walker (pmps_device.cuh) runs for slot i; several bodies fuse into one pass when their inputs coincide.
// Each body cites the reference method whose loop body it replaces.
#pragma once
#include "pmps_device.cuh"

#define VLD(v, i) (v).x[i], (v).y[i], (v).z[i]

// ------------------------------------------------------------------------------------------------ viscosity
// MpsViscosity::calcViscosity, src/MpsViscosity.cpp:20-100: acc = coeff * sum (v_j - v_i) w_L nu_ij over P(reL)
struct BViscosity {
	struct Ctx { double vx, vy, vz, mu, rho, sx, sy, sz; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		const double4 v = S.vel[s.i]; c.vx = v.x; c.vy = v.y; c.vz = v.z; c.mu = S.MEU[s.i]; c.rho = S.RHO[s.i];
		c.sx = c.sy = c.sz = 0.0; return true;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const DevView& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_P<POLY>(q, K.p.reL2)) return;
		const double w = wgt(sqrt(q.d2), K.p.reL, K.p.weightType);
		const double muj = S.MEU[q.j];
		double nu = (c.mu + muj) > K.p.epsilonZero ? 2.0 * c.mu * muj / (c.mu + muj) : 0.0;
		if (pk_type(q.pj) == K.p.wall) nu = 2.0 * c.mu;
		nu = nu / c.rho;
		const double4 vj = S.vel[q.j];
		c.sx += (vj.x - c.vx) * w * nu; c.sy += (vj.y - c.vy) * w * nu; c.sz += (vj.z - c.vz) * w * nu;
	}
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		S.acc.x[s.i] = K.p.coeffViscMultiphase * c.sx; S.acc.y[s.i] = K.p.coeffViscMultiphase * c.sy; S.acc.z[s.i] = K.p.coeffViscMultiphase * c.sz;
	}
};

// MpsViscosity::calcWallNoSlipViscosity / calcWallSlipViscosity, src/MpsViscosity.cpp:1527-1910 (near-wall fluid)
struct BWallViscosity {
	struct Ctx { double v[3], vm[3], mu, rhoDiv, sx, sy, sz; WallFrame f; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		if (!(s.type == K.p.fluid && S.particleNearWall[s.i])) return false;
		const double4 v = S.vel[s.i]; c.v[0] = v.x; c.v[1] = v.y; c.v[2] = v.z; c.mu = S.MEU[s.i];
		wall_frame(c.f, s.pi.x, s.pi.y, s.pi.z, s.mix, s.miy, s.miz, K.p.epsilonZero, true);
		mirror_velocity(K, S.nearMeshType[s.i], c.f, c.v, c.vm);
		c.rhoDiv = s.ptype == 1 ? K.p.DNS_FL1 : K.p.DNS_FL2;
		c.sx = c.sy = c.sz = 0.0; return true;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const DevView& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_M(q, K.p.reL2)) return;
		const double w = wgt(sqrt(q.dm2), K.p.reL, K.p.weightType);
		const double muj = S.MEU[q.j];
		double nu = (c.mu + muj) > K.p.epsilonZero ? 2.0 * c.mu * muj / (c.mu + muj) : 0.0;
		nu = nu / c.rhoDiv;
		const double4 vj = S.vel[q.j];
		c.sx += (vj.x - c.vm[0]) * w * nu; c.sy += (vj.y - c.vm[1]) * w * nu; c.sz += (vj.z - c.vm[2]) * w * nu;
	}
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		const double ex = s.pi.x - s.mix, ey = s.pi.y - s.miy, ez = s.pi.z - s.miz; const double dd = sq3_exact(ex, ey, ez);
		if (dd < K.p.reL2) {   // i itself is a neighbour of its mirror image
			const double w = wgt(sqrt(dd), K.p.reL, K.p.weightType);
			double nu = c.mu > K.p.epsilonZero ? 2.0 * c.mu * c.mu / (c.mu + c.mu) : 0.0;
			nu = nu / c.rhoDiv;
			c.sx += (c.v[0] - c.vm[0]) * w * nu; c.sy += (c.v[1] - c.vm[1]) * w * nu; c.sz += (c.v[2] - c.vm[2]) * w * nu;
		}
		double T[9];
		if (K.p.slipCondition == 0) { for (int k = 0; k < 9; k++) T[k] = 0.0; T[0] = T[4] = T[8] = -1.0; }   // R_inv = -I
		else { for (int k = 0; k < 9; k++) T[k] = c.f.R[k]; }
		const double rho = S.RHO[s.i];
		double* a[3] = { S.acc.x, S.acc.y, S.acc.z }; double* fw[3] = { S.forceWall.x, S.forceWall.y, S.forceWall.z };
		for (int k = 0; k < 3; k++) {
			const double t = MATVEC(T, c.sx, c.sy, c.sz, k);
			a[k][s.i] += t * K.p.coeffViscMultiphase;
			fw[k][s.i] += -t * K.p.coeffViscMultiphase * K.vol * rho;
		}
	}
};

// ------------------------------------------------------------------------------------------------ collisions (PC)
// MpsParticleCollision::checkParticleCollisions, src/MpsParticleCollision.cpp:22-119 (first loop: dvel)
struct BCollision {
	struct Ctx { double vx, vy, vz, mi, dx, dy, dz; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		if (s.type != K.p.fluid) return false;
		const double4 v = S.vel[s.i]; c.vx = v.x; c.vy = v.y; c.vz = v.z;
		c.mi = s.ptype == 1 ? K.p.DNS_FL1 : K.p.DNS_FL2; c.dx = c.dy = c.dz = 0.0; return true;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const DevView& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_P<POLY>(q, K.p.distCollisionLimit2)) return;
		const double4 vj = S.vel[q.j];
		double f = (c.vx - vj.x) * q.rx + (c.vy - vj.y) * q.ry + (c.vz - vj.z) * q.rz;
		if (f > 0.0) {
			const bool jf = pk_type(q.pj) == K.p.fluid;
			const double mj = jf ? (pk_ptype(q.pj) == 1 ? K.p.DNS_FL1 : K.p.DNS_FL2) : K.p.densityWall;
			f *= K.p.restitutionCollision * mj / (c.mi + mj) / q.d2;
			if (jf) { c.dx -= q.rx * f; c.dy -= q.ry * f; c.dz -= q.rz * f; }
			else { c.dx -= 2.0 * q.rx * f; c.dy -= 2.0 * q.ry * f; c.dz -= 2.0 * q.rz * f; }
		}
	}
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		S.dvelCollision.x[s.i] = c.dx; S.dvelCollision.y[s.i] = c.dy; S.dvelCollision.z[s.i] = c.dz;
	}
};

// ------------------------------------------------------------------------------------------------ PND family
// MpsPndNeigh::calcWallNPCD, src/MpsPndNeigh.cpp:105-195
struct BWallNpcd {
	struct Ctx { double sx, sy, sz; WallFrame f; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		if (!(s.type == K.p.fluid && S.particleNearWall[s.i])) return false;
		wall_frame(c.f, s.pi.x, s.pi.y, s.pi.z, s.mix, s.miy, s.miz, K.p.epsilonZero, false);
		c.sx = c.sy = c.sz = 0.0; return true;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const DevView& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_M(q, K.p.reS2)) return;
		const double w = wgt(sqrt(q.dm2), K.p.reS, K.p.weightType);
		c.sx += q.mx * w; c.sy += q.my * w; c.sz += q.mz * w;
	}
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		const double ex = s.pi.x - s.mix, ey = s.pi.y - s.miy, ez = s.pi.z - s.miz; const double dd = sq3_exact(ex, ey, ez);
		if (dd < K.p.reS2) { const double w = wgt(sqrt(dd), K.p.reS, K.p.weightType); c.sx += ex * w; c.sy += ey * w; c.sz += ez * w; }
		S.npcdDeviation.x[s.i] += MATVEC(c.f.R, c.sx, c.sy, c.sz, 0);
		S.npcdDeviation.y[s.i] += MATVEC(c.f.R, c.sx, c.sy, c.sz, 1);
		S.npcdDeviation.z[s.i] += MATVEC(c.f.R, c.sx, c.sy, c.sz, 2);
	}
};

// shared tail of calcPndnNeighNPCD / setInitialPndNumberOfNeighNPCD (src/MpsPndNeigh.cpp:287-305, :81-100)
__device__ __forceinline__ void npcd_finish(const DevView& S, const KParams& K, int i, double dx, double dy, double dz, double wsum, double pndSmall) {
	if (wsum > K.p.epsilonZero) { dx /= pndSmall; dy /= pndSmall; dz /= pndSmall; }
	S.npcdDeviation.x[i] = dx; S.npcdDeviation.y[i] = dy; S.npcdDeviation.z[i] = dz;
	S.npcdDeviation2[i] = dx * dx + dy * dy + dz * dz;
	S.deviationDotPolygonNormal[i] = (dx * S.polygonNormal.x[i] + dy * S.polygonNormal.y[i] + dz * S.polygonNormal.z[i] < 0.0) ? 1 : -1;
}

// MpsPndNeigh::calcPndnNeighNPCD, src/MpsPndNeigh.cpp:199-331
struct BPndNeighNpcd {
	struct Ctx { double ni, wsum, dx, dy, dz; int cnt; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		c.ni = c.wsum = 0.0; c.cnt = S.numNeighWallContribution[s.i];
		c.dx = S.npcdDeviation.x[s.i]; c.dy = S.npcdDeviation.y[s.i]; c.dz = S.npcdDeviation.z[s.i];   // on top of the wall part
		return true;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const DevView& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_P<POLY>(q, K.p.reL2)) return;
		c.cnt += 1;
		if (q.d2 < K.p.reS2) {
			const double w = wgt(sqrt(q.d2), K.p.reS, K.p.weightType);
			c.ni += w; c.dx += q.rx * w; c.dy += q.ry * w; c.dz += q.rz * w; c.wsum += w;
		}
	}
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		S.numNeigh[s.i] = c.cnt;
		const double ps = c.ni + S.pndWallContribution[s.i];
		if (K.p.pndType == 0 || K.p.pndType == 1) { S.pndki[s.i] = S.pndi[s.i]; S.pndi[s.i] = ps; }
		S.pndSmall[s.i] = ps;
		npcd_finish(S, K, s.i, c.dx, c.dy, c.dz, c.wsum, ps);
	}
};

// MpsPndNeigh::setInitialPndNumberOfNeighNPCD, src/MpsPndNeigh.cpp:21-101 (0-th step)
struct BInitialPnd {
	struct Ctx { double acc, wsum, dx, dy, dz; int cnt; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		c.acc = S.pndi[s.i]; c.wsum = 0.0; c.cnt = 0; c.dx = c.dy = c.dz = 0.0; return true;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const DevView& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_P<POLY>(q, K.p.reL2)) return;
		c.cnt += 1;
		if (q.d2 < K.p.reS2) {
			const double w = wgt(sqrt(q.d2), K.p.reS, K.p.weightType);
			c.acc += w; c.dx += q.rx * w; c.dy += q.ry * w; c.dz += q.rz * w; c.wsum += w;
		}
	}
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		double a = c.acc + S.pndWallContribution[s.i];
		if (s.type == K.p.wall) a = K.p.pndSmallZero;
		S.pndi[s.i] = a; S.pndSmall[s.i] = a; S.pndki[s.i] = a;
		S.numNeigh[s.i] = c.cnt + S.numNeighWallContribution[s.i];
		npcd_finish(S, K, s.i, c.dx, c.dy, c.dz, c.wsum, a);
	}
};

// MpsPndNeigh::calcPndDiffusiveTerm, src/MpsPndNeigh.cpp:334-489
struct BPndDiffusive {
	struct Ctx { double ni, vx, vy, vz, Pi, rho, Di, div; double c1[3], c2[3], c3[3]; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		c.ni = S.pndi[s.i];
		if (c.ni < K.p.epsilonZero) return false;
		const double4 v = S.vel[s.i]; c.vx = v.x; c.vy = v.y; c.vz = v.z; c.Pi = S.press[s.i]; c.rho = S.RHO[s.i];
		if (K.p.gradientCorrection) {
			c.c1[0] = S.correcMatrixRow1.x[s.i]; c.c1[1] = S.correcMatrixRow1.y[s.i]; c.c1[2] = S.correcMatrixRow1.z[s.i];
			c.c2[0] = S.correcMatrixRow2.x[s.i]; c.c2[1] = S.correcMatrixRow2.y[s.i]; c.c2[2] = S.correcMatrixRow2.z[s.i];
			c.c3[0] = S.correcMatrixRow3.x[s.i]; c.c3[1] = S.correcMatrixRow3.y[s.i]; c.c3[2] = S.correcMatrixRow3.z[s.i];
		}
		c.Di = c.div = 0.0; return true;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const DevView& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_P<POLY>(q, K.p.reL2)) return;
		const double r = sqrt(q.d2); const double wL = wgt(r, K.p.reL, K.p.weightType);
		if (s.type == K.p.fluid && pk_type(q.pj) == K.p.fluid) {
			const double rhoj = S.RHO[q.j];
			const double pgh = -2.0 * (c.rho * rhoj / (c.rho + rhoj)) * (K.p.gravityX * q.rx + K.p.gravityY * q.ry + K.p.gravityZ * q.rz);
			c.Di += K.p.timeStep / c.rho * K.p.coeffViscMultiphase * (S.press[q.j] - c.Pi - pgh) * wL;
		}
		if (q.d2 < K.p.reS2) {
			const double4 vj = S.vel[q.j];
			const double ux = vj.x - c.vx, uy = vj.y - c.vy, uz = vj.z - c.vz;
			const double wS = wgt(r, K.p.reS, K.p.weightType);
			if (!K.p.gradientCorrection) {
				if (c.ni > K.p.epsilonZero) c.div += ((double)K.p.dim / K.p.pndSmallZero) * (S.pndi[q.j] / c.ni) * (ux * q.rx + uy * q.ry + uz * q.rz) * wS / q.d2;
			} else {
				const double qx = (q.rx * c.c1[0] + q.ry * c.c1[1] + q.rz * c.c1[2]), qy = (q.rx * c.c2[0] + q.ry * c.c2[1] + q.rz * c.c2[2]), qz = (q.rx * c.c3[0] + q.ry * c.c3[1] + q.rz * c.c3[2]);
				c.div += ((double)K.p.dim / K.p.pndSmallZero) * (ux * qx + uy * qy + uz * qz) * wS / q.d2;
			}
		}
	}
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		S.pndAuxVar1[s.i] = c.ni * (1.0 + K.p.timeStep * (-c.div + c.Di * 1.0));
		S.velDivergence[s.i] = c.div; S.diffusiveTerm[s.i] = c.Di;
	}
};

// MpsPndNeigh::calcWallNoSlipPndDiffusiveTerm / calcWallSlipPndDiffusiveTerm, src/MpsPndNeigh.cpp:492-744
struct BWallPndDiffusive {
	struct Ctx { double ni, v[3], vm[3], div; WallFrame f; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		c.ni = S.pndi[s.i];
		if (!(s.type == K.p.fluid && S.particleNearWall[s.i] && c.ni > K.p.epsilonZero)) return false;
		const double4 v = S.vel[s.i]; c.v[0] = v.x; c.v[1] = v.y; c.v[2] = v.z;
		wall_frame(c.f, s.pi.x, s.pi.y, s.pi.z, s.mix, s.miy, s.miz, K.p.epsilonZero, true);
		mirror_velocity(K, S.nearMeshType[s.i], c.f, c.v, c.vm);
		c.div = 0.0; return true;
	}
	__device__ __forceinline__ static double term(const Ctx& c, const KParams& K, double vjx, double vjy, double vjz, double mx, double my, double mz, double pndj, double dd) {
		const double wS = wgt(sqrt(dd), K.p.reS, K.p.weightType);
		double ux = -(vjx - c.vm[0]), uy = -(vjy - c.vm[1]), uz = -(vjz - c.vm[2]);
		double qx = MATVEC(c.f.R, mx, my, mz, 0), qy = MATVEC(c.f.R, mx, my, mz, 1), qz = MATVEC(c.f.R, mx, my, mz, 2);
		if (K.p.slipCondition == 1) { ux = vjx - c.vm[0]; uy = vjy - c.vm[1]; uz = vjz - c.vm[2]; qx = mx; qy = my; qz = mz; }
		return ((double)K.p.dim / K.p.pndSmallZero) * (pndj / c.ni) * (ux * qx + uy * qy + uz * qz) * wS / dd;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const DevView& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_M(q, K.p.reS2)) return;
		const double4 vj = S.vel[q.j];
		c.div += term(c, K, vj.x, vj.y, vj.z, q.mx, q.my, q.mz, S.pndi[q.j], q.dm2);
	}
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		const double ex = s.pi.x - s.mix, ey = s.pi.y - s.miy, ez = s.pi.z - s.miz; const double dd = sq3_exact(ex, ey, ez);
		if (dd < K.p.reS2) c.div += term(c, K, c.v[0], c.v[1], c.v[2], ex, ey, ez, c.ni, dd);
		S.pndAuxVar1[s.i] += -c.ni * K.p.timeStep * c.div;
	}
};

// sum_j pndi_j w / sum_j w over P(reS): first half of MpsPndNeigh::meanPndParticlesWallDummySurface (src/MpsPndNeigh.cpp:777-852)
// and of meanPnd (:855-920).  SEL 0: wall or surface particles, overwrite; SEL 1: every particle, accumulate (meanPnd)
template <int SEL> struct BMeanPnd {
	struct Ctx { double up, dn; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		c.up = c.dn = 0.0;
		if (SEL == 0) {
			S.pndAuxVar1[s.i] = 0.0; S.pndAuxVar2[s.i] = 0.0;
			return s.type == K.p.wall || S.particleBC[s.i] == K.p.surface;
		}
		return true;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const DevView& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_P<POLY>(q, K.p.reS2)) return;
		const double w = wgt(sqrt(q.d2), K.p.reS, K.p.weightType); c.up += S.pndi[q.j] * w; c.dn += w;
	}
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
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const DevView& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_M(q, K.p.reS2)) return;
		const double w = wgt(sqrt(q.dm2), K.p.reS, K.p.weightType); c.up += S.pndi[q.j] * w; c.dn += w;
	}
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
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const DevView& S, const KParams& K, const Self& s, const Pair& q) {
		if (!(pred_P<POLY>(q, K.p.reS2) && pk_type(q.pj) == K.p.fluid)) return;
		const double w = wgt(sqrt(q.d2), K.p.reS, K.p.weightType); c.up += S.pndki[q.j] * w; c.dn += w;
	}
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		S.pndski[s.i] = c.dn > K.p.epsilonZero ? c.up / c.dn : c.up;
	}
};

// ------------------------------------------------------------------------------------------------ pressure gradient
__device__ __forceinline__ double grad_phi(const KParams& K, double w, double Pj, double Pi, double Pmin, double ni, double nj, double dd) {
	switch (K.p.gradientType) {   // src/MpsPressure.cpp:1288-1300
		case 0: return w * ((Pj - Pmin) / dd);
		case 1: return w * ((Pj + Pi) / dd);
		case 2: return w * ((Pj + Pi - 2.0 * Pmin) / dd);
		default: return (ni > K.p.epsilonZero && nj > K.p.epsilonZero) ? w * ((ni * Pj / nj + nj * Pi / ni) / dd) : w;
	}
}

// first pass of calcPressGradient / calcWallPressGradient for gradient types 0 and 2: Pmin over P(reS)
// (src/MpsPressure.cpp:1237-1262, :1405-1430).  Result parked in a scratch plane that is not a reference field.
struct BPressMin {
	struct Ctx { double pmin; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) { c.pmin = S.press[s.i]; return true; }
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const DevView& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_P<POLY>(q, K.p.reS2)) return;
		const double pj = S.press[q.j]; if (c.pmin > pj) c.pmin = pj;
	}
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) { S.scratch1[s.i] = c.pmin; }
};

// MpsPressure::calcPressGradient, src/MpsPressure.cpp:1217-1338
struct BPressGradient {
	struct Ctx { double Pi, ni, Pmin, sx, sy, sz; double c1[3], c2[3], c3[3]; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		c.Pi = S.press[s.i]; c.ni = S.pndi[s.i];
		c.Pmin = (K.p.gradientType == 0 || K.p.gradientType == 2) ? S.scratch1[s.i] : c.Pi;
		if (K.p.gradientCorrection) {
			c.c1[0] = S.correcMatrixRow1.x[s.i]; c.c1[1] = S.correcMatrixRow1.y[s.i]; c.c1[2] = S.correcMatrixRow1.z[s.i];
			c.c2[0] = S.correcMatrixRow2.x[s.i]; c.c2[1] = S.correcMatrixRow2.y[s.i]; c.c2[2] = S.correcMatrixRow2.z[s.i];
			c.c3[0] = S.correcMatrixRow3.x[s.i]; c.c3[1] = S.correcMatrixRow3.y[s.i]; c.c3[2] = S.correcMatrixRow3.z[s.i];
		}
		c.sx = c.sy = c.sz = 0.0; return true;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const DevView& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_P<POLY>(q, K.p.reS2)) return;
		const double w = grad_phi(K, wgt(sqrt(q.d2), K.p.reS, K.p.weightType), S.press[q.j], c.Pi, c.Pmin, c.ni, S.pndi[q.j], q.d2);
		if (!K.p.gradientCorrection) { c.sx += q.rx * w; c.sy += q.ry * w; c.sz += q.rz * w; }
		else {
			c.sx += (q.rx * c.c1[0] + q.ry * c.c1[1] + q.rz * c.c1[2]) * w; c.sy += (q.rx * c.c2[0] + q.ry * c.c2[1] + q.rz * c.c2[2]) * w;
			c.sz += (q.rx * c.c3[0] + q.ry * c.c3[1] + q.rz * c.c3[2]) * w;
		}
	}
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		const double rho = S.RHO[s.i];
		S.acc.x[s.i] = K.p.relaxPress * c.sx * K.p.coeffPressGrad / rho;
		S.acc.y[s.i] = K.p.relaxPress * c.sy * K.p.coeffPressGrad / rho;
		S.acc.z[s.i] = K.p.relaxPress * c.sz * K.p.coeffPressGrad / rho;
	}
};

// wall repulsion, src/MpsPressure.cpp:1554-1611: the scalar that multiplies -n
__device__ __forceinline__ double repulsion(const KParams& K, double r, double rho, double vmax2) {
	const double re = K.p.reRepulsiveForce;
	switch (K.p.repulsiveForceType) {
		case 0: return rho / (K.p.timeStep * K.p.timeStep) * (re - r);
		case 1: return K.p.repForceCoefMitsume * wgt(r, re, K.p.weightType);
		case 2: { const double R1 = (re / r) * (re / r); const double R2 = R1 * R1; return (K.p.repForceCoefLennardJones * vmax2 / r) * (R2 - R1) * rho; }
		default: { const double W1 = (1.0 + 3.0 * 0.5 * r / (re)); const double W2 = (1.0 - r / (re)) * (1.0 - r / (re)) * (1.0 - r / (re));
			return (K.p.repForceCoefMonaghanKajtar * vmax2 / (r - 0.0 * K.p.partDist)) * (1.0 / 8.0) * (W1) * (W2)*rho; }
	}
}

// MpsPressure::calcWallPressGradient, src/MpsPressure.cpp:1341-1550.  velMax of the previous correction is read
// from the device scalar so the step needs no host round trip.
struct BWallPressGradient {
	struct Ctx { double Pi, ni, Pmin, sx, sy, sz; WallFrame f; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		if (!(s.type == K.p.fluid && S.particleNearWall[s.i])) return false;
		c.Pi = S.press[s.i]; c.ni = S.pndi[s.i];
		c.Pmin = (K.p.gradientType == 0 || K.p.gradientType == 2) ? S.scratch1[s.i] : c.Pi;
		wall_frame(c.f, s.pi.x, s.pi.y, s.pi.z, s.mix, s.miy, s.miz, K.p.epsilonZero, false);
		c.sx = c.sy = c.sz = 0.0; return true;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const DevView& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_M(q, K.p.reS2)) return;
		const double w = grad_phi(K, wgt(sqrt(q.dm2), K.p.reS, K.p.weightType), S.press[q.j], c.Pi, c.Pmin, c.ni, S.pndi[q.j], q.dm2);
		c.sx += q.mx * w; c.sy += q.my * w; c.sz += q.mz * w;
	}
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		const double ex = s.pi.x - s.mix, ey = s.pi.y - s.miy, ez = s.pi.z - s.miz; const double dd = sq3_exact(ex, ey, ez);
		if (dd < K.p.reS2) {
			const double w = grad_phi(K, wgt(sqrt(dd), K.p.reS, K.p.weightType), c.Pi, c.Pi, c.Pmin, c.ni, c.ni, dd);
			c.sx += ex * w; c.sy += ey * w; c.sz += ez * w;
		}
		const double rho = S.RHO[s.i];
		double fr[3] = { 0.0, 0.0, 0.0 };
		const double r = c.f.len;
		if (r < K.p.reRepulsiveForce && r > K.p.epsilonZero) {
			const double velMax = sqrt(__longlong_as_double((long long)S.velMax2Bits[1]));
			const double vcap = velMax < K.p.expectMaxVelocity ? velMax : K.p.expectMaxVelocity;
			const double qq = repulsion(K, r, rho, vcap * vcap);
			fr[0] = -qq * c.f.n[0]; fr[1] = -qq * c.f.n[1]; fr[2] = -qq * c.f.n[2];
		}
		double* a[3] = { S.acc.x, S.acc.y, S.acc.z }; double* fw[3] = { S.forceWall.x, S.forceWall.y, S.forceWall.z };
		for (int k = 0; k < 3; k++) {
			const double g = (K.p.relaxPress * MATVEC(c.f.R, c.sx, c.sy, c.sz, k) * K.p.coeffPressGrad - fr[k]);
			a[k][s.i] += g / rho;
			fw[k][s.i] += -g * K.vol;
		}
	}
};

// ------------------------------------------------------------------------------------------------ shifting (type 2)
// MpsShifting::calcConcAndConcGradient, src/MpsShifting.cpp:432-575 — first loop: concentration
struct BConcentration {
	struct Ctx { double c; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) { c.c = 0.0; return true; }
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const DevView& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_P<POLY>(q, K.p.reS2)) return;
		c.c += wgt(sqrt(q.d2), K.p.reS, K.p.weightType);
	}
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		S.concentration[s.i] = (c.c + S.pndWallContribution[s.i]) / K.p.pndSmallZero;
	}
};

// second loop: concentration gradient and the shift.  Jacobi semantics: positions are READ from S.pos (the snapshot)
// and the shifted position is WRITTEN to S.pos2; the reference's in-place update is a race for >1 thread
// (src/MpsShifting.cpp:478-536).
struct BGradConcentration {
	typedef void copy_pos_tag;   // the walker mirrors pos -> pos2 for the slots this body skips
	struct Ctx { double ci, gx, gy, gz; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		if (s.type != K.p.fluid) return false;
		c.ci = S.concentration[s.i]; c.gx = c.gy = c.gz = 0.0; return true;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const DevView& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_P<POLY>(q, K.p.reS2)) return;
		const double w = wgt(sqrt(q.d2), K.p.reS, K.p.weightType); const double cj = S.concentration[q.j];
		c.gx += (c.ci + cj) * q.rx * w / q.d2; c.gy += (c.ci + cj) * q.ry * w / q.d2; c.gz += (c.ci + cj) * q.rz * w / q.d2;
	}
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		const double g0 = c.gx * -K.p.coeffPressGrad, g1 = c.gy * -K.p.coeffPressGrad, g2 = c.gz * -K.p.coeffPressGrad;
		S.gradConcentration.x[s.i] = g0; S.gradConcentration.y[s.i] = g1; S.gradConcentration.z[s.i] = g2;
		if (S.particleBC[s.i] == K.p.inner) {
			double4 p = s.pi; p.x -= K.p.coeffShifting2 * g0; p.y -= K.p.coeffShifting2 * g1; p.z -= K.p.coeffShifting2 * g2;
			S.pos2[s.i] = p;
			const float4 b = S.bpos[s.i];
			const float ddx = (float)(p.x - K.p.domainMinX) - b.x, ddy = (float)(p.y - K.p.domainMinY) - b.y, ddz = (float)(p.z - K.p.domainMinZ) - b.z;
			atomic_max_f(S.maxDisp2, ddx * ddx + ddy * ddy + ddz * ddz);
		}
	}
};

// MpsShifting::calcWallConcAndConcGradient, src/MpsShifting.cpp:578-691 (same read-snapshot / write-alternate rule)
struct BWallGradConcentration {
	typedef void copy_pos_tag;
	struct Ctx { double ci, sx, sy, sz; WallFrame f; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		if (!(s.type == K.p.fluid && S.particleNearWall[s.i])) return false;
		c.ci = S.concentration[s.i];
		wall_frame(c.f, s.pi.x, s.pi.y, s.pi.z, s.mix, s.miy, s.miz, K.p.epsilonZero, false);
		c.sx = c.sy = c.sz = 0.0; return true;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const DevView& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_M(q, K.p.reS2)) return;
		const double w = wgt(sqrt(q.dm2), K.p.reS, K.p.weightType); const double cj = S.concentration[q.j];
		c.sx += (c.ci + cj) * q.mx * w / q.dm2; c.sy += (c.ci + cj) * q.my * w / q.dm2; c.sz += (c.ci + cj) * q.mz * w / q.dm2;
	}
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		const double ex = s.pi.x - s.mix, ey = s.pi.y - s.miy, ez = s.pi.z - s.miz; const double dd = sq3_exact(ex, ey, ez);
		if (dd < K.p.reS2) {
			const double w = wgt(sqrt(dd), K.p.reS, K.p.weightType);
			c.sx += (c.ci + c.ci) * ex * w / dd; c.sy += (c.ci + c.ci) * ey * w / dd; c.sz += (c.ci + c.ci) * ez * w / dd;
		}
		double* g[3] = { S.gradConcentration.x, S.gradConcentration.y, S.gradConcentration.z };
		double p[3] = { s.pi.x, s.pi.y, s.pi.z };
		const bool inner = S.particleBC[s.i] == K.p.inner;
		for (int k = 0; k < 3; k++) {
			const double gw = -MATVEC(c.f.R, c.sx, c.sy, c.sz, k) * K.p.coeffPressGrad;
			g[k][s.i] += gw;
			if (inner) p[k] -= K.p.coeffShifting2 * gw;
		}
		if (inner) {
			double4 pn = s.pi; pn.x = p[0]; pn.y = p[1]; pn.z = p[2]; S.pos2[s.i] = pn;
			const float4 b = S.bpos[s.i];
			const float ddx = (float)(pn.x - K.p.domainMinX) - b.x, ddy = (float)(pn.y - K.p.domainMinY) - b.y, ddz = (float)(pn.z - K.p.domainMinZ) - b.z;
			atomic_max_f(S.maxDisp2, ddx * ddx + ddy * ddy + ddz * ddz);
		}
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
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const DevView& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_P<POLY>(q, K.p.reS2)) return;
		const double w = wgt(sqrt(q.d2), K.p.reS, K.p.weightType); const double inv = 1.0 / q.d2; const double r[3] = { q.rx, q.ry, q.rz };
		for (int a = 0; a < 3; a++) for (int b = 0; b < 3; b++) c.M[3 * a + b] += w * r[a] * r[b] * inv;
	}
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		for (int k = 0; k < 9; k++) c.M[k] *= -K.p.coeffPressGrad;
		invert3(c.M, K.p.dim, K.p.epsilonZero);
		if (S.numNeigh[s.i] < 4) for (int k = 0; k < 9; k++) c.M[k] = (k % 4 == 0) ? 1.0 : 0.0;
		S.correcMatrixRow1.x[s.i] = c.M[0]; S.correcMatrixRow1.y[s.i] = c.M[1]; S.correcMatrixRow1.z[s.i] = c.M[2];
		S.correcMatrixRow2.x[s.i] = c.M[3]; S.correcMatrixRow2.y[s.i] = c.M[4]; S.correcMatrixRow2.z[s.i] = c.M[5];
		S.correcMatrixRow3.x[s.i] = c.M[6]; S.correcMatrixRow3.y[s.i] = c.M[7]; S.correcMatrixRow3.z[s.i] = c.M[8];
	}
};

// ------------------------------------------------------------------------------------------------ particle walls
// MpsPressure::extrapolatePressParticlesWallDummy, src/MpsPressure.cpp:792-846.  Only non-fluid i are written and
// only fluid j are read, so the in-place update has no read/write hazard.
struct BExtrapolatePress {
	struct Ctx { double ni, pr; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		c.ni = c.pr = 0.0;
		return s.type == K.p.dummyWall || (K.p.mpsType == 1 && s.type == K.p.wall);
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const DevView& S, const KParams& K, const Self& s, const Pair& q) {
		if (!(q.d2 < K.p.reS2 && pk_type(q.pj) == K.p.fluid)) return;
		const double w = wgt(sqrt(q.d2), K.p.reS, K.p.weightType); c.ni += w;
		c.pr += (S.press[q.j] - S.RHO[q.j] * (K.p.gravityX * q.rx + K.p.gravityY * q.ry + K.p.gravityZ * q.rz)) * w;
	}
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		if (c.pr < 0.0) return;
		S.press[s.i] = c.ni > 0 ? c.pr / c.ni : c.pr;   // only non-fluid i are written, only fluid j are read: no hazard
	}
};

// MpsParticleVelPos::updateVelocityParticlesWallDummy, src/MpsParticleVelPos.cpp:100-158
struct BWallDummyVel {
	struct Ctx { double ni, ux, uy, uz; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		c.ni = c.ux = c.uy = c.uz = 0.0; return s.type == K.p.dummyWall;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const DevView& S, const KParams& K, const Self& s, const Pair& q) {
		if (!(q.d2 < K.p.reL2 && pk_type(q.pj) == K.p.fluid)) return;
		const double w = wgt(sqrt(q.d2), K.p.reL, K.p.weightType); const double4 vj = S.vel[q.j];
		c.ni += w; c.ux += vj.x * w; c.uy += vj.y * w; c.uz += vj.z * w;
	}
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		double4 v = S.vel[s.i];
		if (c.ni > K.p.epsilonZero) { v.x = 2.0 * 0.0 - c.ux / c.ni; v.y = 2.0 * 0.0 - c.uy / c.ni; v.z = 2.0 * 0.0 - c.uz / c.ni; }
		else { v.x = 2.0 * 0.0 - c.ux; v.y = 2.0 * 0.0 - c.uy; v.z = 2.0 * 0.0 - c.uz; }
		S.vel[s.i] = v;   // only dummy i are written, only fluid j are read: no hazard
	}
};
