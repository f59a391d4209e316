// pmps_bodies.cuh — one stage body per reference neighbour loop.  A body is the {begin, pair, end} triple that the
// list walker (pmps_device.cuh) runs for slot i; several bodies fuse into one pass when their inputs coincide
// (pmps_stages.inc lists the fused groups of the whole-step path).  Each body cites the reference method whose loop
// body it replaces.  Sums follow the arithmetic policy of pmps_device.cuh: predicates exact, weights from the 1-ulp
// reciprocal distance, divisions by per-particle constants hoisted to begin().
#pragma once
#include "pmps_device.cuh"
#include "pmps_wall.cuh"

// ------------------------------------------------------------------------------------------------ viscosity
// MpsViscosity::calcViscosity, src/MpsViscosity.cpp:20-100: acc = coeff * sum (v_j - v_i) w_L nu_ij over P(reL)
struct BViscosity {
	struct Ctx { double vx, vy, vz, mu, invRho, sx, sy, sz; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		if (TV_OWN_STAGED && s.nb) { c.vx = s.nb->vx[s.ms]; c.vy = s.nb->vy[s.ms]; c.vz = s.nb->vz[s.ms]; }
		else { const double4 v = S.vel[s.i]; c.vx = v.x; c.vy = v.y; c.vz = v.z; }
		c.mu = OWN(MEU, S.MEU[s.i]); c.invRho = 1.0 / S.RHO[s.i];
		c.sx = c.sy = c.sz = 0.0; return true;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const Nb& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_P<POLY>(q, K.p.reL2)) return;
		const double w = wgt_r(q.rinv, q.d2, K.p.reL, K.invReL, K.p.weightType);
		const double muj = S.MEU[q.j];
		double nu = (c.mu + muj) > K.p.epsilonZero ? 2.0 * c.mu * muj * fast_rcp(c.mu + muj) : 0.0;
		if (q.tj == K.p.wall) nu = 2.0 * c.mu;
		const double f = w * nu * c.invRho;
		const double vjx = S.vx[q.j], vjy = S.vy[q.j], vjz = S.vz[q.j];
		c.sx += (vjx - c.vx) * f; c.sy += (vjy - c.vy) * f; c.sz += (vjz - c.vz) * f;
	}
	__device__ __forceinline__ static void reduce(Ctx& c, unsigned m) { c.sx = gsum(c.sx, m); c.sy = gsum(c.sy, m); c.sz = gsum(c.sz, m); }
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		S.acc.x[s.i] = K.p.coeffViscMultiphase * c.sx; S.acc.y[s.i] = K.p.coeffViscMultiphase * c.sy; S.acc.z[s.i] = K.p.coeffViscMultiphase * c.sz;
	}
};

// the two integrators, shared by the elementwise stage kernels and the fused tails (exact arithmetic: positions and
// velocities feed every predicate of the next pass).  src/MpsParticleVelPos.cpp:20-50 and :53-97
__device__ __forceinline__ void integrate_predict(double4& p, double4& v, double ax, double ay, double az, double dt) {
	v.x = madd_exact(ax, dt, v.x); v.y = madd_exact(ay, dt, v.y); v.z = madd_exact(az, dt, v.z);
	p.x = madd_exact(v.x, dt, p.x); p.y = madd_exact(v.y, dt, p.y); p.z = madd_exact(v.z, dt, p.z);
}
__device__ __forceinline__ void integrate_correct(double4& p, double4& v, double ax, double ay, double az, double dt) {
	v.x = madd_exact(ax, dt, v.x); v.y = madd_exact(ay, dt, v.y); v.z = madd_exact(az, dt, v.z);
	p.x = __dadd_rn(p.x, __dmul_rn(__dmul_rn(ax, dt), dt)); p.y = __dadd_rn(p.y, __dmul_rn(__dmul_rn(ay, dt), dt)); p.z = __dadd_rn(p.z, __dmul_rn(__dmul_rn(az, dt), dt));
}
// per-step resets of predictionVelocityPosition (src/MpsParticleVelPos.cpp:30-44)
__device__ __forceinline__ void predict_resets(const DevView& S, const KParams& K, int i) {
	S.acc.x[i] = 0.0; S.acc.y[i] = 0.0; S.acc.z[i] = 0.0;
	S.dvelCollision.x[i] = 0.0; S.dvelCollision.y[i] = 0.0; S.dvelCollision.z[i] = 0.0;
	S.npcdDeviation.x[i] = 0.0; S.npcdDeviation.y[i] = 0.0; S.npcdDeviation.z[i] = 0.0;
	S.numNeighWallContribution[i] = 0; S.particleNearWall[i] = 0; S.distParticleWall2[i] = K.big;
	if (K.poly) { S.mirrorParticlePos.x[i] = K.big; S.mirrorParticlePos.y[i] = K.big; S.mirrorParticlePos.z[i] = K.big; }
}

// fused tail of the viscosity pass: addGravity (src/MpsParticle.cpp:111-130) + predictionVelocityPosition.  Reads the
// acceleration BViscosity::end just stored, writes the NEW position/velocity to the alternate buffers (neighbours are
// still reading the old ones in this launch).
struct BPredictTail {
	typedef void copy_pos_tag; typedef void copy_vel_tag; typedef void no_pair_tag;
	struct Ctx {};
	__device__ __forceinline__ static bool begin(Ctx&, const DevView&, const KParams&, const Self&) { return true; }
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx&, const Nb&, const KParams&, const Self&, const Pair&) {}
	__device__ __forceinline__ static void reduce(Ctx& , unsigned m) {  }
	__device__ __forceinline__ static void end(Ctx&, const DevView& S, const KParams& K, const Self& s) {
		const int i = s.i;
		const double ax = S.acc.x[i] + K.p.gravityX, ay = S.acc.y[i] + K.p.gravityY, az = S.acc.z[i] + K.p.gravityZ;
		S.accStar.x[i] = ax; S.accStar.y[i] = ay; S.accStar.z[i] = az;
		if (s.type == K.p.fluid) {
			double4 p = s.pi, v = S.vel[i];
			S.stepCell[i] = ref_bucket_insert(K, s.pi);   // where the reference's lists of this step hold the particle (pmps_domain.cu)
			integrate_predict(p, v, ax, ay, az, K.p.timeStep);
			S.pos2[i] = p; S.vel2[i] = v;
			track_disp(S, K, i, p);
		}
		predict_resets(S, K, i);
	}
};

// MpsViscosity::calcWallNoSlipViscosity / calcWallSlipViscosity, src/MpsViscosity.cpp:1527-1910 (near-wall fluid)
struct BWallViscosity {
	typedef void wall_frame_tag;
	struct Ctx { double v[3], vm[3], mu, invRhoDiv, sx, sy, sz; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		if (!s.nearWallFluid) return false;
		const double4 v = S.vel[s.i]; c.v[0] = v.x; c.v[1] = v.y; c.v[2] = v.z; c.mu = S.MEU[s.i];
		WallFrame f; wall_frame_of(f, s, K.p.epsilonZero, true);
		mirror_velocity(K, S.nearMeshType[s.i], f, c.v, c.vm);
		c.invRhoDiv = 1.0 / (s.ptype == 1 ? K.p.DNS_FL1 : K.p.DNS_FL2);
		c.sx = c.sy = c.sz = 0.0; return true;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const Nb& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_M(q, K.p.reL2)) return;
		const double w = wgt_r(fast_rsqrt(q.dm2), q.dm2, K.p.reL, K.invReL, K.p.weightType);
		const double muj = S.MEU[q.j];
		double nu = (c.mu + muj) > K.p.epsilonZero ? 2.0 * c.mu * muj * fast_rcp(c.mu + muj) : 0.0;
		const double f = w * nu * c.invRhoDiv;
		const double vjx = S.vx[q.j], vjy = S.vy[q.j], vjz = S.vz[q.j];
		c.sx += (vjx - c.vm[0]) * f; c.sy += (vjy - c.vm[1]) * f; c.sz += (vjz - c.vm[2]) * f;
	}
	__device__ __forceinline__ static void reduce(Ctx& c, unsigned m) { c.sx = gsum(c.sx, m); c.sy = gsum(c.sy, m); c.sz = gsum(c.sz, m); }
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		const double ex = s.pi.x - s.mix, ey = s.pi.y - s.miy, ez = s.pi.z - s.miz; const double dd = sq3_exact(ex, ey, ez);
		if (dd < K.p.reL2) {   // i itself is a neighbour of its mirror image
			const double w = wgt(sqrt(dd), K.p.reL, K.p.weightType);
			double nu = c.mu > K.p.epsilonZero ? 2.0 * c.mu * c.mu / (c.mu + c.mu) : 0.0;
			nu = nu * c.invRhoDiv;
			c.sx += (c.v[0] - c.vm[0]) * w * nu; c.sy += (c.v[1] - c.vm[1]) * w * nu; c.sz += (c.v[2] - c.vm[2]) * w * nu;
		}
		double t[3];
		if (K.p.slipCondition == 0) { t[0] = -c.sx; t[1] = -c.sy; t[2] = -c.sz; }   // R_inv = -I
		else { WallFrame f; wall_frame_of(f, s, K.p.epsilonZero, true); for (int k = 0; k < 3; k++) t[k] = MATVEC(f.R, c.sx, c.sy, c.sz, k); }
		const double rho = S.RHO[s.i];
		double* a[3] = { S.acc.x, S.acc.y, S.acc.z }; double* fw[3] = { S.forceWall.x, S.forceWall.y, S.forceWall.z };
		for (int k = 0; k < 3; k++) {
			a[k][s.i] += t[k] * K.p.coeffViscMultiphase;
			fw[k][s.i] += -t[k] * K.p.coeffViscMultiphase * K.vol * rho;
		}
	}
};

// ------------------------------------------------------------------------------------------------ collisions (PC)
// MpsParticleCollision::checkParticleCollisions, src/MpsParticleCollision.cpp:22-119 (first loop: dvel)
struct BCollision {
	typedef void close_pairs_tag;
	struct Ctx { double vx, vy, vz, mi, dx, dy, dz; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		if (s.type != K.p.fluid) return false;
		const double4 v = S.vel[s.i]; c.vx = v.x; c.vy = v.y; c.vz = v.z;
		c.mi = s.ptype == 1 ? K.p.DNS_FL1 : K.p.DNS_FL2; c.dx = c.dy = c.dz = 0.0; return true;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const Nb& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_P<POLY>(q, K.p.distCollisionLimit2)) return;
		const double vjx = S.vx[q.j], vjy = S.vy[q.j], vjz = S.vz[q.j];
		double f = (c.vx - vjx) * q.rx + (c.vy - vjy) * q.ry + (c.vz - vjz) * q.rz;
		if (f > 0.0) {
			const bool jf = q.tj == K.p.fluid;
			const double mj = jf ? (q.ptj == 1 ? K.p.DNS_FL1 : K.p.DNS_FL2) : K.p.densityWall;
			f *= K.p.restitutionCollision * mj / (c.mi + mj) / q.d2;
			if (jf) { c.dx -= q.rx * f; c.dy -= q.ry * f; c.dz -= q.rz * f; }
			else { c.dx -= 2.0 * q.rx * f; c.dy -= 2.0 * q.ry * f; c.dz -= 2.0 * q.rz * f; }
		}
	}
	__device__ __forceinline__ static void reduce(Ctx& c, unsigned m) { c.dx = gsum(c.dx, m); c.dy = gsum(c.dy, m); c.dz = gsum(c.dz, m); }
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		S.dvelCollision.x[s.i] = c.dx; S.dvelCollision.y[s.i] = c.dy; S.dvelCollision.z[s.i] = c.dz;
	}
};
// fused tail: second loop of checkParticleCollisions (:108-116), vel2 = vel + dvel (neighbours still read vel)
struct BCollisionTail {
	typedef void copy_vel_tag; typedef void no_pair_tag;
	struct Ctx {};
	__device__ __forceinline__ static bool begin(Ctx&, const DevView&, const KParams& K, const Self& s) { return s.type == K.p.fluid; }
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx&, const Nb&, const KParams&, const Self&, const Pair&) {}
	__device__ __forceinline__ static void reduce(Ctx& , unsigned m) {  }
	__device__ __forceinline__ static void end(Ctx&, const DevView& S, const KParams& K, const Self& s) {
		double4 v = S.vel[s.i];
		v.x = __dadd_rn(v.x, S.dvelCollision.x[s.i]); v.y = __dadd_rn(v.y, S.dvelCollision.y[s.i]); v.z = __dadd_rn(v.z, S.dvelCollision.z[s.i]);
		S.vel2[s.i] = v;
	}
};
// dynamic particle collision (collision type 1), src/MpsParticleCollision.cpp:124-292
struct BCollisionDPC {
	typedef void close_pairs_tag;
	struct Ctx { double vx, vy, vz, mi, Pi, rho, dx, dy, dz; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		if (s.type != K.p.fluid) return false;
		const double4 v = S.vel[s.i]; c.vx = v.x; c.vy = v.y; c.vz = v.z; c.Pi = S.press[s.i]; c.rho = S.RHO[s.i];
		c.mi = s.ptype == 1 ? K.p.DNS_FL1 : K.p.DNS_FL2; c.dx = c.dy = c.dz = 0.0; return true;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const Nb& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_P<POLY>(q, K.p.partDist * K.p.partDist)) return;
		const bool jf = q.tj == K.p.fluid;
		const double mj = jf ? (q.ptj == 1 ? K.p.DNS_FL1 : K.p.DNS_FL2) : K.p.densityWall;
		const double dst = sqrt(q.d2), invLo = 1.0 / K.p.partDist;
		double Wij = 0.0;
		if (dst > K.p.epsilonZero) { const double w1 = 1.0 - dst * invLo; const double w2 = 4.0 * dst * invLo + 1.0; Wij = w1 * w1 * w1 * w1 * w2; }
		const double chi = sqrt(Wij / (0.5 * 0.5 * 0.5 * 0.5 * 3.0));
		const double kappa = dst < 0.5 * K.p.partDist ? 1.0 : chi;
		double f = (c.vx - S.vx[q.j]) * q.rx + (c.vy - S.vy[q.j]) * q.ry + (c.vz - S.vz[q.j]) * q.rz;
		if (f > 0.0) f *= kappa * 2.0 * mj / (c.mi + mj) / q.d2;
		else {
			const double gmod = sqrt(K.p.gravityX * K.p.gravityX + K.p.gravityY * K.p.gravityY + K.p.gravityZ * K.p.gravityZ);
			const double pmax = 2.0 / 3.0 * c.rho * gmod * 0.2, pmin = c.rho * gmod * K.p.partDist;
			const double a = K.p.lambdaCollision * fabs(c.Pi + S.press[q.j]), b = K.p.lambdaCollision * pmax;
			const double lo = a < b ? a : b; const double ptil = lo > pmin ? lo : pmin;
			f = K.p.timeStep / c.mi * chi * (ptil * chi) / q.d2;
		}
		if (jf) { c.dx -= q.rx * f; c.dy -= q.ry * f; c.dz -= q.rz * f; }
		else { c.dx -= 2.0 * q.rx * f; c.dy -= 2.0 * q.ry * f; c.dz -= 2.0 * q.rz * f; }
	}
	__device__ __forceinline__ static void reduce(Ctx& c, unsigned m) { c.dx = gsum(c.dx, m); c.dy = gsum(c.dy, m); c.dz = gsum(c.dz, m); }
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		S.dvelCollision.x[s.i] = c.dx; S.dvelCollision.y[s.i] = c.dy; S.dvelCollision.z[s.i] = c.dz;
	}
};
// second loop (:262-266): vel += dvel, pos += dvel dt
__device__ __forceinline__ void dpc_apply(const DevView& S, const KParams& K, int i, double4& p, double4& v) {
	const double dx = S.dvelCollision.x[i], dy = S.dvelCollision.y[i], dz = S.dvelCollision.z[i];
	v.x = __dadd_rn(v.x, dx); v.y = __dadd_rn(v.y, dy); v.z = __dadd_rn(v.z, dz);
	p.x = __dadd_rn(p.x, __dmul_rn(dx, K.p.timeStep)); p.y = __dadd_rn(p.y, __dmul_rn(dy, K.p.timeStep)); p.z = __dadd_rn(p.z, __dmul_rn(dz, K.p.timeStep));
}
struct BCollisionDPCTail {
	typedef void copy_vel_tag; typedef void copy_pos_tag; typedef void no_pair_tag;
	struct Ctx {};
	__device__ __forceinline__ static bool begin(Ctx&, const DevView&, const KParams& K, const Self& s) { return s.type == K.p.fluid; }
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx&, const Nb&, const KParams&, const Self&, const Pair&) {}
	__device__ __forceinline__ static void reduce(Ctx& , unsigned m) {  }
	__device__ __forceinline__ static void end(Ctx&, const DevView& S, const KParams& K, const Self& s) {
		double4 v = S.vel[s.i], p = s.pi;
		dpc_apply(S, K, s.i, p, v);
		S.vel2[s.i] = v; S.pos2[s.i] = p; track_disp(S, K, s.i, p);
	}
};
// fused head: the polygon-wall pass of particle i (closest point, mirror, wall weights; pmps_wall.cuh) rides in the
// collision launch — it needs only x_i, and the collision predicate does not look at the mirror (it was reset to
// "infinity" by the prediction, so d2 < dm2 holds trivially: the launch uses the POLY = false walker).
struct BWallClosestHead {
	struct Ctx {};
	__device__ __forceinline__ static bool begin(Ctx&, const DevView& S, const KParams& K, const Self& s) { if (s.lead) wall_closest_of(S, K, s.i, s.pi); return false; }
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx&, const Nb&, const KParams&, const Self&, const Pair&) {}
	__device__ __forceinline__ static void reduce(Ctx& , unsigned m) {  }
	__device__ __forceinline__ static void end(Ctx&, const DevView&, const KParams&, const Self&) {}
};

#include "pmps_bodies2.cuh"
