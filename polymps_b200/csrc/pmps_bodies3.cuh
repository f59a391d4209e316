// pmps_bodies3.cuh — stage bodies, continued: pressure gradient (+ wall terms, + fused correction tail), shifting,
// particle-wall extrapolations.
#pragma once

// ------------------------------------------------------------------------------------------------ pressure gradient
// phi of src/MpsPressure.cpp:1288-1300 times w / d2; invDD = 1/d2, invNi = 1/n_i
__device__ __forceinline__ double grad_phi_r(const KParams& K, double w, double Pj, double Pi, double Pmin, double ni, double invNi, double nj, double invDD) {
	switch (K.p.gradientType) {
		case 0: return w * ((Pj - Pmin) * invDD);
		case 1: return w * ((Pj + Pi) * invDD);
		case 2: return w * ((Pj + Pi - 2.0 * Pmin) * invDD);
		default: return (ni > K.p.epsilonZero && nj > K.p.epsilonZero) ? w * ((ni * Pj * fast_rcp(nj) + nj * Pi * invNi) * invDD) : w;
	}
}

// first pass of calcPressGradient / calcWallPressGradient for gradient types 0 and 2: Pmin over P(reS)
// (src/MpsPressure.cpp:1237-1262, :1405-1430).  Result parked in a scratch plane that is not a reference field.
struct BPressMin {
	struct Ctx { double pmin; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) { c.pmin = S.press[s.i]; return true; }
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const Nb& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_P<POLY>(q, K.p.reS2)) return;
		const double pj = S.press[q.j]; if (c.pmin > pj) c.pmin = pj;
	}
	__device__ __forceinline__ static void reduce(Ctx& c, unsigned m) { c.pmin = gmin(c.pmin, m); }
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) { S.scratch1[s.i] = c.pmin; }
};

// MpsPressure::calcPressGradient, src/MpsPressure.cpp:1217-1338
// PRED = true: MpsPressure::predictionPressGradient (src/MpsPressure.cpp:932-1028, relax_fact < 1): the same sum without the correction
// matrix, added to the acceleration with the factor (1 - relaxPress) instead of overwriting it with relaxPress
template <bool PRED> struct BPressGradientT {
	struct Ctx { double Pi, ni, invNi, Pmin, sx, sy, sz; double c1[3], c2[3], c3[3]; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		c.Pi = OWN(press, S.press[s.i]); c.ni = OWN(pndi, S.pndi[s.i]); c.invNi = c.ni > K.p.epsilonZero ? 1.0 / c.ni : 0.0;
		c.Pmin = (K.p.gradientType == 0 || K.p.gradientType == 2) ? S.scratch1[s.i] : c.Pi;
		if (!PRED && K.p.gradientCorrection) {
			c.c1[0] = S.correcMatrixRow1.x[s.i]; c.c1[1] = S.correcMatrixRow1.y[s.i]; c.c1[2] = S.correcMatrixRow1.z[s.i];
			c.c2[0] = S.correcMatrixRow2.x[s.i]; c.c2[1] = S.correcMatrixRow2.y[s.i]; c.c2[2] = S.correcMatrixRow2.z[s.i];
			c.c3[0] = S.correcMatrixRow3.x[s.i]; c.c3[1] = S.correcMatrixRow3.y[s.i]; c.c3[2] = S.correcMatrixRow3.z[s.i];
		}
		c.sx = c.sy = c.sz = 0.0; return true;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const Nb& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_P<POLY>(q, K.p.reS2)) return;
		const double w = grad_phi_r(K, wgt_r(q.rinv, q.d2, K.p.reS, K.invReS, K.p.weightType), S.press[q.j], c.Pi, c.Pmin, c.ni, c.invNi, S.pndi[q.j], q.rinv * q.rinv);
		if (PRED || !K.p.gradientCorrection) { c.sx += q.rx * w; c.sy += q.ry * w; c.sz += q.rz * w; }
		else {
			c.sx += (q.rx * c.c1[0] + q.ry * c.c1[1] + q.rz * c.c1[2]) * w; c.sy += (q.rx * c.c2[0] + q.ry * c.c2[1] + q.rz * c.c2[2]) * w;
			c.sz += (q.rx * c.c3[0] + q.ry * c.c3[1] + q.rz * c.c3[2]) * w;
		}
	}
	__device__ __forceinline__ static void reduce(Ctx& c, unsigned m) { c.sx = gsum(c.sx, m); c.sy = gsum(c.sy, m); c.sz = gsum(c.sz, m); }
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		const double rho = S.RHO[s.i];
		if (PRED) {
			S.acc.x[s.i] += (1.0 - K.p.relaxPress) * c.sx * K.p.coeffPressGrad / rho;
			S.acc.y[s.i] += (1.0 - K.p.relaxPress) * c.sy * K.p.coeffPressGrad / rho;
			S.acc.z[s.i] += (1.0 - K.p.relaxPress) * c.sz * K.p.coeffPressGrad / rho;
			return;
		}
		S.acc.x[s.i] = K.p.relaxPress * c.sx * K.p.coeffPressGrad / rho;
		S.acc.y[s.i] = K.p.relaxPress * c.sy * K.p.coeffPressGrad / rho;
		S.acc.z[s.i] = K.p.relaxPress * c.sz * K.p.coeffPressGrad / rho;
	}
};
typedef BPressGradientT<false> BPressGradient;
typedef BPressGradientT<true> BPredPressGradient;

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
// PRED = true: predictionWallPressGradient (src/MpsPressure.cpp:1031-1214): factor (1 - relaxPress), no forceWall
template <bool PRED> struct BWallPressGradientT {
	typedef void wall_frame_tag;
	struct Ctx { double Pi, ni, invNi, Pmin, sx, sy, sz; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		if (!s.nearWallFluid) return false;
		c.Pi = S.press[s.i]; c.ni = S.pndi[s.i]; c.invNi = c.ni > K.p.epsilonZero ? 1.0 / c.ni : 0.0;
		c.Pmin = (K.p.gradientType == 0 || K.p.gradientType == 2) ? S.scratch1[s.i] : c.Pi;
		c.sx = c.sy = c.sz = 0.0; return true;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const Nb& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_M(q, K.p.reS2)) return;
		const double mi = fast_rsqrt(q.dm2);
		const double w = grad_phi_r(K, wgt_r(mi, q.dm2, K.p.reS, K.invReS, K.p.weightType), S.press[q.j], c.Pi, c.Pmin, c.ni, c.invNi, S.pndi[q.j], mi * mi);
		c.sx += q.mx * w; c.sy += q.my * w; c.sz += q.mz * w;
	}
	__device__ __forceinline__ static void reduce(Ctx& c, unsigned m) { c.sx = gsum(c.sx, m); c.sy = gsum(c.sy, m); c.sz = gsum(c.sz, m); }
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		const double ex = s.pi.x - s.mix, ey = s.pi.y - s.miy, ez = s.pi.z - s.miz; const double dd = sq3_exact(ex, ey, ez);
		if (dd < K.p.reS2) {
			const double w = grad_phi_r(K, wgt(sqrt(dd), K.p.reS, K.p.weightType), c.Pi, c.Pi, c.Pmin, c.ni, c.invNi, c.ni, 1.0 / dd);
			c.sx += ex * w; c.sy += ey * w; c.sz += ez * w;
		}
		WallFrame f; wall_frame_of(f, s, K.p.epsilonZero, false);
		const double rho = S.RHO[s.i];
		double fr[3] = { 0.0, 0.0, 0.0 };
		const double r = f.len;
		if (r < K.p.reRepulsiveForce && r > K.p.epsilonZero) {
			const double velMax = sqrt(__longlong_as_double((long long)S.velMax2Bits[1]));
			const double vcap = velMax < K.p.expectMaxVelocity ? velMax : K.p.expectMaxVelocity;
			const double qq = repulsion(K, r, rho, vcap * vcap);
			fr[0] = -qq * f.n[0]; fr[1] = -qq * f.n[1]; fr[2] = -qq * f.n[2];
		}
		double* a[3] = { S.acc.x, S.acc.y, S.acc.z }; double* fw[3] = { S.forceWall.x, S.forceWall.y, S.forceWall.z };
		for (int k = 0; k < 3; k++) {
			const double g = ((PRED ? 1.0 - K.p.relaxPress : K.p.relaxPress) * MATVEC(f.R, c.sx, c.sy, c.sz, k) * K.p.coeffPressGrad - fr[k]);
			a[k][s.i] += g / rho;
			if (!PRED) fw[k][s.i] += -g * K.vol;
		}
	}
};
typedef BWallPressGradientT<false> BWallPressGradient;
typedef BWallPressGradientT<true> BPredWallPressGradient;

// fused tail of the gradient pass: correctionVelocityPosition (src/MpsParticleVelPos.cpp:53-97).  Reads the acceleration
// the bodies in front of it stored, writes the new position / velocity to the alternate buffers, feeds velMax.
struct BCorrectTail {
	typedef void copy_pos_tag; typedef void copy_vel_tag; typedef void no_pair_tag;
	struct Ctx {};
	__device__ __forceinline__ static bool begin(Ctx&, const DevView&, const KParams& K, const Self& s) { return s.type == K.p.fluid; }
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx&, const Nb&, const KParams&, const Self&, const Pair&) {}
	__device__ __forceinline__ static void reduce(Ctx& , unsigned m) {  }
	__device__ __forceinline__ static void end(Ctx&, const DevView& S, const KParams& K, const Self& s) {
		const int i = s.i;
		double4 p = s.pi, v = S.vel[i];
		integrate_correct(p, v, S.acc.x[i], S.acc.y[i], S.acc.z[i], K.p.timeStep);
		S.pos2[i] = p; S.vel2[i] = v;
		S.acc.x[i] = 0.0; S.acc.y[i] = 0.0; S.acc.z[i] = 0.0;
		track_disp(S, K, i, p);
		const unsigned long long b = (unsigned long long)__double_as_longlong(v.x * v.x + v.y * v.y + v.z * v.z);
		if (b > S.velMax2Bits[0] && owned_slot(S, i)) atomicMax(S.velMax2Bits, b);
	}
};

// ------------------------------------------------------------------------------------------------ shifting (type 2)
// MpsShifting::calcConcAndConcGradient, src/MpsShifting.cpp:432-575 — first loop: concentration
struct BConcentration {
	struct Ctx { double c; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) { c.c = 0.0; return true; }
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const Nb& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_P<POLY>(q, K.p.reS2)) return;
		c.c += wgt_r(q.rinv, q.d2, K.p.reS, K.invReS, K.p.weightType);
	}
	__device__ __forceinline__ static void reduce(Ctx& c, unsigned m) { c.c = gsum(c.c, m); }
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		S.concentration[s.i] = (c.c + S.pndWallContribution[s.i]) / K.p.pndSmallZero;
	}
};

// second loop: concentration gradient and the shift.  Jacobi semantics: positions are READ from S.pos (the snapshot)
// and the shifted position is WRITTEN to S.pos2; the reference's in-place update is a race for >1 thread
// (src/MpsShifting.cpp:478-536).
struct BGradConcentration {
	typedef void copy_pos_tag;
	struct Ctx { double ci, gx, gy, gz; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		if (s.type != K.p.fluid) return false;
		c.ci = OWN(concentration, S.concentration[s.i]); c.gx = c.gy = c.gz = 0.0; return true;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const Nb& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_P<POLY>(q, K.p.reS2)) return;
		const double f = (c.ci + S.concentration[q.j]) * wgt_r(q.rinv, q.d2, K.p.reS, K.invReS, K.p.weightType) * (q.rinv * q.rinv);
		c.gx += q.rx * f; c.gy += q.ry * f; c.gz += q.rz * f;
	}
	__device__ __forceinline__ static void reduce(Ctx& c, unsigned m) { c.gx = gsum(c.gx, m); c.gy = gsum(c.gy, m); c.gz = gsum(c.gz, m); }
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		const double g0 = c.gx * -K.p.coeffPressGrad, g1 = c.gy * -K.p.coeffPressGrad, g2 = c.gz * -K.p.coeffPressGrad;
		S.gradConcentration.x[s.i] = g0; S.gradConcentration.y[s.i] = g1; S.gradConcentration.z[s.i] = g2;
		if (S.particleBC[s.i] == K.p.inner) {
			double4 p = s.pi;
			p.x = __dsub_rn(p.x, __dmul_rn(K.p.coeffShifting2, g0)); p.y = __dsub_rn(p.y, __dmul_rn(K.p.coeffShifting2, g1)); p.z = __dsub_rn(p.z, __dmul_rn(K.p.coeffShifting2, g2));
			S.pos2[s.i] = p;
			track_disp(S, K, s.i, p); track_shift(S, s.i, p.x - s.pi.x, p.y - s.pi.y, p.z - s.pi.z);
		}
	}
};

// MpsShifting::calcWallConcAndConcGradient, src/MpsShifting.cpp:578-691 (same read-snapshot / write-alternate rule)
struct BWallGradConcentration {
	typedef void copy_pos_tag; typedef void wall_frame_tag;
	struct Ctx { double ci, sx, sy, sz; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		if (!s.nearWallFluid) return false;
		c.ci = S.concentration[s.i];
		c.sx = c.sy = c.sz = 0.0; return true;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const Nb& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_M(q, K.p.reS2)) return;
		const double mi = fast_rsqrt(q.dm2);
		const double f = (c.ci + S.concentration[q.j]) * wgt_r(mi, q.dm2, K.p.reS, K.invReS, K.p.weightType) * (mi * mi);
		c.sx += q.mx * f; c.sy += q.my * f; c.sz += q.mz * f;
	}
	__device__ __forceinline__ static void reduce(Ctx& c, unsigned m) { c.sx = gsum(c.sx, m); c.sy = gsum(c.sy, m); c.sz = gsum(c.sz, m); }
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		const double ex = s.pi.x - s.mix, ey = s.pi.y - s.miy, ez = s.pi.z - s.miz; const double dd = sq3_exact(ex, ey, ez);
		if (dd < K.p.reS2) {
			const double w = wgt(sqrt(dd), K.p.reS, K.p.weightType);
			c.sx += (c.ci + c.ci) * ex * w / dd; c.sy += (c.ci + c.ci) * ey * w / dd; c.sz += (c.ci + c.ci) * ez * w / dd;
		}
		WallFrame f; wall_frame_of(f, s, K.p.epsilonZero, false);
		double* g[3] = { S.gradConcentration.x, S.gradConcentration.y, S.gradConcentration.z };
		double p[3] = { s.pi.x, s.pi.y, s.pi.z };
		const bool inner = S.particleBC[s.i] == K.p.inner;
		for (int k = 0; k < 3; k++) {
			const double gw = -MATVEC(f.R, c.sx, c.sy, c.sz, k) * K.p.coeffPressGrad;
			g[k][s.i] += gw;
			if (inner) p[k] = __dsub_rn(p[k], __dmul_rn(K.p.coeffShifting2, gw));
		}
		if (inner) {
			double4 pn = s.pi; pn.x = p[0]; pn.y = p[1]; pn.z = p[2]; S.pos2[s.i] = pn;
			track_disp(S, K, s.i, pn); track_shift(S, s.i, pn.x - s.pi.x, pn.y - s.pi.y, pn.z - s.pi.z);
		}
	}
};

// ------------------------------------------------------------------------------------------------ particle walls
// MpsPressure::extrapolatePressParticlesWallDummy, src/MpsPressure.cpp:792-846.  Only non-fluid i are written and
// only fluid j are read, so the in-place update has no read/write hazard.
struct BExtrapolatePress {
	typedef void sparse_tag;
	struct Ctx { double ni, pr; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		c.ni = c.pr = 0.0;
		return s.type == K.p.dummyWall || (K.p.mpsType == 1 && s.type == K.p.wall);
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const Nb& S, const KParams& K, const Self& s, const Pair& q) {
		if (!(q.d2 < K.p.reS2 && q.tj == K.p.fluid)) return;
		const double w = wgt_r(q.rinv, q.d2, K.p.reS, K.invReS, K.p.weightType); c.ni += w;
		c.pr += (S.press[q.j] - S.RHO[q.j] * (K.p.gravityX * q.rx + K.p.gravityY * q.ry + K.p.gravityZ * q.rz)) * w;
	}
	__device__ __forceinline__ static void reduce(Ctx& c, unsigned m) { c.ni = gsum(c.ni, m); c.pr = gsum(c.pr, m); }
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		if (c.pr < 0.0) return;
		S.press[s.i] = c.ni > 0 ? c.pr / c.ni : c.pr;
	}
};

// MpsParticleVelPos::updateVelocityParticlesWallDummy, src/MpsParticleVelPos.cpp:100-158 (dummy i written, fluid j read)
struct BWallDummyVel {
	typedef void sparse_tag;
	struct Ctx { double ni, ux, uy, uz; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		c.ni = c.ux = c.uy = c.uz = 0.0; return s.type == K.p.dummyWall;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const Nb& S, const KParams& K, const Self& s, const Pair& q) {
		if (!(q.d2 < K.p.reL2 && q.tj == K.p.fluid)) return;
		const double w = wgt_r(q.rinv, q.d2, K.p.reL, K.invReL, K.p.weightType); const double vjx = S.vx[q.j], vjy = S.vy[q.j], vjz = S.vz[q.j];
		c.ni += w; c.ux += vjx * w; c.uy += vjy * w; c.uz += vjz * w;
	}
	__device__ __forceinline__ static void reduce(Ctx& c, unsigned m) { c.ni = gsum(c.ni, m); c.ux = gsum(c.ux, m); c.uy = gsum(c.uy, m); c.uz = gsum(c.uz, m); }
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		double4 v = S.vel[s.i];
		if (c.ni > K.p.epsilonZero) { v.x = 2.0 * 0.0 - c.ux / c.ni; v.y = 2.0 * 0.0 - c.uy / c.ni; v.z = 2.0 * 0.0 - c.uz / c.ni; }
		else { v.x = 2.0 * 0.0 - c.ux; v.y = 2.0 * 0.0 - c.uy; v.z = 2.0 * 0.0 - c.uz; }
		S.vel[s.i] = v;
	}
};

// MpsShifting::calcShifting (shifting type 1: velocity adjustment), src/MpsShifting.cpp:23-101 with delWeight of src/MpsParticle.cpp:94-108.
// Jacobi semantics like the type-2 pass: neighbours' velocities are READ from S.vel, the adjusted velocity is WRITTEN to S.vel2.
__device__ __forceinline__ double del_wgt(double r, double re, int type) {
	switch (type) {
		case 1: return -re / (r * r) + 1.0 / re;
		case 2: return -re / (r * r) - 1.0 / re;
		case 3: return -3.0 / re * (1.0 - r / re) * (1.0 - r / re);
		case 4: return -2.0 / re * (1.0 - r / re);
		default: return -re / (r * r);
	}
}
struct BShiftVel {
	typedef void copy_vel_tag;
	struct Ctx { double vx, vy, vz, dx, dy, dz; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		if (s.type != K.p.fluid) return false;
		const double4 v = S.vel[s.i]; c.vx = v.x; c.vy = v.y; c.vz = v.z; c.dx = c.dy = c.dz = 0.0; return true;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const Nb& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_P<POLY>(q, K.p.reS2)) return;
		const double dw = del_wgt(sqrt(q.d2), K.p.reS, K.p.weightType);
		c.dx += dw * (S.vx[q.j] - c.vx); c.dy += dw * (S.vy[q.j] - c.vy); c.dz += dw * (S.vz[q.j] - c.vz);
	}
	__device__ __forceinline__ static void reduce(Ctx& c, unsigned m) { c.dx = gsum(c.dx, m); c.dy = gsum(c.dy, m); c.dz = gsum(c.dz, m); }
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		double4 v = S.vel[s.i];
		if (S.particleBC[s.i] == K.p.inner) { v.x -= K.p.coeffShifting1 * c.dx; v.y -= K.p.coeffShifting1 * c.dy; v.z -= K.p.coeffShifting1 * c.dz; }
		S.vel2[s.i] = v;
	}
};
// ------------------------------------------------------------------------------------------------ ARC free-surface detection
// MpsShifting::calcNormalParticles, src/MpsShifting.cpp:106-174: normal_i = -(1/n0S) sum_j r_ij w_S / |r_ij| over P(reS), every particle
struct BNormal {
	struct Ctx { double dx, dy, dz; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) { c.dx = c.dy = c.dz = 0.0; return true; }
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const Nb& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_P<POLY>(q, K.p.reS2)) return;
		const double f = wgt_r(q.rinv, q.d2, K.p.reS, K.invReS, K.p.weightType) * q.rinv;
		c.dx += q.rx * f; c.dy += q.ry * f; c.dz += q.rz * f;
	}
	__device__ __forceinline__ static void reduce(Ctx& c, unsigned m) { c.dx = gsum(c.dx, m); c.dy = gsum(c.dy, m); c.dz = gsum(c.dz, m); }
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		S.normal.x[s.i] = -c.dx / K.p.pndSmallZero; S.normal.y[s.i] = -c.dy / K.p.pndSmallZero; S.normal.z[s.i] = -c.dz / K.p.pndSmallZero;
	}
};
// MpsShifting::calcWallNormalParticles, src/MpsShifting.cpp:319-429: the mirror neighbourhood's part (+ the particle itself), reflected back
struct BWallNormal {
	typedef void wall_frame_tag;
	struct Ctx { double dx, dy, dz; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		if (!s.nearWallFluid) return false;
		c.dx = c.dy = c.dz = 0.0; return true;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const Nb& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_M(q, K.p.reS2)) return;
		const double ri = fast_rsqrt(q.dm2); const double f = wgt_r(ri, q.dm2, K.p.reS, K.invReS, K.p.weightType) * ri;
		c.dx += q.mx * f; c.dy += q.my * f; c.dz += q.mz * f;
	}
	__device__ __forceinline__ static void reduce(Ctx& c, unsigned m) { c.dx = gsum(c.dx, m); c.dy = gsum(c.dy, m); c.dz = gsum(c.dz, m); }
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		const double ex = s.pi.x - s.mix, ey = s.pi.y - s.miy, ez = s.pi.z - s.miz; const double dd = sq3_exact(ex, ey, ez);
		if (dd < K.p.reS2) { const double dst = sqrt(dd); const double f = wgt(dst, K.p.reS, K.p.weightType) / dst; c.dx += ex * f; c.dy += ey * f; c.dz += ez * f; }
		WallFrame f; wall_frame_of(f, s, K.p.epsilonZero, false);
		S.normal.x[s.i] += -MATVEC(f.R, c.dx, c.dy, c.dz, 0) / K.p.pndSmallZero;
		S.normal.y[s.i] += -MATVEC(f.R, c.dx, c.dy, c.dz, 1) / K.p.pndSmallZero;
		S.normal.z[s.i] += -MATVEC(f.R, c.dx, c.dy, c.dz, 2) / K.p.pndSmallZero;
	}
};
// MpsBoundaryCondition::updateParticleBC with PND_ARC, src/MpsBoundaryCondition.cpp:66-88,205-390.  A particle that is not inner by
// its density, neighbour count and normal length is INNER when some neighbour lies in the arc in front of it (within 1.33 lo of the
// point one particle distance along the normal, or — closer than sqrt(2) 1.33 lo — inside the cone of half-angle thetaArc around the
// normal; the same test from the mirror particle for a particle near a polygon wall), SURFACE when it has neighbours but none there,
// and keeps the density-based flag when it has none.  The reference leaves its loops at the first hit: an OR, independent of order.
struct BArcBC {
	struct Ctx { double nx, ny, nz, qx, qy, qz; int bc0, found, seen; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		if (s.type == K.p.dummyWall) { if (s.lead) S.particleBC[s.i] = K.p.other; return false; }
		c.bc0 = S.pndSmall[s.i] < K.p.betaPnd ? K.p.surface : K.p.inner; c.found = 0; c.seen = 0;
		double nx = S.normal.x[s.i], ny = S.normal.y[s.i], nz = S.normal.z[s.i];
		const double n2 = nx * nx + ny * ny + nz * nz;
		if (S.pndi[s.i] >= K.p.betaPnd && S.numNeigh[s.i] >= K.p.betaNeigh && n2 <= K.p.normThreshold2) { if (s.lead) S.particleBC[s.i] = K.p.inner; return false; }
		if (n2 > K.p.epsilonZero) { const double nm = sqrt(n2); nx /= nm; ny /= nm; nz /= nm; } else { nx = ny = nz = 0.0; }
		c.nx = nx; c.ny = ny; c.nz = nz;
		// normal of the mirror particle: the reflection about the wall plane through the closest point (R = I - 2 n_w n_w^T)
		double wx = 0.5 * (s.pi.x - s.mix), wy = 0.5 * (s.pi.y - s.miy), wz = 0.5 * (s.pi.z - s.miz);
		const double w2 = wx * wx + wy * wy + wz * wz;
		if (w2 > K.p.epsilonZero) { const double wm = sqrt(w2); wx /= wm; wy /= wm; wz /= wm; } else { wx = wy = wz = 0.0; }
		const double dn = wx * nx + wy * ny + wz * nz;
		double qx = nx - 2.0 * wx * dn, qy = ny - 2.0 * wy * dn, qz = nz - 2.0 * wz * dn;
		const double q2 = qx * qx + qy * qy + qz * qz;
		if (q2 > K.p.epsilonZero) { const double qm = sqrt(q2); qx /= qm; qy /= qm; qz /= qm; } else { qx = qy = qz = 0.0; }
		c.qx = qx; c.qy = qy; c.qz = qz;
		return true;
	}
	__device__ __forceinline__ static bool in_arc(const KParams& K, double rx, double ry, double rz, double d2, double nx, double ny, double nz) {
		if (d2 >= K.p.dstThreshold2) {   // |x_j - (x_i + lo n)|^2 < hThreshold2
			const double ax = rx - K.p.partDist * nx, ay = ry - K.p.partDist * ny, az = rz - K.p.partDist * nz;
			return ax * ax + ay * ay + az * az < K.p.hThreshold2;
		}
		return acos((rx * nx + ry * ny + rz * nz) / sqrt(d2)) < K.p.thetaArc;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const Nb& S, const KParams& K, const Self& s, const Pair& q) {
		if (pred_P<POLY>(q, K.p.reS2)) { c.seen = 1; if (in_arc(K, q.rx, q.ry, q.rz, q.d2, c.nx, c.ny, c.nz)) c.found = 1; }
		if (POLY && pred_M(q, K.p.reS2)) { c.seen = 1; if (in_arc(K, q.mx, q.my, q.mz, q.dm2, c.qx, c.qy, c.qz)) c.found = 1; }
	}
	__device__ __forceinline__ static void reduce(Ctx& c, unsigned m) { c.found = gsum(c.found, m); c.seen = gsum(c.seen, m); }
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		S.particleBC[s.i] = c.found ? K.p.inner : c.seen ? K.p.surface : c.bc0;
	}
};

#include "pmps_rheo.cuh"
