// pmps_rheo.cuh — non-Newtonian (two-phase granular) rheology of fluidType 1.
//   MpsViscosity::calcVolumeFraction          src/MpsViscosity.cpp:104-190   -> BVolumeFraction (walker body)
//   MpsViscosity::calcViscosityInteractionVal src/MpsViscosity.cpp:194-700   -> k_rheo_maps_* (bed / water level per column),
//                                                                               BStrainRate (walker body), k_rheology (per particle)
//   MpsViscosity::calcWall{Slip,NoSlip}ViscosityInteractionVal src/MpsViscosity.cpp:703-1526 (polygon walls, 3D)
//                                                                            -> BWallStrainRate (wall-only walker body), k_wall_rheology
// The reference fills the column maps in a racy OpenMP loop; what it computes with one thread — the oracle's mode — is a
// running maximum in index order with a strict '>' (src/MpsViscosity.cpp:281-283): the first particle, in host index order,
// that reaches the column's maximum height provides PS.  Here: 64-bit atomic maxima of order-preserving keys, then the
// smallest host index among the particles AT the maximum.
#pragma once
#include "pmps_device.cuh"

__device__ __forceinline__ unsigned long long ord_key(double v) {   // monotone map double -> u64
	const unsigned long long b = (unsigned long long)__double_as_longlong(v);
	return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double ord_val(unsigned long long k) {
	const unsigned long long b = (k >> 63) ? (k & 0x7fffffffffffffffull) : ~k;
	return __longlong_as_double((long long)b);
}
__device__ __forceinline__ int rheo_column(const KParams& K, const double4& p, double& h) {
	// kx = int((x - Xmin)/aaL0) + 1, ky likewise in 3D (1 in 2D); arrays are [kx_max + 1][ky_max + 1]   (:236-238,:279,:293-294)
	const double aa = 2.0 * K.p.partDist;
	const int kx = (int)((p.x - K.p.domainMinX) / aa) + 1;
	int ky = 1;
	if (K.p.dim == 3) { ky = (int)((p.y - K.p.domainMinY) / aa) + 1; h = p.z; } else h = p.y;
	return kx * (K.rheoKy + 1) + ky;
}

static __global__ void k_rheo_maps_init(const __grid_constant__ DevView S, const __grid_constant__ KParams K, int resetMeuMax) {
	const int c = blockIdx.x * blockDim.x + threadIdx.x;
	if (c == 0 && resetMeuMax) S.meuMaxBits[0] = 0ull;   // (the wall variants leave CFLvisc alone: mi_max is taken in calcViscosityInteractionVal only, :637,:690)
	if (c >= (K.rheoKx + 1) * (K.rheoKy + 1)) return;
	const unsigned long long k0 = ord_key(K.p.dim == 3 ? K.p.domainMinZ : K.p.domainMinY);   // :253-255,:263-265
	S.rheoBL[c] = k0; S.rheoWL[c] = k0; S.rheoIdx[c] = 0x7fffffff;
}
static __global__ void k_rheo_maps_max(const __grid_constant__ DevView S, const __grid_constant__ KParams K) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= S.n) return;
	const double4 p = S.pos[i];
	if (pk_type(p) != K.p.fluid) return;
	double h; const int c = rheo_column(K, p, h);
	if (S.Cv[i] > 0.5) atomicMax(S.rheoBL + c, ord_key(h));
	if (pk_ptype(p) == 1) atomicMax(S.rheoWL + c, ord_key(h));
}
static __global__ void k_rheo_maps_arg(const __grid_constant__ DevView S, const __grid_constant__ KParams K) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= S.n) return;
	const double4 p = S.pos[i];
	if (pk_type(p) != K.p.fluid || !(S.Cv[i] > 0.5)) return;
	double h; const int c = rheo_column(K, p, h);
	const double h0 = K.p.dim == 3 ? K.p.domainMinZ : K.p.domainMinY;
	if (ord_key(h) == S.rheoBL[c] && h > h0) atomicMin(S.rheoIdx + c, S.hostIdx[i]);
}

// ------------------------------------------------------------------------------------------------------------------
struct BVolumeFraction {   // src/MpsViscosity.cpp:104-190: no mirror test, fluid neighbours only
	struct Ctx { double s1, s2; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		if (s.type != K.p.fluid) return false;
		c.s1 = c.s2 = 0.0; return true;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const Nb& S, const KParams& K, const Self& s, const Pair& q) {
		if (!(q.d2 < K.p.reS2) || q.tj != K.p.fluid) return;
		const double w = K.p.fractionMethod == 1 ? 1.0 : wgt_r(q.rinv, q.d2, K.p.reS, K.invReS, K.p.weightType);
		c.s1 += w; if (q.ptj >= 2) c.s2 += w;
	}
	__device__ __forceinline__ static void reduce(Ctx& c, unsigned m) { c.s1 = gsum(c.s1, m); c.s2 = gsum(c.s2, m); }
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		S.Cv[s.i] = c.s1 < K.p.epsilonZero ? 0.0 : c.s2 / c.s1;
	}
};

struct BStrainRate {   // src/MpsViscosity.cpp:313-437: velocity gradient over P(reS) (every neighbour type), smoothed pressure
	struct Ctx { double vx, vy, vz, u[9], sp, sw; double m[9]; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		if (s.type != K.p.fluid) return false;
		const double4 v = S.vel[s.i]; c.vx = v.x; c.vy = v.y; c.vz = v.z;
		for (int k = 0; k < 9; k++) c.u[k] = 0.0;
		c.sp = c.sw = 0.0;
		if (K.p.gradientCorrection) {
			c.m[0] = S.correcMatrixRow1.x[s.i]; c.m[1] = S.correcMatrixRow1.y[s.i]; c.m[2] = S.correcMatrixRow1.z[s.i];
			c.m[3] = S.correcMatrixRow2.x[s.i]; c.m[4] = S.correcMatrixRow2.y[s.i]; c.m[5] = S.correcMatrixRow2.z[s.i];
			c.m[6] = S.correcMatrixRow3.x[s.i]; c.m[7] = S.correcMatrixRow3.y[s.i]; c.m[8] = S.correcMatrixRow3.z[s.i];
		}
		return true;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const Nb& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_P<POLY>(q, K.p.reS2)) return;
		const double w = wgt_r(q.rinv, q.d2, K.p.reS, K.invReS, K.p.weightType);
		const double f = w * (q.rinv * q.rinv);
		const double dx = S.vx[q.j] - c.vx, dy = S.vy[q.j] - c.vy, dz = S.vz[q.j] - c.vz;
		double r0 = q.rx, r1 = q.ry, r2 = q.rz;
		if (K.p.gradientCorrection) {
			r0 = q.rx * c.m[0] + q.ry * c.m[1] + q.rz * c.m[2]; r1 = q.rx * c.m[3] + q.ry * c.m[4] + q.rz * c.m[5]; r2 = q.rx * c.m[6] + q.ry * c.m[7] + q.rz * c.m[8];
		}
		c.u[0] += dx * r0 * f; c.u[1] += dx * r1 * f; c.u[2] += dx * r2 * f;
		c.u[3] += dy * r0 * f; c.u[4] += dy * r1 * f; c.u[5] += dy * r2 * f;
		c.u[6] += dz * r0 * f; c.u[7] += dz * r1 * f; c.u[8] += dz * r2 * f;
		c.sp += S.press[q.j] * w; c.sw += w;
	}
	__device__ __forceinline__ static void reduce(Ctx& c, unsigned m) {
		for (int k = 0; k < 9; k++) c.u[k] = gsum(c.u[k], m);
		c.sp = gsum(c.sp, m); c.sw = gsum(c.sw, m);
	}
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		const int i = s.i; const double a = -K.p.coeffPressGrad;
		double ps = c.sw > K.p.epsilonZero ? c.sp / c.sw : S.press[i];
		if (ps < K.p.epsilonZero) ps = 0.0;
		S.p_smooth[i] = ps;
		const double Uxx = a * c.u[0], Uxy = a * c.u[1], Uxz = a * c.u[2], Uyx = a * c.u[3], Uyy = a * c.u[4], Uyz = a * c.u[5], Uzx = a * c.u[6], Uzy = a * c.u[7], Uzz = a * c.u[8];
		const double s11 = 0.5 * (Uxx + Uxx), s12 = 0.5 * (Uxy + Uyx), s13 = 0.5 * (Uxz + Uzx), s22 = 0.5 * (Uyy + Uyy), s23 = 0.5 * (Uyz + Uzy), s33 = 0.5 * (Uzz + Uzz);
		S.S11[i] = s11; S.S12[i] = s12; S.S13[i] = s13; S.S22[i] = s22; S.S23[i] = s23; S.S33[i] = s33;
		double II = 0.5 * (s11 * s11 + 2.0 * s12 * s12 + 2.0 * s13 * s13 + s22 * s22 + 2.0 * s23 * s23 + s33 * s33);
		if (II < K.p.epsilonZero || II * 0 != 0) II = 0.0;
		S.II[i] = II;
	}
};

// per particle: MEU / RHO / VF / Inertia / MEU_Y / p_rheo_new from the strain rate, the smoothed pressure and the column maps
// (src/MpsViscosity.cpp:456-640).  No particle reads another particle's output: the reference's serial loop is order-free.
static __global__ void __launch_bounds__(256) k_rheology(const __grid_constant__ DevView S, const __grid_constant__ KParams K) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	double meu = 0.0;
	if (i < S.n) {
		const double4 p = S.pos[i];
		if (pk_type(p) == K.p.fluid) {
			const PmpsParams& P = K.p;
			const int ptype = pk_ptype(p);
			const double eps = P.epsilonZero;
			const double gmod = sqrt(P.gravityX * P.gravityX + P.gravityY * P.gravityY + P.gravityZ * P.gravityZ);
			if (ptype == 1) S.MEU[i] = P.KNM_VS1 * P.DNS_FL1;
			else if (ptype == 2) {
				const double4 v = S.vel[i]; const double vel2 = v.x * v.x + v.y * v.y + v.z * v.z;
				const double Cv = S.Cv[i];
				double phi = (Cv - 0.25) * P.PHI_1 / (1.0 - 0.25), phi2 = (Cv - 0.25) * P.PHI_2 / (1.0 - 0.25);
				if (Cv <= 0.25) { phi = eps; phi2 = eps; }
				double h; const int c = rheo_column(K, p, h);
				const double BL = ord_val(S.rheoBL[c]), WL = ord_val(S.rheoWL[c]);
				const int ia = S.rheoIdx[c]; const double PS = ia == 0x7fffffff ? 0.0 : S.press[S.slotOfHost[ia]];
				double ps = S.p_smooth[i];
				double ns;
				if (P.dim == 2) {
					if (WL < BL) ns = ps;                // free fall (dry granular material)   :503-505
					else ns = ps - PS;                    // submerged                           :506-512
				} else {
					ns = (BL - h + P.partDist * 0.5) * (P.DNS_FL2 - P.DNS_FL1) * gmod - vel2 * (P.DNS_FL2 - P.DNS_FL1) * 0.5;   // :520
					if (ps - (WL - h) * P.DNS_FL1 * gmod < eps) { ps = (WL - h) * P.DNS_FL1 * gmod; S.p_smooth[i] = ps; }        // :522
					if (P.timeCurrent <= 1.0) ns = (1.0 - P.timeCurrent) * (ps - (WL - h) * P.DNS_FL1 * gmod) + P.timeCurrent * ns;   // :523
				}
				if (ns < 1.0 || Cv < 0.5) ns = 1.0;
				S.p_rheo_new[i] = ns;
				const double II = S.II[i]; const double modE = sqrt(II);
				const bool dry = WL < BL;
				const double inertia = dry ? modE * P.DG / sqrt(ns / P.DNS_SDT) : modE * P.DG / sqrt(ns / (P.DNS_FL1 * P.Cd));
				S.Inertia[i] = inertia;
				double VF = P.VF_max - (P.VF_max - P.VF_min) * inertia;
				if (VF < P.VF_min) VF = P.VF_min;
				S.VF[i] = VF;
				S.RHO[i] = P.DNS_SDT * VF + (1.0 - VF) * P.DNS_FL1;
				phi *= (VF / P.VF_max);
				const double sphi = sin(phi), cphi = cos(phi);
				const double alpha1 = 2.0 * sqrt(3.0) * sphi / (3.0 - sphi), beta1 = 2.0 * sqrt(3.0) * cphi / (3.0 - sphi);
				double ys = P.cohes * beta1 + ns * alpha1;
				if (ys < eps) ys = 0.0;
				const double meuY = ys / (2.0 * sqrt(modE * modE + eps));
				S.MEU_Y[i] = meuY;
				const double tdiff = tan(phi2) - tan(phi);
				const double den = dry ? P.I0 * sqrt(ns / P.DNS_FL2) + modE * P.DG : P.I0 * sqrt(ns / (P.DNS_FL1 * P.Cd)) + modE * P.DG;
				double meu0 = 0.5 * tdiff * ns * P.DG / den * modE / (sqrt(modE * modE + eps));
				if (modE <= eps || (meu0 * 0) != 0) meu0 = P.MEU0;
				const double viscMax = meu0 + ys * P.mm * 0.5;
				meu = meuY + meu0;
				if (II <= eps) meu = viscMax;
				S.MEU[i] = meu;
			}
			if (ptype >= 2) {
				const double Cv = S.Cv[i];
				S.RHO[i] = Cv > 0.5 ? P.DNS_FL2 : Cv * P.DNS_FL2 + (1.0 - Cv) * P.DNS_FL1;
			}
			if (ptype != 2) meu = 0.0;
		}
	}
	// mi_max: only the mixture particles take part (:637)
	unsigned long long b = (unsigned long long)__double_as_longlong(meu > 0.0 ? meu : 0.0);
	for (int o = 16; o; o >>= 1) { const unsigned long long t = __shfl_xor_sync(0xffffffffu, b, o); b = t > b ? t : b; }
	if ((threadIdx.x & 31) == 0 && b > S.meuMaxBits[0]) atomicMax(S.meuMaxBits, b);
}

// ---- polygon walls: the near-wall particles once more (src/MpsViscosity.cpp:703-1111 free-slip, :1112-1526 no-slip) -------------------
// The velocity gradient over the MIRROR neighbourhood M(reS), against the mirror particle's velocity, goes through the wall reflection
// and is ADDED to the strain-rate invariant; the pressure sum over the same neighbourhood is added to the smoothed pressure.
struct BWallStrainRate {
	typedef void wall_frame_tag;
	struct Ctx { double vm[3], u[9], sp; };
	__device__ __forceinline__ static bool begin(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		if (!s.nearWallFluid) return false;
		const double4 v4 = S.vel[s.i]; const double v[3] = { v4.x, v4.y, v4.z };
		WallFrame f; wall_frame_of(f, s, K.p.epsilonZero, true);
		mirror_velocity(K, S.nearMeshType[s.i], f, v, c.vm);
		for (int k = 0; k < 9; k++) c.u[k] = 0.0;
		c.sp = 0.0; return true;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& c, const Nb& S, const KParams& K, const Self& s, const Pair& q) {
		if (!pred_M(q, K.p.reS2)) return;
		const double ri = fast_rsqrt(q.dm2);
		const double w = wgt_r(ri, q.dm2, K.p.reS, K.invReS, K.p.weightType);
		const double f = w * (ri * ri);
		const double dx = S.vx[q.j] - c.vm[0], dy = S.vy[q.j] - c.vm[1], dz = S.vz[q.j] - c.vm[2];
		c.u[0] += dx * q.mx * f; c.u[1] += dx * q.my * f; c.u[2] += dx * q.mz * f;
		c.u[3] += dy * q.mx * f; c.u[4] += dy * q.my * f; c.u[5] += dy * q.mz * f;
		c.u[6] += dz * q.mx * f; c.u[7] += dz * q.my * f; c.u[8] += dz * q.mz * f;
		c.sp += S.press[q.j] * w;
	}
	__device__ __forceinline__ static void reduce(Ctx& c, unsigned m) {
		for (int k = 0; k < 9; k++) c.u[k] = gsum(c.u[k], m);
		c.sp = gsum(c.sp, m);
	}
	__device__ __forceinline__ static void end(Ctx& c, const DevView& S, const KParams& K, const Self& s) {
		const int i = s.i; const double cg = K.p.coeffPressGrad;
		WallFrame f; wall_frame_of(f, s, K.p.epsilonZero, true);
		const double* R = f.R; const double* u = c.u;
		double U[9];
		if (K.p.slipCondition == 1) {   // R (grad u) R, :868-890
			double a[9];
			for (int r = 0; r < 3; r++) for (int k = 0; k < 3; k++) a[3 * r + k] = -cg * (R[3 * r] * u[k] + R[3 * r + 1] * u[3 + k] + R[3 * r + 2] * u[6 + k]);
			for (int r = 0; r < 3; r++) for (int k = 0; k < 3; k++) U[3 * r + k] = a[3 * r] * R[k] + a[3 * r + 1] * R[3 + k] + a[3 * r + 2] * R[6 + k];
		} else {                        // R_inv (grad u) R = -(grad u) R, :1289-1300
			for (int r = 0; r < 3; r++) for (int k = 0; k < 3; k++) U[3 * r + k] = cg * (u[3 * r] * R[k] + u[3 * r + 1] * R[3 + k] + u[3 * r + 2] * R[6 + k]);
		}
		double ps = S.p_smooth[i] + c.sp / K.p.pndGradientZero;
		if (ps < K.p.epsilonZero) ps = 0.0;
		S.p_smooth[i] = ps;
		const double s11 = 0.5 * (U[0] + U[0]), s12 = 0.5 * (U[1] + U[3]), s13 = 0.5 * (U[2] + U[6]), s22 = 0.5 * (U[4] + U[4]), s23 = 0.5 * (U[5] + U[7]), s33 = 0.5 * (U[8] + U[8]);
		S.S11[i] = s11; S.S12[i] = s12; S.S13[i] = s13; S.S22[i] = s22; S.S23[i] = s23; S.S33[i] = s33;
		double II = S.II[i] + 0.5 * (s11 * s11 + 2.0 * s12 * s12 + 2.0 * s13 * s13 + s22 * s22 + 2.0 * s23 * s23 + s33 * s33);
		if (II < K.p.epsilonZero || II * 0 != 0) II = 0.0;
		S.II[i] = II;
	}
};

// per near-wall particle: the viscosity model in the form these two reference functions carry (exponential regularisation of the yield
// term, power-law index N; :1360-1515).  3D only, like the reference's code (its bed-level look-up reads z).
static __global__ void __launch_bounds__(256) k_wall_rheology(const __grid_constant__ DevView S, const __grid_constant__ KParams K) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= S.n) return;
	const double4 p = S.pos[i];
	if (pk_type(p) != K.p.fluid || !S.particleNearWall[i]) return;
	const PmpsParams& P = K.p;
	const int ptype = pk_ptype(p);
	const double eps = P.epsilonZero;
	const double gmod = sqrt(P.gravityX * P.gravityX + P.gravityY * P.gravityY + P.gravityZ * P.gravityZ);
	if (ptype == 1) S.MEU[i] = P.KNM_VS1 * P.DNS_FL1;
	else if (ptype == 2) {
		const double4 v = S.vel[i]; const double vel2 = v.x * v.x + v.y * v.y + v.z * v.z;
		const double Cv = S.Cv[i];
		double phi = (Cv - 0.25) * P.PHI_1 / (1.0 - 0.25), phi2 = (Cv - 0.25) * P.PHI_2 / (1.0 - 0.25);
		if (Cv <= 0.25) { phi = eps; phi2 = eps; }
		double h; const int c = rheo_column(K, p, h);
		const double BL = ord_val(S.rheoBL[c]), WL = ord_val(S.rheoWL[c]);
		double ps = S.p_smooth[i];
		double ns = (BL - h + P.partDist * 0.5) * (P.DNS_FL2 - P.DNS_FL1) * gmod - vel2 * (P.DNS_FL2 - P.DNS_FL1) * 0.5;
		if (ps < (WL - h) * P.DNS_FL1 * gmod) { ps = (WL - h) * P.DNS_FL1 * gmod; S.p_smooth[i] = ps; }
		if (P.timeCurrent <= 1.0) ns = (1.0 - P.timeCurrent) * (ps - (WL - h) * P.DNS_FL1 * gmod) + P.timeCurrent * ns;
		if (ns < 1.0 || Cv < 0.5) ns = 1.0;
		S.p_rheo_new[i] = ns;
		const double II = S.II[i], modE = sqrt(II);
		const double inertia = modE * P.DG / sqrt(ns / (P.DNS_FL1 * P.Cd));
		S.Inertia[i] = inertia;
		double VF = P.VF_max - (P.VF_max - P.VF_min) * inertia;
		if (VF < P.VF_min) VF = P.VF_min;
		S.VF[i] = VF;
		S.RHO[i] = P.DNS_SDT * VF + (1.0 - VF) * P.DNS_FL1;
		phi = phi * VF / P.VF_max;
		double ys = P.cohes * cos(phi) + ns * sin(phi);
		if (ys < (P.slipCondition == 1 ? eps : 0.0)) ys = 0.0;   // :1083 against :1497
		double viscMax = ys * P.mm * 0.5 + P.MEU0;
		const double meuY = II > eps ? ys * (1.0 - exp(-P.mm * modE)) * 0.5 / modE : viscMax;
		S.MEU_Y[i] = meuY;
		double meu0 = 0.5 * (tan(phi2) - tan(phi)) * ns * P.DG / (P.I0 * sqrt(ns / (P.DNS_FL1 * P.Cd)) + modE * P.DG);
		if (II <= eps || (meu0 * 0) != 0) meu0 = P.MEU0;
		viscMax = ys * P.mm * 0.5 + meu0;
		double meu = meuY + P.MEU0 * pow(4.0 * II, (P.powerLawN - 1.0) * 0.5);
		if (II == 0 || meu > viscMax) meu = viscMax;
		S.MEU[i] = meu;
	}
	if (ptype >= 2) {
		const double Cv = S.Cv[i];
		S.RHO[i] = Cv > 0.5 ? P.DNS_FL2 : Cv * P.DNS_FL2 + (1.0 - Cv) * P.DNS_FL1;
	}
}
