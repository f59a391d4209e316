// pmps_device.cuh — device helpers shared by every kernel: exact (non-contracted) predicate arithmetic, weight
// functions, the wall-reflection algebra and the Verlet-list walker that replaces the reference's neighbour-loop
// idiom (canonical instance reference src/MpsViscosity.cpp:22-78).
#pragma once
#include "pmps_internal.h"

// ---- exact arithmetic for everything a PREDICATE depends on ---------------------------------------------------
// The reference is built for baseline x86-64 (no FMA); neighbour sets, bucket ids and table look-ups must match it
// bit for bit, so these never contract.
__device__ __forceinline__ double sq3_exact(double x, double y, double z) {   // (x*x + y*y) + z*z, src/MpsParticle.cpp:35-45
	return __dadd_rn(__dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y)), __dmul_rn(z, z));
}
__device__ __forceinline__ double dot3_exact(double ax, double ay, double az, double bx, double by, double bz) {
	return __dadd_rn(__dadd_rn(__dmul_rn(ax, bx), __dmul_rn(ay, by)), __dmul_rn(az, bz));
}
__device__ __forceinline__ int bucket_coord(double x, double mn, double inv, double eps) {   // src/MpsBucket.cpp:420-425
	return (int)__dadd_rn(__dmul_rn(__dsub_rn(x, mn), inv), eps);
}

__device__ __forceinline__ int pk_type(const double4& p) { return (int)(uint32_t)(__double_as_longlong(p.w) & 0xffffffffll); }
__device__ __forceinline__ int pk_ptype(const double4& p) { return (int)(__double_as_longlong(p.w) >> 32); }
__device__ __forceinline__ double pk_make(int type, int ptype) { return __longlong_as_double(((long long)ptype << 32) | (long long)(uint32_t)type); }

// weight functions, src/MpsParticle.cpp:57-108
__device__ __forceinline__ double wgt(double r, double re, int type) {
	switch (type) {
		case 1: return re / r + r / re - 2.0;
		case 2: return re / r - r / re;
		case 3: return (1.0 - r / re) * (1.0 - r / re) * (1.0 - r / re);
		case 4: return (1.0 - r / re) * (1.0 - r / re);
		default: return re / r - 1.0;
	}
}

// unit normal particle->wall = half of (x_i - mirror_i), reflection R = I - 2 n n^T (e.g. src/MpsPndNeigh.cpp:114-136)
struct WallFrame { double n[3]; double R[9]; double len; };
__device__ __forceinline__ void wall_frame(WallFrame& f, double xi, double yi, double zi, double mx, double my, double mz, double eps, bool squaredTest) {
	f.n[0] = 0.5 * (xi - mx); f.n[1] = 0.5 * (yi - my); f.n[2] = 0.5 * (zi - mz);
	const double n2 = f.n[0] * f.n[0] + f.n[1] * f.n[1] + f.n[2] * f.n[2];
	f.len = sqrt(n2);
	if ((squaredTest ? n2 : f.len) > eps) { f.n[0] = f.n[0] / f.len; f.n[1] = f.n[1] / f.len; f.n[2] = f.n[2] / f.len; }
	else { f.n[0] = f.n[1] = f.n[2] = 0.0; }
	f.R[0] = 1.0 - 2.0 * f.n[0] * f.n[0]; f.R[1] = 0.0 - 2.0 * f.n[0] * f.n[1]; f.R[2] = 0.0 - 2.0 * f.n[0] * f.n[2];
	f.R[3] = 0.0 - 2.0 * f.n[1] * f.n[0]; f.R[4] = 1.0 - 2.0 * f.n[1] * f.n[1]; f.R[5] = 0.0 - 2.0 * f.n[1] * f.n[2];
	f.R[6] = 0.0 - 2.0 * f.n[2] * f.n[0]; f.R[7] = 0.0 - 2.0 * f.n[2] * f.n[1]; f.R[8] = 1.0 - 2.0 * f.n[2] * f.n[2];
}
#define MATVEC(R, a, b, c, k) ((R)[3 * (k)] * (a) + (R)[3 * (k) + 1] * (b) + (R)[3 * (k) + 2] * (c))

// mirror velocity of a near-wall particle (src/MpsViscosity.cpp:1729-1752 no-slip, :1565-1590 free-slip)
__device__ __forceinline__ void mirror_velocity(const KParams& K, int nearMeshType, const WallFrame& f, const double v[3], double vm[3]) {
	double vw[3] = { 0.0, 0.0, 0.0 };
	if (nearMeshType == 2) { vw[0] = K.p.uniformVelWall[0]; vw[1] = K.p.uniformVelWall[1]; vw[2] = K.p.uniformVelWall[2]; }
	const double dn = f.n[0] * vw[0] + f.n[1] * vw[1] + f.n[2] * vw[2];
	if (K.p.slipCondition == 0) {
		double t[3];
		for (int k = 0; k < 3; k++) t[k] = v[k] - 2.0 * (vw[k] - dn * f.n[k]);
		vm[0] = (-1.0 * t[0] + 0.0 * t[1] + 0.0 * t[2]); vm[1] = (0.0 * t[0] + -1.0 * t[1] + 0.0 * t[2]); vm[2] = (0.0 * t[0] + 0.0 * t[1] + -1.0 * t[2]);
	} else {
		for (int k = 0; k < 3; k++) vm[k] = MATVEC(f.R, v[0], v[1], v[2], k);
	}
}

__device__ __forceinline__ void atomic_max_f(float* addr, float v) {   // v >= 0: int order == float order
	if (v > *addr) atomicMax((int*)addr, __float_as_int(v));
}

// ---- the pair geometry handed to every stage body ---------------------------------------------------------------
struct Pair {
	int j; double4 pj;
	double rx, ry, rz, d2;      // r_ij = x_j - x_i
	double mx, my, mz, dm2;     // x_j - mirror_i (polygon walls only)
};

// Per-particle prologue data shared by fused bodies
struct Self {
	int i; double4 pi; int type, ptype;
	double mix, miy, miz;       // mirror position of i
};

// predicates: P(re) real neighbour, M(re) mirror neighbourhood
template <bool POLY> __device__ __forceinline__ bool pred_P(const Pair& q, double re2) { return q.d2 < re2 && (!POLY || q.d2 < q.dm2); }
__device__ __forceinline__ bool pred_M(const Pair& q, double re2) { return q.d2 < re2 && q.dm2 < re2 && q.d2 < q.dm2; }

// Walk the Verlet candidates of slot i.  Bodies... is a pack of stage bodies fused into one pass; each has
//   struct Ctx;  static bool begin(Ctx&, S, K, self)  -> false: body inactive for this i
//   static void pair<POLY>(Ctx&, S, K, self, const Pair&)
//   static void end(Ctx&, S, K, self)
template <class B, class = void> struct CopiesPos { static constexpr bool value = false; };
template <class B> struct CopiesPos<B, typename B::copy_pos_tag> { static constexpr bool value = true; };

template <class... Bodies> struct Fused;
template <> struct Fused<> {
	static constexpr bool kCopyPos = false;
	struct Ctx {};
	__device__ __forceinline__ static bool begin(Ctx&, const DevView&, const KParams&, const Self&) { return false; }
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx&, const DevView&, const KParams&, const Self&, const Pair&) {}
	__device__ __forceinline__ static void end(Ctx&, const DevView&, const KParams&, const Self&) {}
};
template <class B, class... Rest> struct Fused<B, Rest...> {
	static constexpr bool kCopyPos = CopiesPos<B>::value || Fused<Rest...>::kCopyPos;
	struct Ctx { typename B::Ctx c; bool on; typename Fused<Rest...>::Ctx r; };
	__device__ __forceinline__ static bool begin(Ctx& x, const DevView& S, const KParams& K, const Self& s) {
		x.on = B::begin(x.c, S, K, s);
		const bool o = Fused<Rest...>::begin(x.r, S, K, s);
		return x.on || o;
	}
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& x, const DevView& S, const KParams& K, const Self& s, const Pair& q) {
		if (x.on) B::template pair<POLY>(x.c, S, K, s, q);
		Fused<Rest...>::template pair<POLY>(x.r, S, K, s, q);
	}
	__device__ __forceinline__ static void end(Ctx& x, const DevView& S, const KParams& K, const Self& s) {
		if (x.on) B::end(x.c, S, K, s);
		Fused<Rest...>::end(x.r, S, K, s);
	}
};

template <bool POLY, class F>
__global__ void __launch_bounds__(128) k_walk(const __grid_constant__ DevView S, const __grid_constant__ KParams K) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= S.n) return;
	Self s; s.i = i; s.pi = S.pos[i]; s.type = pk_type(s.pi); s.ptype = pk_ptype(s.pi);
	if (F::kCopyPos) S.pos2[i] = s.pi;   // Jacobi position update: bodies overwrite pos2[i] in end()
	if (s.type == K.p.ghost) return;
	if (POLY) { s.mix = S.mirrorParticlePos.x[i]; s.miy = S.mirrorParticlePos.y[i]; s.miz = S.mirrorParticlePos.z[i]; }
	else { s.mix = s.miy = s.miz = 0.0; }
	typename F::Ctx ctx;
	if (!F::begin(ctx, S, K, s)) return;
	const uint32_t* L = S.list + (size_t)(i >> 5) * 32 * (size_t)S.listCap + (i & 31);
	const int cnt = S.listCnt[i];
	for (int k = 0; k < cnt; k++) {
		Pair q; q.j = (int)L[(size_t)k * 32];
		q.pj = S.pos[q.j];
		q.rx = q.pj.x - s.pi.x; q.ry = q.pj.y - s.pi.y; q.rz = q.pj.z - s.pi.z;
		q.d2 = sq3_exact(q.rx, q.ry, q.rz);
		if (POLY) { q.mx = q.pj.x - s.mix; q.my = q.pj.y - s.miy; q.mz = q.pj.z - s.miz; q.dm2 = sq3_exact(q.mx, q.my, q.mz); }
		else { q.mx = q.my = q.mz = 0.0; q.dm2 = 0.0; }
		F::template pair<POLY>(ctx, S, K, s, q);
	}
	F::end(ctx, S, K, s);
}
