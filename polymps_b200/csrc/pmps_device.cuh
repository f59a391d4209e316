// pmps_device.cuh — device helpers shared by every kernel: exact (non-contracted) predicate arithmetic, fast
// reciprocal / reciprocal-square-root for the weighted sums, the wall-reflection algebra and the Verlet-list walker
// that replaces the reference's neighbour-loop idiom (canonical instance src/MpsViscosity.cpp:22-78).
//
// Arithmetic policy:
//   * everything a PREDICATE depends on — squared distances, bucket ids, mirror positions, table look-ups, the
//     position/velocity integrators — is written with __dmul_rn/__dadd_rn/... in the reference's evaluation order, so
//     neighbour sets, bucket assignment and integer outputs match the reference bit for bit;
//   * the weighted SUMS may contract to FMA and use 1-ulp reciprocals instead of IEEE division: their results agree
//     with the reference to ~1e-15 relative, five orders inside the 1e-10 contract, at a third of the FP64 issue slots.
#pragma once
#include "pmps_internal.h"

// ---- exact arithmetic -------------------------------------------------------------------------------------------
__device__ __forceinline__ double sq3_exact(double x, double y, double z) {   // (x*x + y*y) + z*z, src/MpsParticle.cpp:35-45
	return __dadd_rn(__dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y)), __dmul_rn(z, z));
}
__device__ __forceinline__ double dot3_exact(double ax, double ay, double az, double bx, double by, double bz) {
	return __dadd_rn(__dadd_rn(__dmul_rn(ax, bx), __dmul_rn(ay, by)), __dmul_rn(az, bz));
}
__device__ __forceinline__ int bucket_coord(double x, double mn, double inv, double eps) {   // src/MpsBucket.cpp:420-425
	return (int)__dadd_rn(__dmul_rn(__dsub_rn(x, mn), inv), eps);
}
// the bucket updateParticlesID threads a particle on (src/MpsBucket.cpp:310-313 in 2D — with epsilonZero —, :333-336 in 3D — without)
__device__ __forceinline__ int ref_bucket_insert(const KParams& K, const double4& p) {
	const PmpsParams& P = K.p;
	if (P.dim == 2) return bucket_coord(p.y, P.domainMinY, P.invBucketSide, P.epsilonZero) * P.numBucketsX + bucket_coord(p.x, P.domainMinX, P.invBucketSide, P.epsilonZero);
	const int ix = (int)__dmul_rn(__dsub_rn(p.x, P.domainMinX), P.invBucketSide), iy = (int)__dmul_rn(__dsub_rn(p.y, P.domainMinY), P.invBucketSide),
	          iz = (int)__dmul_rn(__dsub_rn(p.z, P.domainMinZ), P.invBucketSide);
	return (iz * P.numBucketsY + iy) * P.numBucketsX + ix;
}
__device__ __forceinline__ double madd_exact(double a, double b, double c) { return __dadd_rn(c, __dmul_rn(a, b)); }   // c + a*b, two roundings

// ---- fast 1-ulp reciprocals (MUFU seed, relative error <= 2^-20, + ONE third-order step; no special-case handling: arguments
// are finite, > 0).  With e = 1 - x y0^2 the exact correction is y0 (1 - e)^-1/2 = y0 (1 + e/2 + 3 e^2/8 + O(e^3)): the O(e^3) left
// over is < 1e-18, i.e. the result carries only the rounding of its last two operations (checked against a 64-bit-mantissa
// evaluation over the whole seed-error range: 1.4e-16).  5 FP64 operations instead of the 7 of two Newton steps.
__device__ __forceinline__ double fast_rsqrt(double x) {
	double y; asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
#ifdef PMPS_NEWTON2
	const double h = 0.5 * x;
	double e = fma(-h * y, y, 0.5); y = fma(y, e, y);
	e = fma(-h * y, y, 0.5); y = fma(y, e, y);
	return y;
#else
	const double e = fma(-(x * y), y, 1.0);
	return fma(y, fma(0.375, e, 0.5) * e, y);
#endif
}
__device__ __forceinline__ double fast_rcp(double x) {   // e = 1 - x y0: 1/x = y0 (1 + e + e^2 + O(e^3))
	double y; asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
#ifdef PMPS_NEWTON2
	double e = fma(-x, y, 1.0); y = fma(y, e, y);
	e = fma(-x, y, 1.0); y = fma(y, e, y);
	return y;
#else
	const double e = fma(-x, y, 1.0);
	return fma(y, fma(e, e, e), y);
#endif
}

__device__ __forceinline__ int pk_type(const double4& p) { return (int)(uint32_t)(__double_as_longlong(p.w) & 0xffffffffll); }
__device__ __forceinline__ int pk_ptype(const double4& p) { return (int)(__double_as_longlong(p.w) >> 32); }
__device__ __forceinline__ double pk_make(int type, int ptype) { return __longlong_as_double(((long long)ptype << 32) | (long long)(uint32_t)type); }

// weight functions, src/MpsParticle.cpp:57-108, from the reciprocal distance: r = d2 * rinv
__device__ __forceinline__ double wgt_r(double rinv, double d2, double re, double invRe, int type) {
	if (type == 0) return re * rinv - 1.0;
	const double q = d2 * rinv * invRe;   // r / re
	switch (type) {
		case 1: return re * rinv + q - 2.0;
		case 2: return re * rinv - q;
		case 3: return (1.0 - q) * (1.0 - q) * (1.0 - q);
		default: return (1.0 - q) * (1.0 - q);
	}
}
// IEEE form (used by the rare self / wall terms and the PPE assembly)
__device__ __forceinline__ double wgt(double r, double re, int type) {
	switch (type) {
		case 1: return re / r + r / re - 2.0;
		case 2: return re / r - r / re;
		case 3: return (1.0 - r / re) * (1.0 - r / re) * (1.0 - r / re);
		case 4: return (1.0 - r / re) * (1.0 - r / re);
		default: return re / r - 1.0;
	}
}

#define MATVEC(R, a, b, c, k) ((R)[3 * (k)] * (a) + (R)[3 * (k) + 1] * (b) + (R)[3 * (k) + 2] * (c))

__device__ __forceinline__ void atomic_max_f(float* addr, float v) {   // v >= 0: int order == float order
	if (v > *addr) atomicMax((int*)addr, __float_as_int(v));
}
// slab decomposition: is slot i a particle this rank owns (not a halo copy)?  Always true on a single GPU.
__device__ __forceinline__ bool owned_slot(const DevView& S, int i) { return S.nOwn == 0x7fffffff || S.hostIdx[i] < S.nOwn; }
// displacement of slot i at its new position p: from the hit-list positions (owned slots when they move; halo copies when the owner's
// position arrives, pmps_slab.cu) and from the bucket-build positions (owned particles; their owner decides about rebuilds)
__device__ __forceinline__ void track_hit_disp(const DevView& S, const KParams& K, int i, const double4& p) {
	const float4 b = S.hpos[i];
	const float dx = (float)(p.x - K.p.domainMinX) - b.x, dy = (float)(p.y - K.p.domainMinY) - b.y, dz = (float)(p.z - K.p.domainMinZ) - b.z;
	atomic_max_f(S.hitMove, dx * dx + dy * dy + dz * dz);
}
__device__ __forceinline__ void track_disp(const DevView& S, const KParams& K, int i, const double4& p) {
	// a halo copy's own integration is garbage (it lacks half its neighbours) and is replaced by the owner's position before any pass
	// reads it (every phase that moves particles ends with a position refresh): only the TRUE position counts, at the unpack
	if (!owned_slot(S, i)) return;
	track_hit_disp(S, K, i, p);
	const float4 b = S.bpos[i];
	const float dx = (float)(p.x - K.p.domainMinX) - b.x, dy = (float)(p.y - K.p.domainMinY) - b.y, dz = (float)(p.z - K.p.domainMinZ) - b.z;
	atomic_max_f(S.maxDisp2, dx * dx + dy * dy + dz * dz);
}

__device__ __forceinline__ void track_shift(const DevView& S, int i, double dx, double dy, double dz) {   // size of a shifting move (hit-list validity)
	if (!owned_slot(S, i)) return;
	const float a = (float)dx, b = (float)dy, c = (float)dz; atomic_max_f(S.maxDisp2 + 1, (a * a + b * b + c * c) * 1.0001f);
}

// ---- periodic domain (domainTypeBC 1, one direction) -------------------------------------------------------------------
// MpsBucket::copyDataBetweenBuckets gives the first / last bucket layer of the periodic axis the particle list of the last / first
// INNER layer (src/MpsBucket.cpp:356-417); a neighbour loop that reaches such a layer sees its members moved by
// bucketPeriodicBC * periodicLength = -/+ L (getPeriodicLengths, src/MpsParticle.cpp:48-54) and forms r = (x_j + pl) - x_i (:35-45).
__device__ __forceinline__ int per_layers(const KParams& K) { return K.p.periodicAxis == 0 ? K.p.numBucketsX : K.p.periodicAxis == 1 ? K.p.numBucketsY : K.p.numBucketsZ; }
// bucket coordinates of a border layer -> the inner layer whose particles it holds
__device__ __forceinline__ void per_alias(const KParams& K, int& ix, int& iy, int& iz) {
	int& c = K.p.periodicAxis == 0 ? ix : K.p.periodicAxis == 1 ? iy : iz;
	const int n = per_layers(K);
	if (c == 0) c = n - 2; else if (c == n - 1) c = 1;
}
// what the members of the layer with coordinate c (along the periodic axis) are moved by
__device__ __forceinline__ double per_shift_of_coord(const KParams& K, int c) { return c == 0 ? -K.p.periodicLength : c == per_layers(K) - 1 ? K.p.periodicLength : 0.0; }
__device__ __forceinline__ void per_move(const KParams& K, double4& p, double sh) {   // x_j + pl, rounded as the reference rounds it
	if (K.p.periodicAxis == 0) p.x = __dadd_rn(p.x, sh); else if (K.p.periodicAxis == 1) p.y = __dadd_rn(p.y, sh); else p.z = __dadd_rn(p.z, sh);
}
__host__ __device__ __forceinline__ void cell_coords(const KParams& K, int c, int& ix, int& iy, int& iz);
// for the consumers that hold GLOBAL slots (PPE assembly, close-pair lists, debug hooks): the move of partner j as particle i sees it,
// from the layers the two were sorted into — partners within reach are either in neighbouring layers or on opposite ends of the axis
__device__ __forceinline__ double per_pair_shift(const DevView& S, const KParams& K, int i, int j) {
	int ci[3], cj[3];
	cell_coords(K, S.cellOf[i], ci[0], ci[1], ci[2]); cell_coords(K, S.cellOf[j], cj[0], cj[1], cj[2]);
	const int d = cj[K.p.periodicAxis] - ci[K.p.periodicAxis], half = per_layers(K) / 2;
	return d > half ? -K.p.periodicLength : d < -half ? K.p.periodicLength : 0.0;
}

// ---- stale bucket lists after an out-of-domain swap ------------------------------------------------------------------
// checkParticleOutDomain copies the LAST live particle over the one that left (src/MpsParticle.cpp:298-396) but the bucket
// lists are only re-threaded at the next step start (src/main.cpp:256): until then array element i — now the moved particle —
// is still listed in the bucket of the particle that left, so a neighbour loop finds it only from there.  hid_visible(): does the
// 27-bucket search of a particle at pi (bucketCoordinates, src/MpsBucket.cpp:420-425: epsilonZero on every axis) reach `cell`?
__device__ __forceinline__ bool hid_visible(const KParams& K, int cell, const double4& pi) {
	const PmpsParams& P = K.p;
	if (cell < 0) return false;
	const int nX = P.numBucketsX, nXY = nX * P.numBucketsY;
	const int cz = cell / nXY, cy = (cell - cz * nXY) / nX, cx = cell - cz * nXY - cy * nX;
	const int bx = bucket_coord(pi.x, P.domainMinX, P.invBucketSide, P.epsilonZero), by = bucket_coord(pi.y, P.domainMinY, P.invBucketSide, P.epsilonZero);
	if (abs(bx - cx) > 1 || abs(by - cy) > 1) return false;
	if (P.dim == 2) return cz == 0;   // minZ = maxZ = 0 (src/MpsViscosity.cpp:35)
	return abs(bucket_coord(pi.z, P.domainMinZ, P.invBucketSide, P.epsilonZero) - cz) <= 1;
}
// the bucket the array element of slot `slot` is listed in when the reference's lists are stale for it, else -1
__device__ __forceinline__ int hid_cell_of_slot(const DevView& S, int slot) {
	const int nh = S.hid[0]; if (nh == 0) return -1;
	const int label = S.hostIdx[slot];
	for (int e = 0; e < nh; e++) if (S.hid[1 + e] == label) return S.hid[1 + PMPS_HID_CAP + e];
	return -1;
}

// ---- lane groups --------------------------------------------------------------------------------------------------
#ifndef TV_G
#define TV_G 4   // lanes per particle in the tile walker (pmps_tv.cuh)
#endif
// Reductions over the TV_G lanes of a particle (an aligned group of a warp).  They are called by ALL 32 lanes of the warp
// in uniform control flow, so the shuffles run with the full mask (partial masks cost a WARPSYNC/vote sequence per
// shuffle: ncu showed them at 10 % of the issued instructions of a pass).  The mask argument is kept for the bodies'
// signatures and ignored.
__device__ __forceinline__ double gsum(double v, unsigned) {
	#pragma unroll
	for (int o = TV_G >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
	return v;
}
__device__ __forceinline__ int gsum(int v, unsigned) {
	#pragma unroll
	for (int o = TV_G >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
	return v;
}
__device__ __forceinline__ double gmin(double v, unsigned) {
	#pragma unroll
	for (int o = TV_G >> 1; o > 0; o >>= 1) { const double t = __shfl_xor_sync(0xffffffffu, v, o); v = t < v ? t : v; }
	return v;
}
// one 32-byte record with a single 256-bit load (LDG.E.256 on sm_100a)
__device__ __forceinline__ double4 ld256(const double4* p) {
	double4 r;
	asm("ld.global.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(r.x), "=d"(r.y), "=d"(r.z), "=d"(r.w) : "l"(p));
	return r;
}

// ---- per-particle prologue shared by the fused bodies --------------------------------------------------------------
struct Nb;
struct Self {
	const Nb* nb; int ms;       // tile walker: the staged neighbour fields and the particle's OWN staged index (nullptr outside the walker), see OWN()
	int i; bool lead; double4 pi; int type, ptype; bool hasMirror;   // hasMirror: the polygon-wall pass gave i a mirror particle   // lead: first lane of the particle's group (does the side effects)
	double mix, miy, miz;       // mirror position of i
	// wall frame of a near-wall fluid particle: unit normal n = (x_i - mirror_i)/|.|, reflection R = I - 2 n n^T,
	// len = |x_i - mirror_i| / 2 (e.g. src/MpsPndNeigh.cpp:114-136); filled only when nearWallFluid
	bool nearWallFluid;
	double wn[3], wR[9], wlen, wn2;
};
__device__ __forceinline__ void self_wall_frame(Self& s) {
	const double a = 0.5 * (s.pi.x - s.mix), b = 0.5 * (s.pi.y - s.miy), c = 0.5 * (s.pi.z - s.miz);
	s.wn2 = a * a + b * b + c * c; s.wlen = sqrt(s.wn2);
	s.wn[0] = a / s.wlen; s.wn[1] = b / s.wlen; s.wn[2] = c / s.wlen;
	s.wR[0] = 1.0 - 2.0 * s.wn[0] * s.wn[0]; s.wR[1] = 0.0 - 2.0 * s.wn[0] * s.wn[1]; s.wR[2] = 0.0 - 2.0 * s.wn[0] * s.wn[2];
	s.wR[3] = 0.0 - 2.0 * s.wn[1] * s.wn[0]; s.wR[4] = 1.0 - 2.0 * s.wn[1] * s.wn[1]; s.wR[5] = 0.0 - 2.0 * s.wn[1] * s.wn[2];
	s.wR[6] = 0.0 - 2.0 * s.wn[2] * s.wn[0]; s.wR[7] = 0.0 - 2.0 * s.wn[2] * s.wn[1]; s.wR[8] = 1.0 - 2.0 * s.wn[2] * s.wn[2];
}
// The reference zeroes the normal when the particle sits on the wall; two variants of the test exist: on |n|
// (calcWallNPCD, wall gradient, wall shifting) and on |n|^2 (wall viscosity, wall diffusive term).
struct WallFrame { double n[3]; double R[9]; double len; };
__device__ __forceinline__ void wall_frame_of(WallFrame& f, const Self& s, double eps, bool squaredTest) {
	f.len = s.wlen;
	if ((squaredTest ? s.wn2 : s.wlen) > eps) {
		for (int k = 0; k < 3; k++) f.n[k] = s.wn[k];
		for (int k = 0; k < 9; k++) f.R[k] = s.wR[k];
	} else {
		f.n[0] = f.n[1] = f.n[2] = 0.0;
		for (int k = 0; k < 9; k++) f.R[k] = (k % 4 == 0) ? 1.0 : 0.0;
	}
}
// mirror velocity of a near-wall particle (src/MpsViscosity.cpp:1729-1752 no-slip, :1565-1590 free-slip)
__device__ __forceinline__ void mirror_velocity(const KParams& K, int nearMeshType, const WallFrame& f, const double v[3], double vm[3]) {
	double vw[3] = { 0.0, 0.0, 0.0 };
	if (nearMeshType == 2) { vw[0] = K.p.uniformVelWall[0]; vw[1] = K.p.uniformVelWall[1]; vw[2] = K.p.uniformVelWall[2]; }
	const double dn = f.n[0] * vw[0] + f.n[1] * vw[1] + f.n[2] * vw[2];
	if (K.p.slipCondition == 0) {
		for (int k = 0; k < 3; k++) vm[k] = -(v[k] - 2.0 * (vw[k] - dn * f.n[k]));
	} else {
		for (int k = 0; k < 3; k++) vm[k] = MATVEC(f.R, v[0], v[1], v[2], k);
	}
}

// ---- the pair geometry handed to every stage body ------------------------------------------------------------------
struct Pair {
	int j;                      // index of the neighbour in the arrays of the Nb view the body receives
	int tj, ptj;                // particleType / PTYPE of the neighbour
	double rx, ry, rz, d2;      // r_ij = x_j - x_i ; d2 exact
	double rinv;                // 1/|r_ij| (1 ulp)
	double mx, my, mz, dm2;     // x_j - mirror_i (polygon walls, i near a wall; otherwise dm2 = +inf) ; dm2 exact
};
// Neighbour-side fields as pair() sees them: planes indexed by Pair::j.  The tile walker points them at the shared-memory
// copies of the staged halo (pmps_tv.cuh); a plane a body does not declare in NbFields is null.
struct Nb {
	const double *vx, *vy, *vz, *press, *pndi, *RHO, *MEU, *concentration, *pndki;
};
// A particle's own value of a field its pack stages for the neighbours (NbFields): from the staged copy in shared memory instead of a
// global load at the start of the round — the staged halo holds the particle itself, and nothing writes these arrays before the pass
// ends.  (ncu: the round-start loads of begin() were the largest part of the long-scoreboard stalls of the fat passes.)
#ifndef TV_OWN_STAGED
#define TV_OWN_STAGED 1
#endif
#define OWN(plane, globalValue) ((TV_OWN_STAGED && s.nb) ? s.nb->plane[s.ms] : (globalValue))

// predicates: P(re) real neighbour, M(re) mirror neighbourhood
template <bool POLY> __device__ __forceinline__ bool pred_P(const Pair& q, double re2) { return q.d2 < re2 && (!POLY || q.d2 < q.dm2); }
__device__ __forceinline__ bool pred_M(const Pair& q, double re2) { return q.d2 < re2 && q.dm2 < re2 && q.d2 < q.dm2; }

// tile-major cell index of bucket (ix, iy, iz): tiles of TX x TY x TZ buckets, x fastest inside a tile and between tiles
__host__ __device__ __forceinline__ int cell_index(const KParams& K, int ix, int iy, int iz) {
	const int tx = ix / K.TX, ty = iy / K.TY, tz = iz / K.TZ;
	const int lx = ix - tx * K.TX, ly = iy - ty * K.TY, lz = iz - tz * K.TZ;
	return ((tz * K.nTY + ty) * K.nTX + tx) * K.TV + (lz * K.TY + ly) * K.TX + lx;
}
// inverse: bucket coordinates of a cell (may lie outside the bucket grid for the padding cells of border tiles)
__host__ __device__ __forceinline__ void cell_coords(const KParams& K, int c, int& ix, int& iy, int& iz) {
	const int t = c / K.TV, l = c - t * K.TV;
	const int tz = t / (K.nTX * K.nTY), ty = (t - tz * K.nTX * K.nTY) / K.nTX, tx = t - (tz * K.nTY + ty) * K.nTX;
	const int lz = l / (K.TX * K.TY), ly = (l - lz * K.TX * K.TY) / K.TX, lx = l - (lz * K.TY + ly) * K.TX;
	ix = tx * K.TX + lx; iy = ty * K.TY + ly; iz = tz * K.TZ + lz;
}

// ---- body packs --------------------------------------------------------------------------------------------------
// A body has  struct Ctx;  static bool begin(Ctx&, S, K, self) (false: inactive for this i);
// template<bool POLY> static void pair(Ctx&, S, K, self, const Pair&);  static void end(Ctx&, S, K, self).
// Tags:  copy_pos_tag / copy_vel_tag — the body writes pos2 / vel2 in end() (Jacobi-style update of the ping-pong
// buffer); the walker mirrors pos -> pos2 / vel -> vel2 first so skipped slots stay valid.
template <class B, class = void> struct CopiesPos { static constexpr bool value = false; };
template <class B> struct CopiesPos<B, typename B::copy_pos_tag> { static constexpr bool value = true; };
template <class B, class = void> struct CopiesVel { static constexpr bool value = false; };
template <class B> struct CopiesVel<B, typename B::copy_vel_tag> { static constexpr bool value = true; };
template <class B, class = void> struct NoPairs { static constexpr bool value = false; };
template <class B> struct NoPairs<B, typename B::no_pair_tag> { static constexpr bool value = true; };   // tail bodies: begin/end only
// close_pairs_tag: the body's pair term needs d < distCollisionLimit only; its launch skips the work items whose closest pair (tracked
// when the hit list was made, k_tv_hits / k_tv_build) cannot have come that close since
template <class B, class = void> struct ClosePairs { static constexpr bool value = false; };
template <class B> struct ClosePairs<B, typename B::close_pairs_tag> { static constexpr bool value = true; };
template <class B, class = void> struct NeedsWall { static constexpr bool value = false; };
template <class B> struct NeedsWall<B, typename B::wall_frame_tag> { static constexpr bool value = true; };

template <class... Bodies> struct Fused;
template <> struct Fused<> {
	static constexpr bool kCopyPos = false, kCopyVel = false, kWall = false, kCloseOnly = true;
	struct Ctx {};
	__device__ __forceinline__ static bool begin(Ctx&, const DevView&, const KParams&, const Self&) { return false; }
	__device__ __forceinline__ static bool need_pairs(const Ctx&) { return false; }
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx&, const Nb&, const KParams&, const Self&, const Pair&) {}
	__device__ __forceinline__ static void reduce(Ctx&, unsigned) {}
	__device__ __forceinline__ static void end(Ctx&, const DevView&, const KParams&, const Self&) {}
};
template <class B, class... Rest> struct Fused<B, Rest...> {
	static constexpr bool kCopyPos = CopiesPos<B>::value || Fused<Rest...>::kCopyPos;
	static constexpr bool kCopyVel = CopiesVel<B>::value || Fused<Rest...>::kCopyVel;
	static constexpr bool kWall = NeedsWall<B>::value || Fused<Rest...>::kWall;
	static constexpr bool kCloseOnly = (ClosePairs<B>::value || NoPairs<B>::value) && Fused<Rest...>::kCloseOnly;   // every pair term is a close-pair term
	struct Ctx { typename B::Ctx c; bool on; typename Fused<Rest...>::Ctx r; };
	__device__ __forceinline__ static bool begin(Ctx& x, const DevView& S, const KParams& K, const Self& s) {
		x.on = B::begin(x.c, S, K, s);
		const bool o = Fused<Rest...>::begin(x.r, S, K, s);
		return x.on || o;
	}
	__device__ __forceinline__ static bool need_pairs(const Ctx& x) { return (x.on && !NoPairs<B>::value) || Fused<Rest...>::need_pairs(x.r); }
	template <bool POLY> __device__ __forceinline__ static void pair(Ctx& x, const Nb& S, const KParams& K, const Self& s, const Pair& q) {
		if (!NoPairs<B>::value) { if (x.on) B::template pair<POLY>(x.c, S, K, s, q); }
		Fused<Rest...>::template pair<POLY>(x.r, S, K, s, q);
	}
	__device__ __forceinline__ static void reduce(Ctx& x, unsigned m) {
		B::reduce(x.c, m);   // every lane of the warp takes part (full-mask shuffles); the sums of an idle body are never read
		Fused<Rest...>::reduce(x.r, m);
	}
	__device__ __forceinline__ static void end(Ctx& x, const DevView& S, const KParams& K, const Self& s) {
		if (x.on) B::end(x.c, S, K, s);
		Fused<Rest...>::end(x.r, S, K, s);
	}
};

