// pmps_api.cu — the C ABI of include/pmps.h: handle lifetime, host<->device transfers, the elementwise stages and
// the stage/step drivers.  Neighbour stages are instantiations of the tile-Verlet walker (pmps_tv.cuh) over the stage
// bodies (pmps_bodies.cuh).  There is no CPU fallback anywhere: without a CUDA device pmps_create fails.
#include "pmps_bodies.cuh"
#include "pmps_tv.cuh"
#include "pmps_optset.h"
#include "pmps_walk_packs.h"
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <cmath>

thread_local std::string g_create_error;

int pmps_fail(pmps_handle* h, int code, const char* fmt, ...) {
	char buf[1024]; va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
	if (h) h->err = buf; else g_create_error = buf;
	return code;
}

ProfScope::ProfScope(pmps_handle* h_, const char* name) : h(h_), idx(-1) {
	if (!h->profile) return;
	for (size_t k = 0; k < h->profNames.size(); k++) if (h->profNames[k] == name) { idx = (int)k; break; }
	if (idx < 0) { idx = (int)h->profNames.size(); h->profNames.push_back(name); h->profMs.push_back(0.f); h->profLaunches.push_back(0); }
	cudaEventRecord(h->evA, h->stream);
}
ProfScope::~ProfScope() {
	if (idx < 0) return;
	cudaEventRecord(h->evB, h->stream); cudaEventSynchronize(h->evB);
	float ms = 0.f; cudaEventElapsedTime(&ms, h->evA, h->evB);
	h->profMs[idx] += ms; h->profLaunches[idx] += 1;
}

// ================================================================================================ elementwise stages
// MpsParticle::addGravity, src/MpsParticle.cpp:111-130
__global__ void k_add_gravity(const __grid_constant__ DevView S, const __grid_constant__ KParams K) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= S.n) return;
	if (pk_type(S.pos[i]) == K.p.ghost) return;
	const double ax = S.acc.x[i] + K.p.gravityX, ay = S.acc.y[i] + K.p.gravityY, az = S.acc.z[i] + K.p.gravityZ;
	S.acc.x[i] = ax; S.acc.y[i] = ay; S.acc.z[i] = az;
	S.accStar.x[i] = ax; S.accStar.y[i] = ay; S.accStar.z[i] = az;
}

// MpsParticleVelPos::predictionVelocityPosition, src/MpsParticleVelPos.cpp:20-50
__global__ void k_predict(const __grid_constant__ DevView S, const __grid_constant__ KParams K) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= S.n) return;
	double4 p = S.pos[i];
	if (pk_type(p) == K.p.ghost) return;
	if (pk_type(p) == K.p.fluid) {
		double4 v = S.vel[i]; const double4 o = p;
		integrate_predict(p, v, S.acc.x[i], S.acc.y[i], S.acc.z[i], K.p.timeStep);
		S.vel[i] = v; S.pos[i] = p;
		track_disp(S, K, i, p);
	}
	predict_resets(S, K, i);
}

// second loop of checkParticleCollisions (src/MpsParticleCollision.cpp:108-116): vel += dvel for fluid particles
__global__ void k_apply_collision(const __grid_constant__ DevView S, const __grid_constant__ KParams K) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= S.n) return;
	if (pk_type(S.pos[i]) != K.p.fluid) return;
	double4 v = S.vel[i];
	v.x = __dadd_rn(v.x, S.dvelCollision.x[i]); v.y = __dadd_rn(v.y, S.dvelCollision.y[i]); v.z = __dadd_rn(v.z, S.dvelCollision.z[i]);
	S.vel[i] = v;
}

__global__ void k_apply_collision_dpc(const __grid_constant__ DevView S, const __grid_constant__ KParams K) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= S.n) return;
	double4 p = S.pos[i];
	if (pk_type(p) != K.p.fluid) return;
	double4 v = S.vel[i];
	dpc_apply(S, K, i, p, v);
	S.vel[i] = v; S.pos[i] = p; track_disp(S, K, i, p);
}

// MpsParticleVelPos::correctionVelocityPosition, src/MpsParticleVelPos.cpp:53-97 (velMax: max over fluid of |v|)
__global__ void k_correct(const __grid_constant__ DevView S, const __grid_constant__ KParams K) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	double v2 = 0.0;
	if (i < S.n) {
		double4 p = S.pos[i];
		if (pk_type(p) == K.p.fluid) {
			double4 v = S.vel[i]; const double4 o = p;
			integrate_correct(p, v, S.acc.x[i], S.acc.y[i], S.acc.z[i], K.p.timeStep);
			S.vel[i] = v; S.pos[i] = p;
			S.acc.x[i] = 0.0; S.acc.y[i] = 0.0; S.acc.z[i] = 0.0;
			if (owned_slot(S, i)) v2 = v.x * v.x + v.y * v.y + v.z * v.z;
			track_disp(S, K, i, p);
		}
	}
	// warp max, then one atomic per warp on the bits (non-negative doubles order like unsigned integers)
	unsigned long long b = (unsigned long long)__double_as_longlong(v2);
	#pragma unroll
	for (int o = 16; o > 0; o >>= 1) { const unsigned long long t = __shfl_xor_sync(0xffffffffu, b, o); b = t > b ? t : b; }
	if ((threadIdx.x & 31) == 0 && b > S.velMax2Bits[0]) atomicMax(S.velMax2Bits, b);
}
__global__ void k_velmax_begin(const __grid_constant__ DevView S) { S.velMax2Bits[0] = 0ull; }
__global__ void k_velmax_end(const __grid_constant__ DevView S) { S.velMax2Bits[1] = S.velMax2Bits[0]; }

// MpsPndNeigh::updatePnd, src/MpsPndNeigh.cpp:747-774
__global__ void k_update_pnd(const __grid_constant__ DevView S, const __grid_constant__ KParams K) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= S.n) return;
	if (pk_type(S.pos[i]) == K.p.ghost) return;
	S.pndi[i] = S.pndAuxVar1[i]; S.pndAuxVar1[i] = 0.0;
}

// second loops of meanPndParticlesWallDummySurface (src/MpsPndNeigh.cpp:836-850) and meanPnd (:905-918)
template <int SEL> __global__ void k_mean_pnd_finish(const __grid_constant__ DevView S, const __grid_constant__ KParams K) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= S.n) return;
	const int t = pk_type(S.pos[i]);
	if (t == K.p.ghost) return;
	const double a1 = S.pndAuxVar1[i], a2 = S.pndAuxVar2[i];
	if (SEL == 0) { if (t == K.p.wall || S.particleBC[i] == K.p.surface) S.pndi[i] = a2 < K.p.epsilonZero ? a1 : a1 / a2; }
	else S.pndi[i] = S.numNeigh[i] < 1 ? a1 : a1 / a2;
	S.pndAuxVar1[i] = 0.0; S.pndAuxVar2[i] = 0.0;
}
__global__ void k_zero_aux(const __grid_constant__ DevView S) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= S.n) return;
	S.pndAuxVar1[i] = 0.0; S.pndAuxVar2[i] = 0.0;
}

// MpsBoundaryCondition::updateParticleBC (PND_NEIGH, PND_NPCD), src/MpsBoundaryCondition.cpp:66-203
__device__ __forceinline__ int classify_bc(const DevView& S, const KParams& K, int i, int type) {
	if (type == K.p.dummyWall) return K.p.other;
	const double ps = S.pndSmall[i]; const int nn = S.numNeigh[i];
	int bc = ps < K.p.betaPnd ? K.p.surface : K.p.inner;
	if (K.p.freeSurfType == 0) bc = (ps < K.p.betaPnd && nn < K.p.betaNeigh) ? K.p.surface : K.p.inner;
	else if (K.p.freeSurfType == 1) { if (bc == K.p.surface && nn > 4 && S.npcdDeviation2[i] < K.p.delta2) bc = K.p.inner; }
	return bc;
}
__global__ void k_update_bc(const __grid_constant__ DevView S, const __grid_constant__ KParams K) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= S.n) return;
	const int t = pk_type(S.pos[i]);
	if (t == K.p.ghost) return;
	S.particleBC[i] = classify_bc(S, K, i, t);
}

// MpsPressure::calcPressEMPS / calcPressWCMPS, src/MpsPressure.cpp:62-163
__global__ void k_press_eos(const __grid_constant__ DevView S, const __grid_constant__ KParams K) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= S.n) return;
	const double4 pi = S.pos[i];
	if (pk_type(pi) == K.p.ghost) return;
	const double mi = pk_ptype(pi) == 1 ? K.p.DNS_FL1 : K.p.DNS_FL2; double p = 0.0;
	if (S.particleBC[i] == K.p.inner) {
		if (K.p.mpsType == 0) p = (S.pndi[i] - K.p.pndSmallZero) * K.p.coeffPressEMPS * mi;
		else p = (mi * K.p.coeffPressWCMPS / K.p.gamma) * (pow(S.pndi[i] / K.p.pndSmallZero, K.p.gamma) - 1);
	}
	if (p < 0.0) p = 0.0;
	S.press[i] = p;
}

// fused elementwise pair of the whole-step path: updateParticleBC then the equation of state
__global__ void k_bc_eos(const __grid_constant__ DevView S, const __grid_constant__ KParams K) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= S.n) return;
	const double4 pi = S.pos[i]; const int t = pk_type(pi);
	if (t == K.p.ghost) return;
	const int bc = classify_bc(S, K, i, t);
	S.particleBC[i] = bc;
	const double mi = pk_ptype(pi) == 1 ? K.p.DNS_FL1 : K.p.DNS_FL2; double p = 0.0;
	if (bc == K.p.inner) {
		if (K.p.mpsType == 0) p = (S.pndi[i] - K.p.pndSmallZero) * K.p.coeffPressEMPS * mi;
		else p = (mi * K.p.coeffPressWCMPS / K.p.gamma) * (pow(S.pndi[i] / K.p.pndSmallZero, K.p.gamma) - 1);
	}
	if (p < 0.0) p = 0.0;
	S.press[i] = p;
}

// ================================================================================================ handle lifetime
template <class T> static int dalloc(pmps_handle* h, T** p, size_t count) {
	void* q = nullptr;
	cudaError_t e = cudaMalloc(&q, sizeof(T) * (count ? count : 1));
	if (e != cudaSuccess) return pmps_fail(h, PMPS_E_CUDA, "cudaMalloc(%zu bytes) failed: %s", sizeof(T) * count, cudaGetErrorString(e));
	cudaMemsetAsync(q, 0, sizeof(T) * (count ? count : 1), h->stream);
	h->allocs.push_back(q); *p = (T*)q; return PMPS_OK;
}

static int alloc_view(pmps_handle* h, DevView& V, int n0, bool primary) {
	int rc = 0;
	rc |= dalloc(h, &V.pos, n0); rc |= dalloc(h, &V.vel, n0); rc |= dalloc(h, &V.hostIdx, n0); rc |= dalloc(h, &V.particleNearWall, n0); rc |= dalloc(h, &V.stepCell, n0); rc |= dalloc(h, &V.wrapHid, h->K.p.domainTypeBC == 1 ? n0 : 1);
#define X(n) rc |= dalloc(h, &V.n, n0);
	PMPS_I32_FIELDS(X)
	PMPS_F64_FIELDS(X)
#undef X
#define X(n) rc |= dalloc(h, &V.n, h->K.p.fluidType == 1 ? n0 : 1);
	PMPS_RHEO_FIELDS(X)
#undef X
#define X(n) rc |= dalloc(h, &V.n.x, n0); rc |= dalloc(h, &V.n.y, n0); rc |= dalloc(h, &V.n.z, n0);
	PMPS_V3_FIELDS(X)
#undef X
	(void)primary;
	return rc ? PMPS_E_CUDA : PMPS_OK;
}

__global__ void k_iota(int32_t* a, int32_t* b, int n) { const int i = blockIdx.x * blockDim.x + threadIdx.x; if (i < n) { a[i] = i; b[i] = i; } }

extern "C" int pmps_abi_version(void) { return PMPS_ABI_VERSION; }

extern "C" const char* pmps_last_error(const pmps_t* h) { return h ? h->err.c_str() : g_create_error.c_str(); }

// numParticlesZero is taken from the first pmps_upload; pmps_create only validates and binds the device.
extern "C" int pmps_create(const PmpsParams* params, int device, pmps_t** out) {
	if (!params || !out) return pmps_fail(nullptr, PMPS_E_ARG, "pmps_create: NULL argument");
	*out = nullptr;
	if (params->abi_version != PMPS_ABI_VERSION) return pmps_fail(nullptr, PMPS_E_ARG, "pmps_create: ABI version %d, library is %d", params->abi_version, PMPS_ABI_VERSION);
	if (params->dim != 2 && params->dim != 3) return pmps_fail(nullptr, PMPS_E_ARG, "pmps_create: dim must be 2 or 3");
	if (params->fluidType != 0 && params->fluidType != 1) return pmps_fail(nullptr, PMPS_E_ARG, "pmps_create: fluidType %d", params->fluidType);
	if (params->fluidType == 1 && params->wallType != 0 && params->dim != 3) return pmps_fail(nullptr, PMPS_E_UNSUPPORTED, "pmps_create: the polygon-wall variants of the non-Newtonian rheology (src/MpsViscosity.cpp:703-1526) exist in 3D only — the reference's code reads the bed level along z");
	if (params->fluidType == 1 && params->wallType != 0 && params->slipCondition != 0 && params->slipCondition != 1) return pmps_fail(nullptr, PMPS_E_ARG, "pmps_create: slipCondition %d", params->slipCondition);
	if (params->fluidType == 1 && params->fractionMethod != 1 && params->fractionMethod != 2) return pmps_fail(nullptr, PMPS_E_ARG, "pmps_create: fractionMethod %d", params->fractionMethod);
	if (params->domainTypeBC != 0 && params->domainTypeBC != 1) return pmps_fail(nullptr, PMPS_E_ARG, "pmps_create: domainTypeBC %d", params->domainTypeBC);
	if (params->domainTypeBC == 1) {   // periodic in one direction (src/MpsBucket.cpp:142-289)
		const int a = params->periodicAxis;
		if (a < 0 || a > 2 || (a == 2 && params->dim == 2)) return pmps_fail(nullptr, PMPS_E_ARG, "pmps_create: periodicAxis %d in a %dD case", a, params->dim);
		const int nA = a == 0 ? params->numBucketsX : a == 1 ? params->numBucketsY : params->numBucketsZ;
		if (!(params->periodicLength > 0.0) || !(params->physDomMax > params->physDomMin)) return pmps_fail(nullptr, PMPS_E_ARG, "pmps_create: periodic domain without periodicLength / physDomMin / physDomMax");
		// the two border layers alias inner layers: a tile must not meet a bucket and its alias at once
		if (nA < 8) return pmps_fail(nullptr, PMPS_E_UNSUPPORTED, "pmps_create: %d bucket layers along the periodic axis; the device path needs at least 8 (6 inner layers)", nA);
		if (params->freeSurfType == 2) return pmps_fail(nullptr, PMPS_E_UNSUPPORTED, "pmps_create: ARC free-surface detection in a periodic domain (the reference's arc test mixes moved and unmoved neighbour positions, src/MpsBoundaryCondition.cpp:205-390)");
		if (params->fluidType == 1) return pmps_fail(nullptr, PMPS_E_UNSUPPORTED, "pmps_create: non-Newtonian rheology in a periodic domain is not on the device path");
	}
	if (params->freeSurfType < 0 || params->freeSurfType > 2) return pmps_fail(nullptr, PMPS_E_ARG, "pmps_create: freeSurfType %d", params->freeSurfType);
	if (params->collisionType != 0 && params->collisionType != 1) return pmps_fail(nullptr, PMPS_E_ARG, "pmps_create: collisionType %d", params->collisionType);
	if (params->mpsType < 0 || params->mpsType > 3) return pmps_fail(nullptr, PMPS_E_ARG, "pmps_create: mpsType %d", params->mpsType);
	if (params->numBucketsX < 1 || params->numBucketsY < 1 || params->numBucketsZ < 1 || !(params->bucketSide > 0.0)) return pmps_fail(nullptr, PMPS_E_ARG, "pmps_create: bucket geometry not set");
	const long long nB = (long long)params->numBucketsX * params->numBucketsY * params->numBucketsZ;
	if (nB > 2000000000ll) return pmps_fail(nullptr, PMPS_E_OVERFLOW, "pmps_create: %lld buckets overflow int32", nB);
	int ndev = 0;
	if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) { cudaGetLastError(); return pmps_fail(nullptr, PMPS_E_CUDA, "pmps_create: no CUDA device (this library has no CPU path)"); }
	if (device < 0 || device >= ndev) return pmps_fail(nullptr, PMPS_E_ARG, "pmps_create: device %d of %d", device, ndev);
	if (cudaSetDevice(device) != cudaSuccess) return pmps_fail(nullptr, PMPS_E_CUDA, "pmps_create: cudaSetDevice failed");
	pmps_handle* h = new pmps_handle();
	h->device = device;
	if (cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking) != cudaSuccess) { delete h; return pmps_fail(nullptr, PMPS_E_CUDA, "pmps_create: cudaStreamCreate failed"); }
	cudaEventCreate(&h->evA); cudaEventCreate(&h->evB);
	KParams& K = h->K; memset(&K, 0, sizeof K);
	K.p = *params;
	K.big = 10e8 * params->partDist;
	K.vol = pow(params->partDist, (double)params->dim);
	K.velMax = params->velMax;
	K.poly = params->wallType == 1;
	K.reMax2 = fmax(params->reS2, params->reL2); K.invReS = 1.0 / params->reS; K.invReL = 1.0 / params->reL;
	h->optset = getenv("PMPS_GENERIC_KERNELS") ? PMPS_OPT_GENERIC : pmps_optset_of(*params);   // walker kernels with this option set's enums as constants (pmps_optset.h)
	memset(&h->V, 0, sizeof(DevView)); memset(&h->A, 0, sizeof(DevView));
	h->V.nB = (int)nB;
	// tiles of the tile-Verlet walker (pmps_tv.cuh): 2x2x2 buckets in 3D (halo 4x4x4), 4x4 in 2D (halo 6x6)
	K.TX = params->dim == 3 ? 2 : 4; K.TY = K.TX; K.TZ = params->dim == 3 ? 2 : 1;
	if (const char* e = getenv("PMPS_TILE")) { int a = 0, b = 0, c = 0; if (sscanf(e, "%d,%d,%d", &a, &b, &c) == 3 && a > 0 && b > 0 && c > 0 && (a + 2) * (b + 2) * (c + 2) <= TV_HMAX) { K.TX = a; K.TY = b; K.TZ = params->dim == 3 ? c : 1; } }
	K.TV = K.TX * K.TY * K.TZ;
	K.HVt = (K.TX + 2) * (K.TY + 2) * (params->dim == 3 ? K.TZ + 2 : 1); if (K.HVt < 27) K.HVt = 27;
	K.tabStride = (2 * K.HVt + 6 + 3) & ~3;   // hOff[HVt + 1], hBase[HVt], own0, P, box origin (3), rounded up to a multiple of 4
	K.nTX = (params->numBucketsX + K.TX - 1) / K.TX; K.nTY = (params->numBucketsY + K.TY - 1) / K.TY; K.nTZ = (params->numBucketsZ + K.TZ - 1) / K.TZ;
	const long long nC = (long long)K.nTX * K.nTY * K.nTZ * K.TV;
	if (nC > 2000000000ll) { delete h; return pmps_fail(nullptr, PMPS_E_OVERFLOW, "pmps_create: %lld cells overflow int32", nC); }
	h->V.nC = (int)nC; h->nTilesAll = K.nTX * K.nTY * K.nTZ;
	*out = h;
	return PMPS_OK;
}

// margin of the hit lists made from now on: a list reaches hitDelta beyond every cut-off on both sides and stays valid until somebody
// may have moved hitDelta since it was made (k_hits_decide); a larger margin means fewer re-makes and more rejected candidates per
// pass, (re + 2 delta)^3.  The close-pair radius of the collision skip moves with it.
void pmps_set_hit_delta(pmps_handle* h, double delta) {
	KParams& K = h->K; const PmpsParams& P = K.p;
	const double maxRe = fmax(P.reL, P.reS), rl = P.bucketSide;
	K.hitDelta = delta;
	{ const double r = (maxRe + 2.0 * K.hitDelta) * (1.0 + 1e-5); K.hitR2 = (float)(r * r); }
	{ const double r = ((P.collisionType == 1 ? P.partDist : sqrt(P.distCollisionLimit2)) + 2.0 * K.hitDelta) * (1.0 + 1e-5) + 8.0 * 1.2e-7 * 4.0 * rl; K.closeR2 = (float)(r * r); }
}
static int ensure_capacity(pmps_handle* h, int n0) {
	if (h->numParticlesZero == n0) return PMPS_OK;
	if (h->numParticlesZero != 0) return pmps_fail(h, PMPS_E_ARG, "numParticlesZero changed from %d to %d; create a new handle", h->numParticlesZero, n0);
	if (n0 < 1) return pmps_fail(h, PMPS_E_ARG, "numParticlesZero must be positive");
	cudaSetDevice(h->device);
	DevView& S = h->V; KParams& K = h->K; const PmpsParams& P = K.p;
	int rc = 0;
	rc |= alloc_view(h, S, n0, true); rc |= alloc_view(h, h->A, n0, false);
	rc |= dalloc(h, &S.scratch1, n0); rc |= dalloc(h, &S.bpos, n0); rc |= dalloc(h, &S.hpos, n0); rc |= dalloc(h, &S.slotOfHost, n0); rc |= dalloc(h, &S.cellOf, n0); rc |= dalloc(h, &S.hid, 1 + 2 * PMPS_HID_CAP);
	rc |= dalloc(h, &S.cellStart, (size_t)S.nC + 2); rc |= dalloc(h, &h->cellCount, (size_t)S.nC + 2);
	rc |= dalloc(h, &h->perm, n0); rc |= dalloc(h, &h->arbRank, n0); rc |= dalloc(h, &h->newCell, n0);
	rc |= dalloc(h, &h->tmpHost, 2 * (size_t)n0);
	h->nScanBlocks = cdiv((long long)S.nC + 2, 512 * 8) + 1;
	rc |= dalloc(h, &h->scanBlock, h->nScanBlocks);
	{	// work items of the tile-Verlet walker: at most one per non-empty cell
		const size_t items = (size_t)(S.nC < n0 ? S.nC : n0) + 1;
		rc |= dalloc(h, &S.tileList, items); rc |= dalloc(h, &S.itemMinD2, items); rc |= dalloc(h, &S.actList, 3 * items); rc |= dalloc(h, &S.actCount, 4); rc |= dalloc(h, &S.tileSparse, (size_t)h->nTilesAll + 1); S.actCap = (int32_t)items; rc |= dalloc(h, &S.tileClose, (size_t)h->nTilesAll + 1); rc |= dalloc(h, &S.itemTab, items * (size_t)K.tabStride); rc |= dalloc(h, &S.nTiles, 1); rc |= dalloc(h, &S.tileCursor, PMPS_CURSORS); rc |= dalloc(h, &S.tileWall, (size_t)h->nTilesAll + 1);
		int sm = 148; cudaDeviceGetAttribute(&sm, cudaDevAttrMultiProcessorCount, h->device); h->smCount = sm;
	}
	{	// column maps of the rheology (src/MpsViscosity.cpp:236-238)
		const double aa = 2.0 * K.p.partDist;
		K.rheoKx = (int)((K.p.domainMaxX - K.p.domainMinX) / aa) + 1;
		K.rheoKy = K.p.dim == 3 ? (int)((K.p.domainMaxY - K.p.domainMinY) / aa) + 1 : 1;
		const size_t nc = K.p.fluidType == 1 ? (size_t)(K.rheoKx + 1) * (K.rheoKy + 1) : 1;
		rc |= dalloc(h, &S.rheoBL, nc); rc |= dalloc(h, &S.rheoWL, nc); rc |= dalloc(h, &S.rheoIdx, nc); rc |= dalloc(h, &S.meuMaxBits, 1);
	}
	rc |= dalloc(h, &S.velMax2Bits, 2); rc |= dalloc(h, &S.maxDisp2, 4); rc |= dalloc(h, &S.hitMove, 4); rc |= dalloc(h, &S.hitFlag, 2); rc |= dalloc(h, &S.flags, NFLAGS); rc |= dalloc(h, &S.oodList, PMPS_OOD_CAP);
	// Verlet list: radius = bucketSide (the reference's own search slack, src/MpsBucket.cpp:23); capacity from the
	// lattice count with head room (PMPS_LIST_CAP overrides)
	const double rl = P.bucketSide, lo = P.partDist;
	const double lat = P.dim == 3 ? 4.18879 * pow(rl / lo, 3.0) : 3.14159 * pow(rl / lo, 2.0);
	int cap = (int)(lat * 1.6) + 8;
	if (const char* e = getenv("PMPS_LIST_CAP")) cap = atoi(e);
	cap = (cap + 3) & ~3; if (cap < 8) cap = 8; if (cap > 4096) cap = 4096;
	S.listCap = cap;   // the global-slot list itself is allocated on first use (pmps_ensure_list)
	// tile-local list: whole chunks of 64 entries; staging capacity from the shared-memory budget of the fattest pass
	S.tlCap = ((cap + TV_CHUNK - 1) / TV_CHUNK) * TV_CHUNK; if (S.tlCap > 255 * TV_CHUNK) S.tlCap = 255 * TV_CHUNK;
	rc |= dalloc(h, &S.hlist, (size_t)n0 * (size_t)S.tlCap); rc |= dalloc(h, &S.hcnt, n0); rc |= dalloc(h, &S.clist, (size_t)n0 * PMPS_CLOSE_CAP); rc |= dalloc(h, &S.ccnt, n0);
	K.tvMaxM = TV_MAXM;   // follow passes; a scanning pass of the fattest pack gets one block per SM
	if (K.tvMaxM > 65534) K.tvMaxM = 65534;
	if (const char* e = getenv("PMPS_TILE_MAXM")) { const int v = atoi(e); if (v > 0 && v < K.tvMaxM) K.tvMaxM = v; }   // tests: force the split path
	if (K.tvMaxM < 64) return pmps_fail(h, PMPS_E_UNSUPPORTED, "candidate lists of %d entries leave no room for the staged halo", S.tlCap);
	h->stageBytes = 64 * (size_t)n0;
	{ void* q = nullptr; if (cudaMalloc(&q, h->stageBytes) != cudaSuccess) rc |= 1; else { h->allocs.push_back(q); h->stage = q; } }
	if (cudaMallocHost((void**)&h->h_flags, sizeof(int32_t) * NFLAGS) != cudaSuccess) rc |= 1;
	if (rc) return pmps_fail(h, PMPS_E_CUDA, "device allocation for %d particles failed: %s", n0, cudaGetErrorString(cudaGetLastError()));
	S.n = n0; S.n0 = n0; S.nOwn = 0x7fffffff; S.pos2 = h->A.pos; S.vel2 = h->A.vel;
	h->A.n = n0; h->A.n0 = n0;
	// float margin of the list test: positions are stored relative to domainMin in float
	const double ext = fmax(fmax(P.domainMaxX - P.domainMinX, P.domainMaxY - P.domainMinY), P.domainMaxZ - P.domainMinZ);
	const double maxRe = fmax(P.reL, P.reS);
	const double rlf = rl * (1.0 + 1e-6) + 8.0 * 1.2e-7 * ext;
	K.rl2 = (float)(rlf * rlf);
	const double guard = 0.5 * (rl - maxRe) - 8.0 * 1.2e-7 * ext;
	K.guard2 = (float)(guard > 0 ? guard * guard : 0.0);
	// reach of the hit lists beyond the cut-offs: a hit list stays valid until somebody may have moved hitDelta since it was made
	// (k_hits_decide); a larger reach means fewer rebuilds and more rejected candidates per pass, (re + 2 delta)^3
	h->hitDeltaFixed = -1.0;
	if (const char* e = getenv("PMPS_HIT_DELTA")) { const double v = atof(e); if (v > 0.0 && v < 0.2) h->hitDeltaFixed = v * lo; }
	pmps_set_hit_delta(h, h->hitDeltaFixed > 0.0 ? h->hitDeltaFixed : 0.06 * lo);
	// |x_j - mirror_i| >= 2 dist(i, wall) - displacement - |x_j - x_i| > re whenever dist > re + displacement/2: margin = the full guard
	{ const double far = maxRe * (1.0 + 1e-6) + (guard > 0 ? guard : 0.0) + 1e-6 * lo; K.mirrorFar2 = far * far; }
	k_iota<<<cdiv(n0, 256), 256, 0, h->stream>>>(S.hostIdx, S.slotOfHost, n0);
	{ unsigned long long vb[2]; double v2 = K.velMax * K.velMax; memcpy(&vb[0], &v2, 8); vb[1] = vb[0];
	  cudaMemcpyAsync(S.velMax2Bits, vb, 16, cudaMemcpyHostToDevice, h->stream); cudaStreamSynchronize(h->stream); }
	h->numParticlesZero = n0;
	cudaError_t e = cudaGetLastError();
	if (e != cudaSuccess) return pmps_fail(h, PMPS_E_CUDA, "initialisation failed: %s", cudaGetErrorString(e));
	return PMPS_OK;
}
int pmps_ensure_capacity(pmps_handle* h, int n0) { return ensure_capacity(h, n0); }

extern "C" int pmps_destroy(pmps_t* h) {
	if (!h) return PMPS_E_ARG;
	cudaSetDevice(h->device);
	cudaStreamSynchronize(h->stream);
	for (auto& r : h->registered) cudaHostUnregister(r.first);
	for (void* p : h->allocs) cudaFree(p);
	for (DevMesh& m : h->meshes) { cudaFree(m.V); cudaFree(m.F); cudaFree(m.N); cudaFree(m.bbMin); cudaFree(m.bbMax); cudaFree(m.prim);
		cudaFree(m.xDataPND); cudaFree(m.xDataNeigh); cudaFree(m.Zij); cudaFree(m.nNeighWall); }
	if (h->d_meshes) cudaFree(h->d_meshes);
	if (h->h_flags) cudaFreeHost(h->h_flags);
	cudaEventDestroy(h->evA); cudaEventDestroy(h->evB);
	cudaStreamDestroy(h->stream);
	delete h;
	return PMPS_OK;
}

extern "C" int pmps_set_mesh(pmps_t* h, int mesh_id, int nV, const double* V, int nF, const int32_t* F, const double* faceNormals,
                             int nTable, const double* xDataPND, const double* xDataNeigh, const double* Zij, const int32_t* nNeighWall) {
	if (!h || !V || !F || !faceNormals || !xDataPND || !xDataNeigh || !Zij || !nNeighWall) return h ? pmps_fail(h, PMPS_E_ARG, "pmps_set_mesh: NULL argument") : PMPS_E_ARG;
	cudaSetDevice(h->device);
	cudaStreamSynchronize(h->stream);
	return pmps_mesh_upload(h, mesh_id, nV, V, nF, F, faceNormals, nTable, xDataPND, xDataNeigh, Zij, nNeighWall);
}

// ---- walker dispatch: option set x pack group -> the translation unit that holds the kernels (pmps_walk_inst.inc) ---------------
#define PMPS_WALK_DECL(set) int pmps_walk_##set##_g0(int, pmps_handle*, const char*, int); int pmps_walk_##set##_g1(int, pmps_handle*, const char*, int); \
	int pmps_walk_##set##_g2(int, pmps_handle*, const char*, int); int pmps_walk_##set##_g3(int, pmps_handle*, const char*, int);
PMPS_WALK_DECL(generic) PMPS_WALK_DECL(wc3d) PMPS_WALK_DECL(inc3d)
int pmps_walk_dispatch(int optset, int packId, int group, pmps_handle* h, const char* name, int polyOverride) {
	typedef int (*fn)(int, pmps_handle*, const char*, int);
	static const fn table[PMPS_OPT_SETS][PMPS_WALK_GROUPS] = {
		{ pmps_walk_generic_g0, pmps_walk_generic_g1, pmps_walk_generic_g2, pmps_walk_generic_g3 },
		{ pmps_walk_wc3d_g0, pmps_walk_wc3d_g1, pmps_walk_wc3d_g2, pmps_walk_wc3d_g3 },
		{ pmps_walk_inc3d_g0, pmps_walk_inc3d_g1, pmps_walk_inc3d_g2, pmps_walk_inc3d_g3 } };
	if (optset < 0 || optset >= PMPS_OPT_SETS || group < 0 || group >= PMPS_WALK_GROUPS) return pmps_fail(h, PMPS_E_ARG, "%s: option set %d / group %d", name, optset, group);
	return table[optset][group](packId, h, name, polyOverride);
}

#include "pmps_stages.inc"
