// pmps_internal.h — device-side data model of libpmps.so (the B200-native PolyMPS particle hot path).
//
// Layout: all per-particle state lives in HBM in SLOT order = particles sorted by the reference's
// bucket id (src/MpsBucket.cpp:299-353), ascending host index inside a bucket (= the reference's list order).
//   pos, vel      : 32-byte records {x,y,z,w}; pos.w carries the bits of {particleType, PTYPE}; one sector per gather
//   every other field : one 8-byte (or 4/1-byte) plane per component, coalesced own-index access
//   hlist         : per-epoch candidates as 16-bit indices into the staged halo of the particle's tile (pmps_tv.cuh);
//                   list: Verlet candidates as global slot ids, built on demand (PPE assembly)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>
#include "../../include/pmps.h"

// ---- field tables (names = reference MpsParticle members, include/MpsParticle.h:139-241) ----------------------
#define PMPS_I32_FIELDS(X) X(particleBC) X(numNeigh) X(nearMeshType) X(numNeighWallContribution) X(elementID)
#define PMPS_F64_FIELDS(X) X(press) X(pndi) X(pndki) X(pndski) X(pndSmall) X(npcdDeviation2) X(concentration) \
	X(velDivergence) X(diffusiveTerm) X(pndWallContribution) X(distParticleWall2) X(deviationDotPolygonNormal) \
	X(RHO) X(MEU) X(pndAuxVar1) X(pndAuxVar2) X(pressurePPE) X(sourceTerm)
// non-Newtonian fields (include/MpsParticle.h, allocated by the reference only when fluidType == 1, src/MpsInputOutput.cpp:729-744):
// full-size planes only on handles created with fluidType 1, one element otherwise
#define PMPS_RHEO_FIELDS(X) X(Cv) X(II) X(MEU_Y) X(Inertia) X(p_rheo_new) X(p_smooth) X(VF) X(S11) X(S12) X(S13) X(S22) X(S23) X(S33)
#define PMPS_V3_FIELDS(X) X(acc) X(accStar) X(npcdDeviation) X(gradConcentration) X(correcMatrixRow1) \
	X(correcMatrixRow2) X(correcMatrixRow3) X(particleAtWallPos) X(mirrorParticlePos) X(polygonNormal) X(forceWall) \
	X(dvelCollision) X(normal)

struct V3 { double* x; double* y; double* z; };

// Pointers to the CURRENT buffers, passed by value to every kernel.
struct DevView {
	double4* pos; double4* vel;          // current
	double4* pos2; double4* vel2;        // alternate (ping-pong target of integrators / permute)
	float4* bpos;                        // position at list-build time, float, relative to domainMin (guard + build)
	float4* hpos;                        // position when the hit list was made, same format (hit-list validity, k_hits_decide)
	int32_t* hostIdx;                    // slot -> host (reference) index
	int32_t* slotOfHost;                 // host index -> slot
	int32_t* cellOf;                     // slot -> bucket id at the last rebuild
	int32_t* stepCell;                   // slot -> the reference bucket (iz*nXY + iy*nX + ix) the particle was threaded on at the START of the current step
	                                     // (updateParticlesID, src/MpsBucket.cpp:299-353); read when a particle leaves the domain (pmps_domain.cu)
	int32_t* hid;                        // particles the reference's stale bucket lists hide for the rest of a step after an out-of-domain swap:
	                                     // [0] entries, [1 .. CAP] host label (-1: dead), [1 + CAP .. 2 CAP] the bucket its array element sits in (hid_visible())
	int32_t* cellStart;                  // nC+2, slots of cell c (tile-major bucket index, cell_index()) are [cellStart[c], cellStart[c+1]); cell nC = ghosts
	uint32_t* list; uint16_t* listCnt;   // Verlet candidates (slot ids), row-major, listCap per particle; built on demand
	uint32_t* clist; uint8_t* ccnt;      // close-pair list of every fluid slot (k_hl_build): GLOBAL slots of the neighbours within closeR when the hit list was
	                                       // made, PMPS_CLOSE_CAP per slot; ccnt 255 = more than that (the collision kernel then searches the buckets itself)
	uint16_t* hlist; uint16_t* hcnt;       // hit list (pmps_tv.cuh, k_hl_build): the staged records within reach at the positions of the current epoch, tlCap entries
	                                       // per slot in chunks of 64 (tv_list_pos), as 16-bit indices into the staged halo of the slot's work item; entries used
	uint8_t* particleNearWall;
	uint8_t* wrapHid;                    // periodic domain: 1 = moved through a periodic face since the step start (found by nobody until the lists are re-threaded); n0 bytes, 1 byte when not periodic
	double* ppeP;                        // CG search direction (pmps_ppe.cu), here so that the halo pack / unpack kernels reach it
	double* scratch1;                    // per-slot scratch (Pmin of the gradient passes); not a reference field
#define X(n) int32_t* n;
	PMPS_I32_FIELDS(X)
#undef X
#define X(n) double* n;
	PMPS_F64_FIELDS(X)
	PMPS_RHEO_FIELDS(X)
#undef X
#define X(n) V3 n;
	PMPS_V3_FIELDS(X)
#undef X
	int32_t n;                           // slots in use (= live particles at the last rebuild incl. later ghosts)
	int32_t n0;                          // capacity (numParticlesZero)
	int32_t nOwn;                        // slab decomposition (pmps_slab.cu): host indices >= nOwn are HALO COPIES of particles another
	                                     // rank owns; they take part in every neighbour sum but not in velMax / displacement / out-of-domain
	                                     // bookkeeping (their owner does that).  INT32_MAX on a single GPU.
	int32_t listCap;
	int32_t nB;                          // reference buckets (numBucketsX*Y*Z)
	int32_t nC;                          // cells of the tile-major index (>= nB: tiles are padded to full size)
	int32_t tlCap;                       // entries per slot in hlist (multiple of 64)
	// device scalars
	unsigned long long* velMax2Bits;     // max |v|^2 of fluid particles (bits of a non-negative double)
	float* maxDisp2;                     // [0] max squared displacement since the list build, [1] max squared shift of the shifting passes of the step (float, >= 0),
	                                     // [2] copy of [0] taken at the end of the last step (one GPU: mid_step_check derives the size of a step's small moves from it)
	float* hitMove;                      // hit-list validity (pmps_tv.cuh): [0] largest squared displacement of ANY slot (halo copies included)
	                                     // from its position when the hit list was made (running maximum, reset by a rebuild), [1] the margin
	                                     // hitDelta the current list was made with
	int32_t* hitFlag;                    // [0] 1: k_tv_hits must rebuild the hit list now (set by k_hits_decide), [1] rebuilds so far
	int32_t* flags;                      // [0] list overflow, [1] out-of-domain count, [2] guard violated, [3] live count
	int32_t* oodList;                    // host ids flagged by the out-of-domain test
	const struct DevMesh* meshes; int32_t nMeshes;   // polygon meshes (device array), set by pmps_set_mesh
	int32_t listValid;                   // the Verlet list of the current bucket order has been built
	int32_t* tileList; int32_t* nTiles;  // work items of the last rebuild: tile id (whole tile) or ~cell (one bucket of a tile whose halo
	                                     // does not fit the staging area); count on device
	uint8_t* tileClose;                  // per tile: some pair of the tile was closer than closeR when the hit list was made (split tiles: always 1)
	uint8_t* tileSparse;                 // per tile: some particle of it wants pairs in the sparse pack about to run (k_sparse_mark)
	int32_t* actList; int32_t* actCount; // work items a wall-only ([0, cap)) / close-pair ([cap, 2 cap)) / sparse ([2 cap, 3 cap)) pass has to visit, compacted before the pass (k_item_flags); counts [3]
	int32_t actCap;
	float* itemMinD2;                    // per work item: smallest squared pair distance (float positions) when the hit list was made
	int32_t* itemTab;                    // per work item: precomputed halo table (K.tabStride ints, pmps_tv.cuh)
	int32_t* tileCursor;                 // work-stealing cursors, one per walk launch of a step
	uint8_t* tileWall;                   // per tile: some particle of the tile is near a polygon wall (set by the wall pass)
	// rheology (pmps_rheo.cuh): bed level / water level per column of 2 lo x 2 lo (src/MpsViscosity.cpp:236-311) as order-preserving
	// 64-bit keys, the host index of the particle that set the bed level (its pressure is PS), max(MEU) of the mixture particles
	unsigned long long* rheoBL; unsigned long long* rheoWL; int32_t* rheoIdx; unsigned long long* meuMaxBits;
};

enum { FLAG_OVERFLOW = 0, FLAG_OOD_COUNT = 1, FLAG_GUARD = 2, FLAG_LIVE = 3, FLAG_MAXCELL = 4, FLAG_OOD_TOTAL = 5, FLAG_TILE_OVERFLOW = 6, FLAG_UPLOAD_TYPES = 7, FLAG_WRAPPED = 8, FLAG_WRAP_LAST = 9, FLAG_WRAP_STEP = 10, NFLAGS = 11 };
#define PMPS_OOD_CAP 65536
#define PMPS_HID_CAP 8192    // particles moved into a hole by the out-of-domain swap, or to the other end of a periodic domain, within one step
#define PMPS_CLOSE_CAP 16 // close-pair list entries per slot (collision pass)
#define PMPS_CURSORS 64   // work-stealing cursors per step (one per tile-walker launch)

// Kernel-side parameters: the ABI params plus derived constants
// (a template on the type of `p`: the walker kernels are also compiled with option sets whose enums are compile-time constants,
// pmps_optset.h — same layout, the constants hide the run-time members)
template <class PT> struct KParamsT {
	PT p;
	double big;          // 10e8*partDist, the reference's "infinity" (src/MpsParticleVelPos.cpp:37)
	double vol;          // partDist^dim
	double velMax;       // velMax of the previous correction (src/MpsPressure.cpp:1585)
	float rl2;           // (list radius)^2 incl. float margin
	float guard2;        // (allowed displacement)^2
	float closeR2;       // (sqrt(distCollisionLimit2) + 2 hitDelta)^2 with the float margin: work items whose closest pair was farther skip the collision walk
	float hitR2;         // (reach of the hit lists)^2 = (max(reS, reL) + 2 hitDelta)^2 with the float margin
	double hitDelta;     // a hit list stays valid while no particle moved farther than this since it was made (covers the shifting)
	int32_t poly;        // wallType == polygon
	int32_t TX, TY, TZ, TV;       // buckets per tile along x, y, z and their product
	int32_t nTX, nTY, nTZ;        // tiles per axis
	int32_t HVt, tabStride;       // halo buckets of a whole tile; ints per item table
	int32_t tvMaxM;               // staged records a tile may need (smallest capacity over the passes): larger tiles are split
	int32_t rheoKx, rheoKy;       // columns of the bed / water level maps (kx_max, ky_max of src/MpsViscosity.cpp:237-238)
	double reMax2;       // max(reS, reL)^2: no predicate of the reference reaches farther
	double mirrorFar2;   // a particle farther than this from every wall (distParticleWall2) cannot fail d2 < dm2 nor pass M(re): its mirror is skipped
	double invReS, invReL;
};
typedef KParamsT<PmpsParams> KParams;

#define MAX_TRI_CONST 64
struct DevMesh {       // one polygon mesh on the device + libigl-order AABB hierarchy in heap order
	int nV, nF, nNodes, nTable;
	double* V; int32_t* F; double* N; double* bbMin; double* bbMax; int32_t* prim;
	double* xDataPND; double* xDataNeigh; double* Zij; int32_t* nNeighWall;
};

struct pmps_handle {
	int device = 0;
	cudaStream_t stream = nullptr;
	KParams K;
	DevView V;
	std::vector<void*> allocs;           // everything cudaMalloc'd
	std::vector<DevMesh> meshes; DevMesh* d_meshes = nullptr;
	// alternate buffers for the permute (same shapes as V's planes)
	DevView A;
	int32_t* perm = nullptr;             // new slot -> old slot
	int32_t* cellCount = nullptr;        // nB+1 (histogram / running scatter offsets)
	int32_t* arbRank = nullptr; int32_t* tmpHost = nullptr; int32_t* newCell = nullptr;
	int32_t* scanBlock = nullptr; int nScanBlocks = 0;
	void* stage = nullptr; size_t stageBytes = 0;    // device staging for upload/download (host order)
	void* pinned = nullptr; size_t pinnedBytes = 0;
	int32_t* h_flags = nullptr;          // pinned mirror of V.flags
	int numParticles = 0, numParticlesZero = 0;
	bool haveState = false, haveList = false, listBuilt = false;   // haveList: bucket order valid; listBuilt: Verlet list valid
	int nTilesAll = 0, smCount = 148, walkSeq = 0; bool tileWallValid = false;
	double dispNow = 0.0, dispInc = 0.0, dispIncPrev = 0.0; int stepsSinceBuild = 0; bool needRebuild = true;   // displacement since the list build (step_begin's rebuild decision)
	double lastShift = 0.0;              // largest shift of the last step's shifting passes
	double dispMid = 0.0, tailMove = 0.0, hitDeltaFixed = -1.0;   // one GPU: displacement right after the last mid-step check; largest move of a step's phases after it; PMPS_HIT_DELTA
	double* slabWork = nullptr; pmps_sum_fn slabSum = nullptr; pmps_refresh_fn slabRefresh = nullptr; void* slabUser = nullptr;   // pmps_slab_set_hooks
	double meuMax = 0.0;                 // max MEU of the last rheology pass (CFLvisc)
	cudaStream_t ioStream = nullptr; cudaEvent_t evIoUp = nullptr, evIoMid = nullptr; void* stage2 = nullptr;   // pmps_step_host: copies overlapped with the step
	bool deferSync = false;              // download_impl: leave the stream running (the caller synchronises later)
	bool slabMidChecks = false;          // slab mode: the driver takes the rebuild decision after the prediction (pmps_slab_mid_check)
	bool slab = false;                   // slab mode: rebuilds and the step-end scalars are decided by the caller from GLOBAL maxima (pmps_slab.cu)
	bool shiftListOk = false;            // the shifting passes of this step followed a complete hit list
	bool hitForce = false;               // positions were replaced from outside (upload): the next check rebuilds the hit list
	int optset = 0;                      // option set of the walker kernels this handle runs (pmps_optset.h), chosen by pmps_create
	bool staleReplay = true;             // replay the reference's stale bucket lists after an out-of-domain swap (pmps_set_stale_list_replay)
	bool hitValid = false;               // hlist matches the current positions (set by the list build / k_tv_hits, cleared when positions change)
	size_t listAllocated = 0;
	int solverIter = 0; double solverError = 0.0, velMax = 0.0;
	long long launches = 0;
	long long nWrapped = 0, nWrappedCleared = 0;   // particles moved through a periodic face so far; the count at the last clearing of their hidden flags
	long long nRebuilds = 0, nStepsDone = 0;   // pmps_get_counters: bucket-order rebuilds (sort + Verlet list) and fused steps so far
	double guardWorst = 0.0;             // sticky: largest displacement since a rebuild that exceeded the search slack (0: never); cleared by pmps_upload
	int oodTotal = 0, maxCell = 0;
	std::string err;
	std::vector<std::pair<void*, size_t>> registered;   // host ranges pinned with cudaHostRegister
	// PPE workspace
	uint32_t* ppe_col = nullptr; uint16_t* ppe_cnt = nullptr;
	double* ppe_bicg[5] = { nullptr, nullptr, nullptr, nullptr, nullptr };   // r0, v, s, y, z of the BiCGSTAB solver
	double *ppe_val = nullptr, *ppe_diag = nullptr, *ppe_r = nullptr, *ppe_p = nullptr, *ppe_t = nullptr;
	double *ppe_partial = nullptr, *ppe_scal = nullptr, *ppe_host = nullptr;
	// profiling
	bool profile = false;
	std::vector<std::string> profNames; std::vector<float> profMs; std::vector<int> profLaunches;
	std::vector<const char*> profNamePtrs;
	cudaEvent_t evA = nullptr, evB = nullptr;
};

extern thread_local std::string g_create_error;
int pmps_fail(pmps_handle* h, int code, const char* fmt, ...);
#define CUDA_TRY(h, call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return pmps_fail((h), PMPS_E_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); } while (0)

static inline int cdiv(long long a, int b) { return (int)((a + b - 1) / b); }

// Launch helper: counts launches and (optionally) accumulates per-kernel CUDA-event time on the step stream.
struct ProfScope {
	pmps_handle* h; int idx;
	ProfScope(pmps_handle* h_, const char* name);
	~ProfScope();
};
#define LAUNCH(h, name, kernel, grid, block, ...) do { ProfScope ps_((h), name); kernel<<<(grid), (block), 0, (h)->stream>>>(__VA_ARGS__); (h)->launches++; } while (0)

// implemented in the .cu files
int pmps_ensure_capacity(pmps_handle* h, int n0);
int pmps_sort_and_list(pmps_handle* h);
int pmps_rebuild_now(pmps_handle* h);   // bucket order + work items + hit list of the current positions, in the middle of a step
int pmps_ensure_list(pmps_handle* h);
int pmps_tv_prepare(pmps_handle* h);
int pmps_tv_hits(pmps_handle* h);
int pmps_collide_lists(pmps_handle* h);   // pmps_collide.cu: the PC collision model from the close-pair lists (dvel + vel2 = vel + dvel)
void pmps_set_hit_delta(pmps_handle* h, double delta);
int pmps_ppe_solve(pmps_handle* h);
int pmps_step_end_host(pmps_handle* h);   // step_end of the fused step + the time advance (pmps_stages.inc), for pmps_step_host
int pmps_step_end_apply(pmps_handle* h, double disp2, double shift2);   // pmps_stages.inc: the host half of step_end
int pmps_mesh_upload(pmps_handle* h, int mesh_id, int nV, const double* V, int nF, const int32_t* F, const double* N,
                     int nTable, const double* xP, const double* xN, const double* Z, const int32_t* nW);
