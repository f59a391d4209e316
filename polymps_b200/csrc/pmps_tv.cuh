// pmps_tv.cuh — the tile-Verlet walker: every neighbour loop of the reference as one kernel family.
//
// Buckets are grouped in TILES of TX x TY x TZ buckets; the sort puts the particles of a tile in consecutive slots
// (cell_index()).  A thread block takes one tile and copies what its particles can possibly need — the particles of the
// (TX+2)(TY+2)(TZ+2) surrounding buckets, the HALO — into shared memory with coalesced, independent loads: a float
// position relative to the tile origin (+ the type word), the exact FP64 position, and the neighbour fields the fused
// bodies declare (NbFields).  Candidates come from a per-step Verlet list (k_tv_build) whose entries are 16-bit
// indices INTO THAT STAGED HALO, so the inner loop never touches global memory:
//   scan  : TV_G lanes per particle; lane g tests candidates g, g+TV_G, ... of a 64-entry chunk with the float positions
//           (the threshold carries the float error) and the survivors of the lanes are compacted into the particle's
//           queue with a full-warp ballot (control flow is kept warp-uniform so the ballots need no partial masks:
//           ncu showed partial-mask votes costing a quarter of all issued instructions);
//   pairs : the lanes drain the queue round-robin: exact FP64 distance, the reference's predicates, the stage bodies;
//   end   : partial sums are combined with shuffles, the first lane runs the body epilogues.
// Consecutive candidates of a list are (nearly) consecutive staged records, so the lanes of a particle read one
// contiguous stretch of shared memory per scan step.
//
// The list is valid while no particle has moved farther than (listRadius - re)/2 since the build (tracked, checked in
// pmps_get_scalars); listRadius = bucketSide is the reference's own search slack (src/MpsBucket.cpp:23).
#pragma once
#include "pmps_bodies.cuh"

#ifndef TV_T
#define TV_T 256                 // threads per block
#endif
// TV_G (lanes per particle in the walker) is defined in pmps_device.cuh next to the group reductions
#define TV_NG (TV_T / TV_G)      // particles per round
#define TV_BG 8                  // lanes per particle in the list build
#define TV_HMAX 160              // halo buckets a work item may have
#define TV_CHUNK 64              // list entries per chunk: walker lane g holds entries g, g+TV_G, ... of the chunk, contiguously
#define TV_PER (TV_CHUNK / TV_G) // entries per lane and chunk
#define TV_MAXOWN 512            // particles a work item may own (ownMap, sCnt, sHas)
#define TV_SMEM_BUDGET (111 * 1024)   // dynamic shared memory per block: two blocks per SM with their static part (< 2 KB) and the 1 KB the driver reserves per block
                                      // (228 KB per SM; walk() in pmps_walk.cuh checks the resident blocks of every pass against TV_BLOCKS)
#define TV_LROW (TV_CHUNK * 2 + 16)   // bytes per prefetched list row in shared memory: 16 bytes of padding stagger the lane groups over the banks
// fixed dynamic shared memory of the walker besides the staged records: one round of list rows (each warp fetches the rows of its
// own lane groups, so a round needs no block barrier and one buffer is enough), ownMap, sCnt, sHas
#define TV_FIXED_BYTES (TV_NG * TV_LROW + TV_MAXOWN * 2 + TV_MAXOWN * 2 + TV_MAXOWN)
// position of list entry e inside the particle's list: chunk-major, then lane-major
__host__ __device__ __forceinline__ int tv_list_pos(int e) { const int w = e & (TV_CHUNK - 1); return (e & ~(TV_CHUNK - 1)) + (w % TV_G) * TV_PER + w / TV_G; }

enum { NF_VEL = 1, NF_PRESS = 2, NF_PNDI = 4, NF_RHO = 8, NF_MEU = 16, NF_CONC = 32, NF_PNDKI = 64 };
// neighbour fields each body reads in pair() (positions and the type word are always staged)
template <class B> struct NbFields { static constexpr unsigned value = 0; };
template <> struct NbFields<BViscosity> { static constexpr unsigned value = NF_VEL | NF_MEU; };
template <> struct NbFields<BWallViscosity> { static constexpr unsigned value = NF_VEL | NF_MEU; };
template <> struct NbFields<BCollision> { static constexpr unsigned value = NF_VEL; };
template <> struct NbFields<BCollisionDPC> { static constexpr unsigned value = NF_VEL | NF_PRESS; };
template <> struct NbFields<BPndDiffusive> { static constexpr unsigned value = NF_VEL | NF_PRESS | NF_PNDI | NF_RHO; };
template <> struct NbFields<BWallPndDiffusive> { static constexpr unsigned value = NF_VEL | NF_PNDI; };
template <> struct NbFields<BWallVelDivergence> { static constexpr unsigned value = NF_VEL | NF_PNDI; };
template <> struct NbFields<BVelDivergence> { static constexpr unsigned value = NF_VEL | NF_PNDI; };
template <> struct NbFields<BMeanPnd<0>> { static constexpr unsigned value = NF_PNDI; };
template <> struct NbFields<BMeanPnd<1>> { static constexpr unsigned value = NF_PNDI; };
template <> struct NbFields<BMeanWallPnd> { static constexpr unsigned value = NF_PNDI; };
template <> struct NbFields<BMeanNeighFluidPnd> { static constexpr unsigned value = NF_PNDKI; };
template <> struct NbFields<BPressMin> { static constexpr unsigned value = NF_PRESS; };
template <> struct NbFields<BPressGradient> { static constexpr unsigned value = NF_PRESS | NF_PNDI; };
template <> struct NbFields<BWallPressGradient> { static constexpr unsigned value = NF_PRESS | NF_PNDI; };
template <> struct NbFields<BPredPressGradient> { static constexpr unsigned value = NF_PRESS | NF_PNDI; };
template <> struct NbFields<BPredWallPressGradient> { static constexpr unsigned value = NF_PRESS | NF_PNDI; };
template <> struct NbFields<BGradConcentration> { static constexpr unsigned value = NF_CONC; };
template <> struct NbFields<BWallGradConcentration> { static constexpr unsigned value = NF_CONC; };
template <> struct NbFields<BExtrapolatePress> { static constexpr unsigned value = NF_PRESS | NF_RHO; };
template <> struct NbFields<BWallDummyVel> { static constexpr unsigned value = NF_VEL; };
template <> struct NbFields<BShiftVel> { static constexpr unsigned value = NF_VEL; };
template <> struct NbFields<BStrainRate> { static constexpr unsigned value = NF_VEL | NF_PRESS; };
template <> struct NbFields<BWallStrainRate> { static constexpr unsigned value = NF_VEL | NF_PRESS; };
template <class... B> struct PackFields { static constexpr unsigned value = 0; };
template <class B, class... R> struct PackFields<B, R...> { static constexpr unsigned value = NbFields<B>::value | PackFields<R...>::value; };
// bodies that act only on particles near a polygon wall: their launches skip tiles without any (tileWall)
template <class B, class = void> struct WallOnly { static constexpr bool value = false; };
template <class B> struct WallOnly<B, typename B::wall_frame_tag> { static constexpr bool value = true; };
template <class... B> struct PackWallOnly { static constexpr bool value = true; };
template <class B, class... R> struct PackWallOnly<B, R...> { static constexpr bool value = (WallOnly<B>::value || NoPairs<B>::value) && PackWallOnly<R...>::value; };
// (tail bodies — NoPairs — still run for every particle of a skipped tile: see the light path of k_tv)
// bodies that act on a sparse subset of the particles (free surface, wall/dummy particles): their launches first ask
// begin() of every particle of the tile, one lane per particle, and skip the staging when nobody wants pairs
template <class B, class = void> struct Sparse { static constexpr bool value = false; };
template <class B> struct Sparse<B, typename B::sparse_tag> { static constexpr bool value = true; };
template <class... B> struct PackSparse { static constexpr bool value = true; };
template <class B, class... R> struct PackSparse<B, R...> { static constexpr bool value = (Sparse<B>::value || NoPairs<B>::value) && PackSparse<R...>::value; };

__host__ __device__ constexpr int popc8(unsigned f) { int n = 0; for (int k = 0; k < 8; k++) n += (f >> k) & 1; return n; }
__host__ __device__ constexpr int tv_rec_bytes(unsigned f) {   // staged bytes per halo record: {x, y} and {z, type word} as two 16-byte planes, then the fields
	return 32 + ((f & NF_VEL) ? 24 : 0) + 8 * popc8(f & ~(unsigned)NF_VEL);
}
#define TV_REC_MAX (32 + 24 + 8 * 3)   // the fattest pack of the step paths: F3 (vel, press, pndi, RHO)
// blocks per SM by the bytes a pack stages per record (shared memory decides: TV_MAXM * rec + TV_FIXED_BYTES per block, 228 KB per SM):
// up to TV_OCC4_REC four blocks (64 registers per thread), up to TV_OCC3_REC three (80 registers), two otherwise (128 registers)
// (measured on the B200, 10 M particles: four / three blocks against two everywhere -28 % on the concentration pass, -14 % on the
// concentration gradient, -12 % on the pressure gradient)
#ifndef TV_OCC4_REC
#define TV_OCC4_REC 32
#endif
#ifndef TV_OCC3_REC
#define TV_OCC3_REC 48
#endif
#define TV_BLOCKS(f) (tv_rec_bytes(f) <= TV_OCC4_REC ? 4 : tv_rec_bytes(f) <= TV_OCC3_REC ? 3 : 2)
// staged records per work item: a compile-time constant, so every plane of the shared-memory image sits at a constant
// offset (no registers spent on plane pointers).  Sized for two blocks per SM of the fattest pack in follow mode.
#define TV_MAXM (((TV_SMEM_BUDGET - TV_FIXED_BYTES) / TV_REC_MAX) & ~3)   // multiple of 4: every plane and the fixed part stay 16-byte aligned
#ifndef TV_ILP2_FAT
#define TV_ILP2_FAT 0             // 1: two list entries per trip in the passes that run two blocks per SM (their registers allow it)
#endif
#ifndef TV_ILP2
#define TV_ILP2 0                 // 1: follow passes take two list entries per trip (interleaved geometry chains); measured: no gain
#endif

// ---- work item -> boxes ----------------------------------------------------------------------------------------
struct TvBox {
	int ox, oy, oz;        // bucket coordinates of the first own bucket
	int HX, HY, HZ, HV;    // halo box extents in buckets (own extent + 2; HZ = 1 in 2D) and their product
	int c0, c1;            // own cells [c0, c1)
};
__device__ __forceinline__ TvBox tv_box(const KParams& K, int item) {
	TvBox b; const bool three = K.p.dim == 3;
	int ex, ey, ez;
	if (item >= 0) {
		const int tz = item / (K.nTX * K.nTY), ty = (item - tz * K.nTX * K.nTY) / K.nTX, tx = item - (tz * K.nTY + ty) * K.nTX;
		b.ox = tx * K.TX; b.oy = ty * K.TY; b.oz = tz * K.TZ; ex = K.TX; ey = K.TY; ez = K.TZ; b.c0 = item * K.TV; b.c1 = b.c0 + K.TV;
	} else {
		const int c = ~item; cell_coords(K, c, b.ox, b.oy, b.oz); ex = ey = ez = 1; b.c0 = c; b.c1 = c + 1;
	}
	b.HX = ex + 2; b.HY = ey + 2; b.HZ = three ? ez + 2 : 1; b.HV = b.HX * b.HY * b.HZ;
	return b;
}
// cell of halo bucket hb of the box, or -1 outside the bucket grid
__device__ __forceinline__ int tv_halo_cell(const KParams& K, const TvBox& b, int hb) {
	const int hz = hb / (b.HX * b.HY), hy = (hb - hz * b.HX * b.HY) / b.HX, hx = hb - (hz * b.HY + hy) * b.HX;
	int ix = b.ox - 1 + hx, iy = b.oy - 1 + hy, iz = K.p.dim == 3 ? b.oz - 1 + hz : 0;
	if (ix < 0 || ix >= K.p.numBucketsX || iy < 0 || iy >= K.p.numBucketsY || iz < 0 || iz >= K.p.numBucketsZ) return -1;
	if (K.p.domainTypeBC == 1) per_alias(K, ix, iy, iz);   // periodic: a border layer holds the particles of the opposite inner layer
	return cell_index(K, ix, iy, iz);
}
// periodic domain: what the records of halo bucket hb are moved by along the periodic axis (pmps_device.cuh, per_shift_of_coord)
__device__ __forceinline__ double tv_halo_shift(const KParams& K, const TvBox& b, int hb) {
	const int hz = hb / (b.HX * b.HY), hy = (hb - hz * b.HX * b.HY) / b.HX, hx = hb - (hz * b.HY + hy) * b.HX;
	const int a = K.p.periodicAxis;
	return per_shift_of_coord(K, a == 0 ? b.ox - 1 + hx : a == 1 ? b.oy - 1 + hy : b.oz - 1 + hz);
}
// halo table in shared memory: hBase[hb] first slot, hOff[hb] first staged index (hOff[HV] = staged records)
__device__ __forceinline__ void tv_halo_table(const DevView& S, const KParams& K, const TvBox& b, int* hBase, int* hOff) {
	const int tid = threadIdx.x;
	if (tid < b.HV) {
		const int c = tv_halo_cell(K, b, tid);
		int s0 = 0, n = 0;
		if (c >= 0) { s0 = S.cellStart[c]; n = S.cellStart[c + 1] - s0; }
		hBase[tid] = s0; hOff[tid + 1] = n;
	}
	__syncthreads();
	if (tid < 32) {   // exclusive scan of <= TV_HMAX counts by one warp: 5 consecutive entries per lane
		constexpr int PER = (TV_HMAX + 31) / 32;
		int v[PER]; int sum = 0;
		#pragma unroll
		for (int k = 0; k < PER; k++) { const int e = tid * PER + k; v[k] = e < b.HV ? hOff[e + 1] : 0; sum += v[k]; }
		int incl = sum;
		#pragma unroll
		for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (tid >= o) incl += t; }
		int run = incl - sum;
		__syncwarp();
		#pragma unroll
		for (int k = 0; k < PER; k++) { const int e = tid * PER + k; if (e < b.HV) hOff[e] = run; run += v[k]; if (e == b.HV - 1) hOff[b.HV] = run; }
	}
	__syncthreads();
}
__device__ __forceinline__ float tv_type_word(const double4& p) { return __int_as_float((pk_type(p) & 0xffff) | (pk_ptype(p) << 16)); }

// ---- work items of the current bucket order ---------------------------------------------------------------------
// One thread per tile: a tile whose halo fits the staging area is one item; otherwise each of its non-empty buckets is.
// Every item gets its halo table precomputed (K.tabStride ints: hOff[0..HVt], hBase[0..HVt), own0, P — HVt = halo
// buckets of a whole tile) so the passes read it with one coalesced load instead of chasing cellStart.
#define TV_TABMAX (2 * TV_HMAX + 8)
__device__ __forceinline__ void tv_write_table(const DevView& S, const KParams& K, const TvBox& b, int slot) {
	int* tab = S.itemTab + (size_t)slot * K.tabStride;
	int run = 0;
	for (int hb = 0; hb < b.HV; hb++) {
		const int c = tv_halo_cell(K, b, hb);
		int s0 = 0, n = 0;
		if (c >= 0) { s0 = S.cellStart[c]; n = S.cellStart[c + 1] - s0; }
		tab[hb] = run; tab[K.HVt + 1 + hb] = s0; run += n;
	}
	tab[b.HV] = run;
	const int own0 = S.cellStart[b.c0];
	tab[2 * K.HVt + 1] = own0; tab[2 * K.HVt + 2] = S.cellStart[b.c1] - own0;
	tab[2 * K.HVt + 3] = b.ox; tab[2 * K.HVt + 4] = b.oy; tab[2 * K.HVt + 5] = b.oz;   // (the passes read the box from here: no divisions per item)
}
// the box of a work item from its table in shared memory
__device__ __forceinline__ TvBox tv_box_tab(const KParams& K, int item, const int* tab) {
	TvBox b; const bool three = K.p.dim == 3, whole = item >= 0;
	b.ox = tab[2 * K.HVt + 3]; b.oy = tab[2 * K.HVt + 4]; b.oz = tab[2 * K.HVt + 5];
	const int ex = whole ? K.TX : 1, ey = whole ? K.TY : 1, ez = whole ? K.TZ : 1;
	b.c0 = whole ? item * K.TV : ~item; b.c1 = whole ? b.c0 + K.TV : b.c0 + 1;
	b.HX = ex + 2; b.HY = ey + 2; b.HZ = three ? ez + 2 : 1; b.HV = b.HX * b.HY * b.HZ;
	return b;
}
// Threads enumerate the tiles along a Z-order curve (bits of the thread index dealt round-robin to the axes that still have bits
// left), so the work items come out in roughly that order: the blocks working at any one time — they take items from a cursor —
// then sit in a compact 3-D neighbourhood and find each other's halo records in L2 (a raster order re-reads every record from
// DRAM for the tiles of the next z-layer: ncu, 2.1 - 4.1 x the algorithmic bytes in round 1).
#ifdef PMPS_TV_LISTBUILD
static __global__ void k_tv_items(const __grid_constant__ DevView S, const __grid_constant__ KParams K, int nCodes, int bx, int by, int bz) {
	const int code = blockIdx.x * blockDim.x + threadIdx.x;
	if (code >= nCodes) return;
	int tx = 0, ty = 0, tz = 0;
	{ int c = code, ax = 0, ay = 0, az = 0;
	  while (ax < bx || ay < by || az < bz) {
		if (ax < bx) { tx |= (c & 1) << ax; c >>= 1; ax++; }
		if (ay < by) { ty |= (c & 1) << ay; c >>= 1; ay++; }
		if (az < bz) { tz |= (c & 1) << az; c >>= 1; az++; }
	  } }
	if (tx >= K.nTX || ty >= K.nTY || tz >= K.nTZ) return;
	const int t = (tz * K.nTY + ty) * K.nTX + tx;
	if (S.cellStart[(t + 1) * K.TV] == S.cellStart[t * K.TV]) return;
	const TvBox b = tv_box(K, t);
	int M = 0;
	for (int hb = 0; hb < b.HV; hb++) { const int c = tv_halo_cell(K, b, hb); if (c >= 0) M += S.cellStart[c + 1] - S.cellStart[c]; }
	if (M + 1 <= K.tvMaxM) { const int slot = atomicAdd(S.nTiles, 1); S.tileList[slot] = t; tv_write_table(S, K, b, slot); return; }
	for (int c = b.c0; c < b.c1; c++) if (S.cellStart[c + 1] > S.cellStart[c]) {
		const int slot = atomicAdd(S.nTiles, 1); S.tileList[slot] = ~c; tv_write_table(S, K, tv_box(K, ~c), slot);
	}
}
#endif
// item scheduling shared by the tile kernels: a work-stealing cursor whose NEXT value is requested at the start of an
// item, so the atomic's latency is hidden behind the item's work
__device__ __forceinline__ int tv_first_item(const DevView& S, int cursorIdx, int* sItem) {
	if (threadIdx.x == 0) *sItem = atomicAdd(&S.tileCursor[cursorIdx], 1);
	__syncthreads();
	return *sItem;
}
__device__ __forceinline__ int tv_advance(int nxt, int* sItem) {
	__syncthreads();   // the item is done with the tables and the staged data
	if (threadIdx.x == 0) *sItem = nxt;
	__syncthreads();
	return *sItem;
}
template <int NT = TV_T> __device__ __forceinline__ void tv_load_table(const DevView& S, const KParams& K, int tt, int* sTab) {
	for (int k = threadIdx.x; k < K.tabStride; k += NT) sTab[k] = S.itemTab[(size_t)tt * K.tabStride + k];
	__syncthreads();
}
static __global__ void k_copy_ghost_tail(const __grid_constant__ DevView S, int copyPos, int copyVel) {
	const int i = S.cellStart[S.nC] + blockIdx.x * blockDim.x + threadIdx.x;   // ghosts are parked behind the last cell
	if (i >= S.n0) return;
	if (copyPos) S.pos2[i] = S.pos[i];
	if (copyVel) S.vel2[i] = S.vel[i];
}


// ---- shared-memory image of the halo ------------------------------------------------------------------------------
#define TV_W_HIDDEN 0x40000000   // bit of the float type word (list build only): the record is found by nobody (a particle the periodic wrap moved in this step, DevView::wrapHid)
#define TV_HIDMAX 32   // staged records of one work item the reference's stale bucket lists may hide (pmps_device.cuh, hid_visible)
struct TvStage {   // shared-memory planes of the staged halo: the walker stages {x, y} / {z, type word} pairs of doubles (two 16-byte planes: a candidate
	// is two 128-bit loads, and 8 consecutive records cover all 32 banks), the list build float positions
	float4* fpos; double2 *pxy, *pzw; Nb nb;
};
// Hidden records of a work item (rare: only in a step in which a particle left the domain; pmps_device.cuh, hid_visible): staged
// index and the bucket the record's array element is listed in.  Only the hit-list build looks at them (k_hl_build): a particle whose
// search does not reach that bucket does not get the record into its list, and the list is re-made at every position epoch while
// hidden entries exist (k_hits_decide) — the walkers carry no test.
struct TvHid { int n; int m[TV_HIDMAX]; int cell[TV_HIDMAX]; };
template <bool B> struct TvBool { static constexpr bool value = B; };
template <unsigned FIELDS> __device__ __forceinline__ TvStage tv_carve(unsigned char* sp) {
	TvStage t; constexpr size_t m = TV_MAXM;
	t.fpos = nullptr;
	t.pxy = (double2*)sp; sp += 16 * m; t.pzw = (double2*)sp; sp += 16 * m;
	double *vx = nullptr, *vy = nullptr, *vz = nullptr, *pr = nullptr, *pn = nullptr, *rh = nullptr, *mu = nullptr, *cc = nullptr, *pk = nullptr;
	if (FIELDS & NF_VEL) { vx = (double*)sp; sp += 8 * m; vy = (double*)sp; sp += 8 * m; vz = (double*)sp; sp += 8 * m; }
	if (FIELDS & NF_PRESS) { pr = (double*)sp; sp += 8 * m; }
	if (FIELDS & NF_PNDI) { pn = (double*)sp; sp += 8 * m; }
	if (FIELDS & NF_RHO) { rh = (double*)sp; sp += 8 * m; }
	if (FIELDS & NF_MEU) { mu = (double*)sp; sp += 8 * m; }
	if (FIELDS & NF_CONC) { cc = (double*)sp; sp += 8 * m; }
	if (FIELDS & NF_PNDKI) { pk = (double*)sp; sp += 8 * m; }
	t.nb.vx = vx; t.nb.vy = vy; t.nb.vz = vz; t.nb.press = pr; t.nb.pndi = pn; t.nb.RHO = rh; t.nb.MEU = mu; t.nb.concentration = cc; t.nb.pndki = pk;
	return t;
}


// ---- staging the halo ------------------------------------------------------------------------------------------------
// One warp per halo bucket, one lane per record.  Everything the WALKER stages travels with cp.async (global -> shared, no registers,
// nothing waits until the end): the 32-byte position record as two 16-byte copies into the {x, y} and {z, w} planes, the neighbour
// fields as 8-byte copies into theirs.  (ncu, round 1/2: with the positions going through registers the staging loads and the
// stores waiting for them were a quarter of all stall samples of a light pass.)  The list build stages float positions relative to
// the tile origin and does go through registers, four buckets in flight per thread.
__device__ __forceinline__ void tv_cp8(const double* dstShared, const double* srcGlobal) {   // 8-byte asynchronous global -> shared copy
	asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" :: "r"((unsigned)__cvta_generic_to_shared(dstShared)), "l"(srcGlobal));
}
__device__ __forceinline__ void tv_cp16(const void* dstShared, const void* srcGlobal) {   // 16-byte asynchronous global -> shared copy, L2 only
	asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"((unsigned)__cvta_generic_to_shared(dstShared)), "l"(srcGlobal));
}
template <unsigned FIELDS, int NT = TV_T>
__device__ __forceinline__ void tv_stage(const DevView& S, const TvStage& T, const int* hOff, const int* hBase, int HV, int own0, int P, uint16_t* ownMap) {
	constexpr int NW = NT / 32;
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	for (int hb = warp; hb < HV; hb += NW) {
		const int off = hOff[hb], n = hOff[hb + 1] - off, s0 = hBase[hb];
		for (int r = lane; r < n; r += 32) {
			const int j = s0 + r, m = off + r;
			const double2* pj = (const double2*)(S.pos + j);
			tv_cp16(T.pxy + m, pj); tv_cp16(T.pzw + m, pj + 1);
			if (j >= own0 && j < own0 + P) ownMap[j - own0] = (uint16_t)m;
			if (FIELDS & NF_VEL) { const double* v = (const double*)(S.vel + j); tv_cp8(T.nb.vx + m, v); tv_cp8(T.nb.vy + m, v + 1); tv_cp8(T.nb.vz + m, v + 2); }
			if (FIELDS & NF_PRESS) tv_cp8(T.nb.press + m, S.press + j);
			if (FIELDS & NF_PNDI) tv_cp8(T.nb.pndi + m, S.pndi + j);
			if (FIELDS & NF_RHO) tv_cp8(T.nb.RHO + m, S.RHO + j);
			if (FIELDS & NF_MEU) tv_cp8(T.nb.MEU + m, S.MEU + j);
			if (FIELDS & NF_CONC) tv_cp8(T.nb.concentration + m, S.concentration + j);
			if (FIELDS & NF_PNDKI) tv_cp8(T.nb.pndki + m, S.pndki + j);
		}
	}
	asm volatile("cp.async.wait_all;" ::: "memory");
}
// Periodic domain: the records of the aliased border layers are seen moved by -/+ periodicLength, x_j + pl with its own rounding as
// in the reference (src/MpsParticle.cpp:35-45).  Every thread moves the records it copied itself (same warp -> bucket, lane -> record
// mapping as tv_stage, whose cp.async.wait_all made them visible to it); the block barrier after the staging publishes them.
template <int NT = TV_T>
__device__ __forceinline__ void tv_stage_periodic(const KParams& K, const TvBox& b, const TvStage& T, const int* hOff) {
	constexpr int NW = NT / 32;
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, a = K.p.periodicAxis;
	for (int hb = warp; hb < b.HV; hb += NW) {
		const double sh = tv_halo_shift(K, b, hb);
		if (sh == 0.0) continue;
		const int off = hOff[hb], n = hOff[hb + 1] - off;
		for (int r = lane; r < n; r += 32) {
			const int m = off + r;
			if (a == 0) T.pxy[m].x = __dadd_rn(T.pxy[m].x, sh); else if (a == 1) T.pxy[m].y = __dadd_rn(T.pxy[m].y, sh); else T.pzw[m].x = __dadd_rn(T.pzw[m].x, sh);
		}
	}
}
// float positions (+ type word) of the halo for the list build
template <int NT, bool PER>
__device__ __forceinline__ void tv_stage_float(const DevView& S, const KParams& K, const TvBox& bx, float4* fpos, int32_t* slotOf, const int* hOff, const int* hBase, int HV, double ox, double oy, double oz) {
	constexpr int NB = 4;   // buckets per thread whose position loads are in flight together
	constexpr int NW = NT / 32;
	const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	for (int hb0 = warp; hb0 < HV; hb0 += NW * NB) {
		double4 p[NB]; int m[NB]; bool ok[NB];
		#pragma unroll
		for (int b = 0; b < NB; b++) {
			const int hb = hb0 + b * NW;
			ok[b] = false; m[b] = 0;
			if (hb < HV) { const int off = hOff[hb]; ok[b] = lane < hOff[hb + 1] - off; m[b] = off + lane; if (ok[b]) p[b] = ld256(S.pos + hBase[hb] + lane); }
			if (PER && hb < HV && ok[b]) per_move(K, p[b], tv_halo_shift(K, bx, hb));
		}
		#pragma unroll
		for (int b = 0; b < NB; b++) if (ok[b]) {
			const int slot = hBase[hb0 + b * NW] + lane;
			float w = tv_type_word(p[b]);
			if (PER && S.wrapHid[slot]) w = __int_as_float(__float_as_int(w) | TV_W_HIDDEN);   // a particle the periodic wrap moved in this step: found by nobody
			fpos[m[b]] = make_float4((float)(p[b].x - ox), (float)(p[b].y - oy), (float)(p[b].z - oz), w); slotOf[m[b]] = slot;
		}
		#pragma unroll
		for (int b = 0; b < NB; b++) {   // buckets with more than 32 particles (far above the lattice density)
			const int hb = hb0 + b * NW;
			if (hb >= HV) continue;
			const int off = hOff[hb], n = hOff[hb + 1] - off, s0 = hBase[hb];
			const double sh = PER ? tv_halo_shift(K, bx, hb) : 0.0;
			for (int r = lane + 32; r < n; r += 32) {
				double4 q = ld256(S.pos + s0 + r); if (PER) per_move(K, q, sh);
				float w = tv_type_word(q); if (PER && S.wrapHid[s0 + r]) w = __int_as_float(__float_as_int(w) | TV_W_HIDDEN);
				fpos[off + r] = make_float4((float)(q.x - ox), (float)(q.y - oy), (float)(q.z - oz), w); slotOf[off + r] = s0 + r;
			}
		}
	}
}

// ---- hit list in tile-local indices ------------------------------------------------------------------------------------
// ONE kernel makes the candidate list of every neighbour pass: for each own particle of a work item, the staged records within
// hitR = max(reS, reL) + 2 hitDelta at the current positions, as 16-bit indices into the staged halo (same tiling and staging
// order as the walker, so the indices mean the same records there).  It runs after a bucket rebuild (force = 1) and whenever
// k_hits_decide finds that somebody has moved hitDelta since the list was made (positions change three times per step).
//
// One LANE per own particle: the lane sweeps the 27 (9 in 2D) buckets around its particle's bucket — nine (three) contiguous
// stretches of the staged float positions — and appends the hits to its own row in shared memory.  The lanes of a warp are
// particles of the same one or two buckets, so they read the same staged record at the same time (a shared-memory broadcast),
// there are no votes and no cross-lane compaction; a row is an odd number of 32-bit words long, so the lanes' appends fall
// into different banks.  The rows are then copied to hlist chunk by chunk (128 contiguous bytes per particle) in the lane-major
// order the walker reads (tv_list_pos); the unused tail of the last chunk points at the far-away dummy record M.
// 197 warp-instructions per particle against 1 218 (two-level build: Verlet list + hit list) + 197 (refilter) of the
// round-1 kernels (ncu, profiles/r02_ncu_*): the Verlet list of radius bucketSide is gone.
#define HL_T 128
#define HL_ROW (TV_CHUNK + 2)   // 16-bit entries per shared-memory row: the first chunk of a list ((HL_ROW / 2) is odd: lanes fall into different banks)
#ifdef PMPS_TV_LISTBUILD   // (compiled only where it is launched: pmps_sort.cu)
template <bool PER>   // PER: periodic domain (aliased border layers, their records moved by -/+ periodicLength)
static __global__ void __launch_bounds__(HL_T, 5) k_hl_build(const __grid_constant__ DevView S, const __grid_constant__ KParams K, int cursorIdx, int force) {
	if (!force && !S.hitFlag[0]) return;   // k_hits_decide: the hit list still covers the current positions
	extern __shared__ __align__(16) unsigned char smemRaw[];
	__shared__ int sTab[TV_TABMAX], sItem;
	__shared__ unsigned sMinBits;   // smallest squared pair distance of the item (bits of a positive float)
	__shared__ TvHid sHid;          // hidden records of the item (only in a step in which a particle left the domain)
	const int nHidAll = S.hid[0];   // uniform over the grid
	float4* const fpos = (float4*)smemRaw;
	const int cap = S.tlCap;
	const int tid = threadIdx.x, lane = tid & 31, w0 = tid & ~31;
	uint16_t* const rows = (uint16_t*)(smemRaw + 16 * (size_t)(TV_MAXM + 4));
	uint16_t* const row = rows + (size_t)tid * HL_ROW;
	int32_t* const slotOf = (int32_t*)(smemRaw + 16 * (size_t)(TV_MAXM + 4) + (size_t)HL_T * HL_ROW * 2);   // staged index -> global slot (close-pair lists)
	const float C2 = K.closeR2;
	const bool three = K.p.dim == 3;
	const int nItems = *S.nTiles;
	const float R2 = K.hitR2;
	const int* hOff = sTab; const int* hBase = sTab + K.HVt + 1;
	if (blockIdx.x == 0 && tid == 0) S.hitMove[1] = (float)K.hitDelta;   // the margin THIS list was made with (k_hits_decide)
	for (int tt = tv_first_item(S, cursorIdx, &sItem), nxt = 0; tt < nItems; tt = tv_advance(nxt, &sItem)) {
		if (tid == 0) nxt = atomicAdd(&S.tileCursor[cursorIdx], 1);
		const int item = S.tileList[tt];
		tv_load_table<HL_T>(S, K, tt, sTab);
		const TvBox b = tv_box_tab(K, item, sTab);
		if (tid == 0) { sMinBits = 0x7f800000u; S.itemMinD2[tt] = 0.f; S.tileClose[item >= 0 ? item : (~item) / K.TV] = 1; }
		const int M = hOff[b.HV], own0 = sTab[2 * K.HVt + 1], P = sTab[2 * K.HVt + 2];
		if (M + 1 > TV_MAXM || P > TV_MAXOWN) {   // a single bucket whose 27-bucket neighbourhood does not fit: reported by pmps_get_scalars
			if (tid == 0) atomicMax(&S.flags[FLAG_TILE_OVERFLOW], M + 1 > TV_MAXM ? M + 1 : P);
			for (int p = tid; p < P; p += HL_T) S.hcnt[own0 + p] = 0;
			continue;
		}
		const double ox = K.p.domainMinX + (double)(b.ox - 1) * K.p.bucketSide, oy = K.p.domainMinY + (double)(b.oy - 1) * K.p.bucketSide,
		             oz = three ? K.p.domainMinZ + (double)(b.oz - 1) * K.p.bucketSide : 0.0;
		tv_stage_float<HL_T, PER>(S, K, b, fpos, slotOf, hOff, hBase, b.HV, ox, oy, oz);
		if (tid == 0) fpos[M] = make_float4(1e18f, 1e18f, 1e18f, 0.f);   // the far-away dummy record: what a lane reads beyond the end of its stretch
		if (nHidAll > 0) {   // which staged records of this item are hidden entries (S.hid)?  One thread per entry.
			if (tid == 0) sHid.n = 0;
			__syncthreads();
			for (int e = tid; e < nHidAll; e += HL_T) {
				const int label = S.hid[1 + e];
				if (label < 0) continue;
				const int slot = S.slotOfHost[label], c = S.cellOf[slot];
				if (c >= S.nC) continue;
				int cc[3]; cell_coords(K, c, cc[0], cc[1], cc[2]);
				// (periodic domain: the record is staged where its bucket is and where the border layer that aliases its bucket is)
				const int nAx = K.p.periodicAxis == 0 ? K.p.numBucketsX : K.p.periodicAxis == 1 ? K.p.numBucketsY : K.p.numBucketsZ;
				const int alias = !PER ? -1 : cc[K.p.periodicAxis] == nAx - 2 ? 0 : cc[K.p.periodicAxis] == 1 ? nAx - 1 : -1;
				for (int v = 0; v < (alias >= 0 ? 2 : 1); v++) {
					if (v == 1) cc[K.p.periodicAxis] = alias;
					const int gx = cc[0] - b.ox + 1, gy = cc[1] - b.oy + 1, gz = three ? cc[2] - b.oz + 1 : 0;
					if (gx < 0 || gx >= b.HX || gy < 0 || gy >= b.HY || gz < 0 || gz >= b.HZ) continue;
					const int hb = (gz * b.HY + gy) * b.HX + gx, r = slot - hBase[hb];
					if (r < 0 || r >= hOff[hb + 1] - hOff[hb]) continue;
					const int k = atomicAdd(&sHid.n, 1);
					if (k < TV_HIDMAX) { sHid.m[k] = hOff[hb] + r; sHid.cell[k] = S.hid[1 + PMPS_HID_CAP + e]; }
					else atomicOr(&S.flags[FLAG_GUARD], 8);
				}
			}
		}
		__syncthreads();
		const int nHid = nHidAll > 0 ? min(sHid.n, TV_HIDMAX) : 0;   // uniform over the block
		// (the particle loop exists twice: with the hidden-record test — only in a step in which a particle left the domain — and without)
		auto sweep = [&](auto hidTag) {
		constexpr bool HID = decltype(hidTag)::value;
		for (int p0 = 0; p0 < P; p0 += HL_T) {   // uniform over the block
			const int p = p0 + tid;
			const bool act = p < P;
			const int i = own0 + (act ? p : 0);
			{	// every entry of the row points at the dummy record until a hit overwrites it
				uint32_t* r32 = (uint32_t*)row; const uint32_t pad = (uint32_t)M * 0x10001u;
				#pragma unroll
				for (int q = 0; q < TV_CHUNK / 2; q++) r32[q] = pad;
			}
			int hx = 1, hy = 1, hz = three ? 1 : 0, self = -1;
			float4 fi = make_float4(0.f, 0.f, 0.f, 0.f);
			if (act) {
				int ix, iy, iz; cell_coords(K, S.cellOf[i], ix, iy, iz);
				hx = ix - b.ox + 1; hy = iy - b.oy + 1; hz = three ? iz - b.oz + 1 : 0;
				const int hbs = (hz * b.HY + hy) * b.HX + hx;
				self = hOff[hbs] + (i - hBase[hbs]);
				fi = fpos[self];
			}
			// the sweep: hits are appended to the row in the order found (the copy-out below applies the lane-major order of the
			// walker); only the first chunk lives in shared memory — a longer list (rare: 64 hits need 1.4 x the lattice density)
			// gets its further entries from a second sweep that writes straight to global memory
			int cnt = 0; float dmin = __int_as_float(0x7f800000);
			const unsigned rowS = (unsigned)__cvta_generic_to_shared(row);
			for (int dz = three ? -1 : 0; dz <= (three ? 1 : 0); dz++) for (int dy = -1; dy <= 1; dy++) {
				const int hb0 = ((hz + dz) * b.HY + hy + dy) * b.HX + hx - 1;
				const int r0 = hOff[hb0], r1 = act ? hOff[hb0 + 3] : r0;
				const int trips = __reduce_max_sync(0xffffffffu, r1 - r0);
				for (int k = 0; k < trips; k += 4) {
					// four candidates at a time: the loads first (independent), then the four tests, then the appends
					int m[4]; float4 f[4]; float d2[4];
					#pragma unroll
					for (int u = 0; u < 4; u++) { const int mm = r0 + k + u; m[u] = (mm < r1 && mm != self) ? mm : M; f[u] = fpos[m[u]]; }
					#pragma unroll
					for (int u = 0; u < 4; u++) { const float dx = f[u].x - fi.x, dy2 = f[u].y - fi.y, dz2 = f[u].z - fi.z; d2[u] = dx * dx + dy2 * dy2 + dz2 * dz2; }
					dmin = fminf(fminf(dmin, fminf(d2[0], d2[1])), fminf(d2[2], d2[3]));
					if (HID) {   // (rare) a hidden record is a hit only for a particle whose search reaches the bucket it is listed in
						#pragma unroll
						for (int u = 0; u < 4; u++) if (d2[u] < R2) {
							for (int q = 0; q < nHid; q++) if (sHid.m[q] == m[u] && !hid_visible(K, sHid.cell[q], S.pos[i])) d2[u] = 3e38f;
						}
					}
					#pragma unroll
					for (int u = 0; u < 4; u++) {   // (the row address is kept in a register by hand: under the 80-register cap ptxas re-derives it from %tid for every store)
						const bool hit = d2[u] < R2 && !(PER && (__float_as_int(f[u].w) & TV_W_HIDDEN));
						if (hit && cnt < TV_CHUNK) asm volatile("st.shared.u16 [%0], %1;" :: "r"(rowS + 2u * (unsigned)cnt), "h"((unsigned short)m[u]) : "memory");
						cnt += hit ? 1 : 0;
					}
				}
			}
			// close pairs (the collision pass needs nothing else, pmps_collide.cu): the first-chunk hits within closeR, as GLOBAL slots
			int nClose = 0;
			if (act && (int)(short)(__float_as_int(fi.w) & 0xffff) == K.p.fluid) {
				const int ne = min(cnt, TV_CHUNK);
				for (int e = 0; e < ne; e++) {
					const int mm = row[e]; const float4 f = fpos[mm];
					const float dx = f.x - fi.x, dy2 = f.y - fi.y, dz2 = f.z - fi.z;
					if (dx * dx + dy2 * dy2 + dz2 * dz2 < C2) { if (nClose < PMPS_CLOSE_CAP) S.clist[(size_t)i * PMPS_CLOSE_CAP + nClose] = (uint32_t)slotOf[mm]; nClose++; }
				}
			}
			if (cnt > TV_CHUNK) {   // second sweep (divergent, rare): entries 64, 65, ... in the walker's order, then the padding of the last chunk
				uint16_t* const grow = S.hlist + (size_t)i * (size_t)cap;
				int c2 = 0;
				for (int dz = three ? -1 : 0; dz <= (three ? 1 : 0); dz++) for (int dy = -1; dy <= 1; dy++) {
					const int hb0 = ((hz + dz) * b.HY + hy + dy) * b.HX + hx - 1;
					for (int mm = hOff[hb0]; mm < hOff[hb0 + 3]; mm++) {
						if (mm == self) continue;
						const float4 f = fpos[mm]; const float dx = f.x - fi.x, dy2 = f.y - fi.y, dz2 = f.z - fi.z;
						bool hit = dx * dx + dy2 * dy2 + dz2 * dz2 < R2 && !(PER && (__float_as_int(f.w) & TV_W_HIDDEN));
						if (HID && hit) for (int q = 0; q < nHid; q++) if (sHid.m[q] == mm && !hid_visible(K, sHid.cell[q], S.pos[i])) hit = false;
						if (hit) {
							if (c2 >= TV_CHUNK && c2 < cap) grow[tv_list_pos(c2)] = (uint16_t)mm;
							if (c2 >= TV_CHUNK && dx * dx + dy2 * dy2 + dz2 * dz2 < C2) { if (nClose < PMPS_CLOSE_CAP) S.clist[(size_t)i * PMPS_CLOSE_CAP + nClose] = (uint32_t)slotOf[mm]; nClose++; }
							c2++;
						}
					}
				}
				for (int e = min(cnt, cap); (e & (TV_CHUNK - 1)) != 0; e++) grow[tv_list_pos(e)] = (uint16_t)M;
			}
			if (act) {
				if (cnt > cap) { atomicMax(&S.flags[FLAG_OVERFLOW], cnt); cnt = cap; }
				S.hcnt[i] = (uint16_t)cnt; S.ccnt[i] = (uint8_t)(nClose > PMPS_CLOSE_CAP ? 255 : nClose);
				S.hpos[i] = make_float4(fi.x + (float)(ox - K.p.domainMinX), fi.y + (float)(oy - K.p.domainMinY), fi.z + (float)(oz - (three ? K.p.domainMinZ : 0.0)), 0.f);
			} else cnt = 0;
			__syncwarp();
			// first chunks: rows -> hlist in the lane-major order (entry e of the list sits at tv_list_pos(e)): destination word `lane`
			// holds positions 2 lane and 2 lane + 1 = entries e0 and e0 + TV_G; one particle's 128 contiguous bytes per instruction
			const int e0 = ((2 * lane) % TV_PER) * TV_G + (2 * lane) / TV_PER;
			for (int q = 0; q < 32; q++) {
				const int pq = p0 + w0 + q;
				if (pq >= P) break;   // uniform over the warp
				if (__shfl_sync(0xffffffffu, cnt, q) == 0) continue;   // uniform
				const uint16_t* src = rows + (size_t)(w0 + q) * HL_ROW;
				uint32_t* dst = (uint32_t*)(S.hlist + (size_t)(own0 + pq) * (size_t)cap);
				dst[lane] = (uint32_t)src[e0] | ((uint32_t)src[e0 + TV_G] << 16);
			}
			{ const unsigned mb = __reduce_min_sync(0xffffffffu, __float_as_uint(dmin)); if (lane == 0) atomicMin(&sMinBits, mb); }
			__syncwarp();
		}
		};
		if (nHid > 0) sweep(TvBool<true>()); else sweep(TvBool<false>());
		__syncthreads();
		if (tid == 0) { const float md = __uint_as_float(sMinBits); S.itemMinD2[tt] = md; if (item >= 0) S.tileClose[item] = md < K.closeR2 ? 1 : 0; }
	}
}
#endif   // PMPS_TV_LISTBUILD

// ---- passes that act on few tiles ----------------------------------------------------------------------------------
// Wall-only packs touch the tiles with a particle near a polygon wall, close-pair packs (collisions) the tiles that held a
// close pair when the hit list was made.  For the other tiles only the TAILS of the pack run (integrators, buffer copies):
// k_tails does that with one thread per particle, fully coalesced, and k_tv leaves those work items at once.
static __global__ void k_item_flags(const __grid_constant__ DevView S, const __grid_constant__ KParams K) {
	const int tt = blockIdx.x * blockDim.x + threadIdx.x;
	if (tt >= *S.nTiles) return;
	const int item = S.tileList[tt]; const int tile = item >= 0 ? item : (~item) / K.TV;
	if (S.tileWall[tile]) S.actList[atomicAdd(&S.actCount[0], 1)] = tt;
	if (S.tileClose[tile]) S.actList[S.actCap + atomicAdd(&S.actCount[1], 1)] = tt;
	if (S.tileSparse[tile]) S.actList[2 * S.actCap + atomicAdd(&S.actCount[2], 1)] = tt;
}
// sparse packs (free-surface / wall-particle means): which tiles hold a particle that wants pairs?  begin() of these bodies is idempotent.
template <class F>
static __global__ void __launch_bounds__(256) k_sparse_mark(const __grid_constant__ DevView S, const __grid_constant__ KParams K) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= S.n) return;
	const int c = S.cellOf[i];
	if (c >= S.nC) return;
	Self s; s.nb = nullptr; s.ms = 0; s.i = i; s.lead = true; s.nearWallFluid = false; s.hasMirror = false; s.mix = s.miy = s.miz = 0.0;
	s.pi = ld256(S.pos + i); s.type = pk_type(s.pi); s.ptype = pk_ptype(s.pi);
	if (s.type == K.p.ghost) return;
	typename F::Ctx ctx = {};
	if (F::begin(ctx, S, K, s) && F::need_pairs(ctx)) S.tileSparse[c / K.TV] = 1;
}
template <int KIND, class F>   // KIND 0 wall-only, 1 close-pair, 2 sparse
static __global__ void __launch_bounds__(256) k_tails(const __grid_constant__ DevView S, const __grid_constant__ KParams K) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= S.n) return;
	const int c = S.cellOf[i];
	if (c >= S.nC) return;   // parked ghost slots: k_copy_ghost_tail
	const int tile = c / K.TV;
	if ((KIND == 0 ? S.tileWall[tile] : KIND == 1 ? S.tileClose[tile] : S.tileSparse[tile]) != 0) return;   // k_tv does this tile
	Self s; s.nb = nullptr; s.ms = 0; s.i = i; s.lead = true; s.nearWallFluid = false; s.hasMirror = false; s.mix = s.miy = s.miz = 0.0;
	s.pi = ld256(S.pos + i); s.type = pk_type(s.pi); s.ptype = pk_ptype(s.pi);
	if (F::kCopyPos) S.pos2[i] = s.pi;
	if (F::kCopyVel) S.vel2[i] = S.vel[i];
	if (s.type == K.p.ghost) return;
	typename F::Ctx ctx = {};
	if (F::begin(ctx, S, K, s)) F::end(ctx, S, K, s);
}

// ---- the walker ---------------------------------------------------------------------------------------------------
// exact pair geometry of staged record m for particle s, then the bodies
template <bool POLY, class F>
__device__ __forceinline__ void tv_pair(typename F::Ctx& ctx, const TvStage& T, const KParams& K, const Self& s, int m) {
	Pair q; q.j = m;
	const double2 xy = T.pxy[m], zw = T.pzw[m];
	const double xj = xy.x, yj = xy.y, zj = zw.x;
	q.rx = xj - s.pi.x; q.ry = yj - s.pi.y; q.rz = zj - s.pi.z;
	q.d2 = sq3_exact(q.rx, q.ry, q.rz);
	if (!(q.d2 < K.reMax2)) return;
	q.tj = __double2loint(zw.y); q.ptj = __double2hiint(zw.y);   // the type word of the position record (pk_make)
	if (POLY && s.hasMirror) { q.mx = xj - s.mix; q.my = yj - s.miy; q.mz = zj - s.miz; q.dm2 = sq3_exact(q.mx, q.my, q.mz); }
	else { q.mx = q.my = q.mz = 0.0; q.dm2 = __longlong_as_double(0x7ff0000000000000ll); }   // no mirror in reach: d2 < dm2 holds, M(re) fails (pmps_internal.h, mirrorFar2)
	q.rinv = fast_rsqrt(q.d2);
	F::template pair<POLY>(ctx, T.nb, K, s, q);
}
// two staged records at once: the two geometry chains (shared-memory loads, exact distance, reciprocal square root) are
// independent and interleave
template <bool POLY, class F>
__device__ __forceinline__ void tv_pair2(typename F::Ctx& ctx, const TvStage& T, const KParams& K, const Self& s, int m0, int m1) {
	Pair a, b; a.j = m0; b.j = m1;
	const double2 xya = T.pxy[m0], zwa = T.pzw[m0], xyb = T.pxy[m1], zwb = T.pzw[m1];
	const double xa = xya.x, ya = xya.y, za = zwa.x, xb = xyb.x, yb = xyb.y, zb = zwb.x;
	a.rx = xa - s.pi.x; a.ry = ya - s.pi.y; a.rz = za - s.pi.z; b.rx = xb - s.pi.x; b.ry = yb - s.pi.y; b.rz = zb - s.pi.z;
	a.d2 = sq3_exact(a.rx, a.ry, a.rz); b.d2 = sq3_exact(b.rx, b.ry, b.rz);
	const bool ha = a.d2 < K.reMax2, hb = b.d2 < K.reMax2;
	if (!(ha || hb)) return;
	a.tj = __double2loint(zwa.y); a.ptj = __double2hiint(zwa.y); b.tj = __double2loint(zwb.y); b.ptj = __double2hiint(zwb.y);
	if (POLY && s.hasMirror) {
		a.mx = xa - s.mix; a.my = ya - s.miy; a.mz = za - s.miz; a.dm2 = sq3_exact(a.mx, a.my, a.mz);
		b.mx = xb - s.mix; b.my = yb - s.miy; b.mz = zb - s.miz; b.dm2 = sq3_exact(b.mx, b.my, b.mz);
	} else {
		a.mx = a.my = a.mz = b.mx = b.my = b.mz = 0.0; a.dm2 = b.dm2 = __longlong_as_double(0x7ff0000000000000ll);
	}
	a.rinv = fast_rsqrt(a.d2); b.rinv = fast_rsqrt(b.d2);
	if (ha) F::template pair<POLY>(ctx, T.nb, K, s, a);
	if (hb) F::template pair<POLY>(ctx, T.nb, K, s, b);
}
// The first chunk (64 entries, 128 bytes) of the hit lists of one round on its way into shared memory.  Every WARP fetches the rows
// of its own lane groups and is the only one to read them: a round needs no block barrier, and because a lane group takes its row
// into registers at the start of the round, the next round's rows can land in the same buffer while the round computes.
// A row is 128 contiguous, 128-byte aligned bytes of global memory: ONE bulk copy of the TMA unit per row (cp.async.bulk, SASS
// UBLKCP), completing on the warp's mbarrier — 8 instructions per warp and round instead of 64 16-byte cp.async.  (The staged
// halo cannot travel that way: a bucket's stretch of an 8-byte plane starts at an arbitrary slot, i.e. is 8-byte aligned, and the
// 32-byte position records would land as an array of structures whose 128-bit reads conflict two ways on the banks; they go as
// 16-byte / 8-byte cp.async, tv_stage.)
#ifndef TV_ROWS_TMA
#define TV_ROWS_TMA 0   // measured on the B200 (10 M particles, A/B on one box): every walker pass 4-6 % SLOWER with the bulk copies + mbarrier wait than
                      // with the 16-byte cp.async it replaces (F3 5.71 -> 6.07 ms, F1 4.59 -> 4.79, F5 2.82 -> 2.96): kept as an option, off
#endif
__device__ __forceinline__ void tv_bar_init(uint64_t* bar) {
	asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"((unsigned)__cvta_generic_to_shared(bar)) : "memory");
}
__device__ __forceinline__ void tv_bar_wait(uint64_t* bar, unsigned phase) {
	asm volatile("{\n\t.reg .pred P1;\n\tTV_WAIT:\n\tmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t@P1 bra TV_DONE;\n\tbra TV_WAIT;\n\tTV_DONE:\n\t}"
	             :: "r"((unsigned)__cvta_generic_to_shared(bar)), "r"(phase) : "memory");
}
// returns true when the warp has rows on their way (it must then wait for them at the start of the round that uses them)
__device__ __forceinline__ bool tv_fetch_rows(const DevView& S, unsigned char* rowsBuf, uint64_t* bar, int own0, int p0, int P) {
	const int lane = threadIdx.x & 31, g0 = (threadIdx.x >> 5) * (32 / TV_G);
#if TV_ROWS_TMA
	const int nrows = min(32 / TV_G, P - (p0 + g0));   // rows of this warp in the round (uniform over the warp)
	if (nrows <= 0) return false;
	const unsigned barS = (unsigned)__cvta_generic_to_shared(bar);
	if (lane == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(barS), "r"((unsigned)(nrows * TV_CHUNK * 2)) : "memory");
	__syncwarp();
	if (lane < nrows) {
		const unsigned dst = (unsigned)__cvta_generic_to_shared(rowsBuf + (size_t)(g0 + lane) * TV_LROW);
		const void* src = S.hlist + (size_t)(own0 + p0 + g0 + lane) * (size_t)S.tlCap;
		asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" :: "r"(dst), "l"(src), "r"((unsigned)(TV_CHUNK * 2)), "r"(barS) : "memory");
	}
	return true;
#else
	#pragma unroll
	for (int q = lane; q < (32 / TV_G) * (TV_CHUNK * 2 / 16); q += 32) {
		const int r = q >> 3, c = q & 7, p = p0 + g0 + r;
		if (p < P) tv_cp16(rowsBuf + (size_t)(g0 + r) * TV_LROW + c * 16, (const unsigned char*)(S.hlist + (size_t)(own0 + p) * (size_t)S.tlCap) + c * 16);
	}
	asm volatile("cp.async.commit_group;" ::: "memory");
	return true;
#endif
}

// Every lane group takes its particle's HIT list (k_hl_build: the staged records within reach at the positions of the current
// epoch) — no scan, no queue, no votes.  What a round needs from global memory besides the staged halo — the list rows, the list
// lengths, the has-a-mirror flags — is fetched into shared memory ahead of time (lists: cp.async, one round ahead, per warp; the
// per-particle scalars with the staging loads), because ncu showed the round-start loads of these as 25 % of all stall samples.
// Block barriers per work item: table loaded, halo staged, item done (twice) — none inside the rounds.
template <bool POLY, unsigned FIELDS, bool WALLONLY, bool SPARSE, class F>
__global__ void __launch_bounds__(TV_T, TV_BLOCKS(FIELDS)) k_tv(const __grid_constant__ DevView S, const __grid_constant__ KParams K, int cursorIdx) {
	extern __shared__ __align__(16) unsigned char smemRaw[];
	__shared__ int sTab[TV_TABMAX], sItem;
	__shared__ __align__(8) uint64_t sRowBar[TV_T / 32];   // one mbarrier per warp: its list rows have landed
	const TvStage T = tv_carve<FIELDS>(smemRaw);
	const int tid = threadIdx.x, grp = tid / TV_G, g = tid % TV_G;
	uint64_t* const rowBar = sRowBar + (tid >> 5);
	unsigned rowPhase = 0u;   // parity of the warp's next completed row fetch (uniform over the warp)
#if TV_ROWS_TMA
	if ((tid & 31) == 0) tv_bar_init(rowBar);
	asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
	__syncthreads();
#endif
	unsigned char* const fx = smemRaw + (size_t)TV_MAXM * tv_rec_bytes(FIELDS);
	unsigned char* const rowsBuf = fx;                                       // TV_NG list rows of TV_LROW bytes
	uint16_t* const ownMap = (uint16_t*)(fx + TV_NG * TV_LROW);          // own particle -> staged index
	uint16_t* const sCnt = ownMap + TV_MAXOWN;                               // own particle -> entries of its hit list
	uint8_t* const sHas = (uint8_t*)(sCnt + TV_MAXOWN);                      // own particle -> a mirror particle is in reach
	const unsigned gshift = (tid & 31) & ~(TV_G - 1);
	const unsigned gbits = (1u << TV_G) - 1u, gmask = gbits << gshift;
	const bool three = K.p.dim == 3;
	const int nItems = *S.nTiles;
	const int* hOff = sTab; const int* hBase = sTab + K.HVt + 1;
	// wall-only / close-pair packs visit only the work items k_item_flags listed for them (the tails of the other tiles: k_tails)
	constexpr bool LISTED = WALLONLY || F::kCloseOnly || SPARSE;
	constexpr int KIND = WALLONLY ? 0 : SPARSE ? 2 : 1;
	const int nVisit = LISTED ? S.actCount[KIND] : nItems;
	const int32_t* visit = S.actList + KIND * S.actCap;
	for (int it = tv_first_item(S, cursorIdx, &sItem), nxt = 0; it < nVisit; it = tv_advance(nxt, &sItem)) {
		if (tid == 0) nxt = atomicAdd(&S.tileCursor[cursorIdx], 1);
		const int tt = LISTED ? visit[it] : it;
		const int item = S.tileList[tt];
		tv_load_table(S, K, tt, sTab);
		const TvBox b = tv_box_tab(K, item, sTab);
		const int M = hOff[b.HV], own0 = sTab[2 * K.HVt + 1], P = sTab[2 * K.HVt + 2];
		if (M + 1 > TV_MAXM || P > TV_MAXOWN) continue;   // reported by k_hl_build (FLAG_TILE_OVERFLOW)
		if (SPARSE) {   // does any particle of the tile want pairs?  (begin() of these bodies is idempotent)
			bool want = false;
			for (int p = tid; p < P; p += TV_T) {
				const int i = own0 + p;
				Self s; s.nb = nullptr; s.ms = 0; s.i = i; s.lead = true; s.nearWallFluid = false; s.hasMirror = false; s.mix = s.miy = s.miz = 0.0;
				s.pi = ld256(S.pos + i); s.type = pk_type(s.pi); s.ptype = pk_ptype(s.pi);
				if (s.type == K.p.ghost) continue;
				typename F::Ctx ctx = {};
				if (F::begin(ctx, S, K, s) && F::need_pairs(ctx)) want = true;
			}
			if (!__syncthreads_or(want)) {
				for (int p = tid; p < P; p += TV_T) {   // nobody: the tails still run, one lane per particle
					const int i = own0 + p;
					Self s; s.nb = nullptr; s.ms = 0; s.i = i; s.lead = true; s.nearWallFluid = false; s.hasMirror = false; s.mix = s.miy = s.miz = 0.0;
					s.pi = ld256(S.pos + i); s.type = pk_type(s.pi); s.ptype = pk_ptype(s.pi);
					if (F::kCopyPos) S.pos2[i] = s.pi;
					if (F::kCopyVel) S.vel2[i] = S.vel[i];
					if (s.type == K.p.ghost) continue;
					typename F::Ctx ctx = {};
					if (F::begin(ctx, S, K, s)) F::end(ctx, S, K, s);
				}
				continue;   // uniform over the block
			}
		}
		const double ox = K.p.domainMinX + (double)(b.ox - 1) * K.p.bucketSide, oy = K.p.domainMinY + (double)(b.oy - 1) * K.p.bucketSide,
		             oz = three ? K.p.domainMinZ + (double)(b.oz - 1) * K.p.bucketSide : 0.0;
		bool rowsComing = tv_fetch_rows(S, rowsBuf, rowBar, own0, 0, P);   // first round's lists on their way while the staging loads land
		for (int p = tid; p < P; p += TV_T) {     // per-particle scalars of the whole item, independent loads issued together with the staging
			sCnt[p] = S.hcnt[own0 + p];
			if (POLY) sHas[p] = S.distParticleWall2[own0 + p] < K.mirrorFar2 ? 1 : 0;
		}
		tv_stage<FIELDS>(S, T, hOff, hBase, b.HV, own0, P, ownMap);
		if (K.p.domainTypeBC == 1) tv_stage_periodic(K, b, T, hOff);
		if (tid == 0) { T.pxy[M] = make_double2(1e30, 1e30); T.pzw[M] = make_double2(1e30, 0.0); }   // the record list padding points at
		__syncthreads();   // the staged halo, ownMap and the per-particle scalars are visible to every thread
		const uint4* const Ls = (const uint4*)(rowsBuf + (size_t)grp * TV_LROW) + g * (TV_PER / 8);   // this lane's part of its group's row
		// ---- rounds of TV_NG particles; control flow stays uniform over a warp (full-mask shuffles) ----
		for (int p0 = 0; p0 < P; p0 += TV_NG) {
#if TV_ROWS_TMA
			if (rowsComing) { tv_bar_wait(rowBar, rowPhase); rowPhase ^= 1u; }   // this warp's rows of the round have landed
#else
			asm volatile("cp.async.wait_group 0;" ::: "memory");
			__syncwarp();
#endif
			unsigned w[TV_PER / 2];   // chunk 0 of the list: this lane's entries, into registers
			#pragma unroll
			for (int v = 0; v < TV_PER / 8; v++) { const uint4 e = Ls[v]; w[4 * v] = e.x; w[4 * v + 1] = e.y; w[4 * v + 2] = e.z; w[4 * v + 3] = e.w; }
			{ unsigned t_ = w[TV_PER / 2 - 1] ^ w[0]; t_ = __shfl_xor_sync(0xffffffffu, t_, 1); asm volatile("" :: "r"(t_)); }   // every lane HAS its words before the buffer is reused
			rowsComing = p0 + TV_NG < P ? tv_fetch_rows(S, rowsBuf, rowBar, own0, p0 + TV_NG, P) : false;   // next round's rows land while this round computes
			const int p = p0 + grp;
			bool act = p < P;
			const int i = own0 + (act ? p : 0);
			Self s; s.nb = &T.nb; s.ms = M; s.i = i; s.lead = g == 0; s.nearWallFluid = false; s.hasMirror = false; s.mix = s.miy = s.miz = 0.0;
			s.type = K.p.ghost; s.ptype = 0;
			int cntE = 0;
			if (act) {
				cntE = sCnt[p];
				const int ms = ownMap[p]; s.ms = ms;
				const double2 xy = T.pxy[ms], zw = T.pzw[ms];
				s.pi.x = xy.x; s.pi.y = xy.y; s.pi.z = zw.x; s.pi.w = zw.y;
				s.type = pk_type(s.pi); s.ptype = pk_ptype(s.pi);
				if (F::kCopyPos && s.lead) S.pos2[i] = s.pi;
				if (F::kCopyVel && s.lead) S.vel2[i] = S.vel[i];
				act = s.type != K.p.ghost;
			}
			typename F::Ctx ctx = {};   // idle lanes still take part in the (full-mask) reductions
			bool pairs = false;
			if (act) {
				if (POLY) {
					s.hasMirror = sHas[p] != 0;
					if (s.hasMirror) { s.mix = S.mirrorParticlePos.x[i]; s.miy = S.mirrorParticlePos.y[i]; s.miz = S.mirrorParticlePos.z[i]; }
					else { s.mix = s.miy = s.miz = K.big; }
				}
				if (POLY && F::kWall) { s.nearWallFluid = s.type == K.p.fluid && S.particleNearWall[i]; if (s.nearWallFluid) self_wall_frame(s); }
				act = F::begin(ctx, S, K, s);
				pairs = act && F::need_pairs(ctx);
			}
			if (!pairs) cntE = 0;
			const uint4* Lg = (const uint4*)(S.hlist + (size_t)i * (size_t)S.tlCap) + g * (TV_PER / 8);   // further chunks (rare): global
			for (int ch = 0; ch * TV_CHUNK < cntE; ch++) {
				if (ch > 0) {
					#pragma unroll
					for (int v = 0; v < TV_PER / 8; v++) { const uint4 e = __ldg(Lg + ch * (TV_CHUNK / 8) + v); w[4 * v] = e.x; w[4 * v + 1] = e.y; w[4 * v + 2] = e.z; w[4 * v + 3] = e.w; }
				}
				const int rem = cntE - ch * TV_CHUNK - g;   // entries u * TV_G + g of this chunk exist for u * TV_G < rem
				if (TV_ILP2 || (TV_ILP2_FAT && TV_BLOCKS(FIELDS) == 2)) {
					#pragma unroll
					for (int u = 0; u < TV_PER; u += 2) {   // two entries per trip (one 32-bit word of the list); a missing one points at the dummy record
						if (u * TV_G < rem) tv_pair2<POLY, F>(ctx, T, K, s, (int)(w[u >> 1] & 0xffffu), (u + 1) * TV_G < rem ? (int)(w[u >> 1] >> 16) : M);
					}
				} else {
					#pragma unroll
					for (int u = 0; u < TV_PER; u++) {
						if (u * TV_G < rem) tv_pair<POLY, F>(ctx, T, K, s, (int)((w[u >> 1] >> ((u & 1) * 16)) & 0xffffu));
					}
				}
			}
			F::reduce(ctx, gmask);   // all lanes, uniform
			if (act && s.lead) F::end(ctx, S, K, s);
		}
	}
}
