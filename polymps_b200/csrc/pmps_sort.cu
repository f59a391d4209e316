// pmps_sort.cu — G0: bucket ordering + Verlet candidate list (replaces MpsBucket::updateParticlesID,
// reference src/MpsBucket.cpp:299-353).
//
// The reference threads particles onto per-bucket linked lists serially.  Here the particles themselves are put in
// bucket order by a counting sort (one atomic per particle into the bucket histogram, a device-wide exclusive scan
// = the cell-start table, a scatter, and an in-bucket rank fix that makes the order canonical: ascending host index
// inside a bucket, i.e. exactly the reference's list order).  All per-particle state is then gathered into the new
// order so every later kernel reads its neighbours from nearby addresses.  Finally ONE pass over the 27 (9 in 2D)
// surrounding buckets builds a candidate list with radius bucketSide = reL(1+CFL) in float; the ~10 neighbour passes
// of the step walk that list (≈67 candidates) instead of re-walking 27 buckets (≈430 candidates) each.
#define PMPS_TV_LISTBUILD   // this translation unit launches k_tv_items / k_hl_build
#include "pmps_tv.cuh"
#include <cstdio>

// ---- 1. bucket id + histogram --------------------------------------------------------------------------------
__global__ void k_cell_hist(const __grid_constant__ DevView S, const __grid_constant__ KParams K, int32_t* newCell, int32_t* arbRank, int32_t* cellCount) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= S.n0) return;
	const double4 p = S.pos[i];
	int cell = S.nC;   // ghosts are parked behind the last cell (the reference skips them, src/MpsBucket.cpp:332)
	if (pk_type(p) != K.p.ghost) {
		const PmpsParams& P = K.p;
		int ix, iy, iz = 0;
		if (P.dim == 2) {   // 2D insertion adds epsilonZero, 3D insertion does not (src/MpsBucket.cpp:312-313 vs :333-335)
			ix = bucket_coord(p.x, P.domainMinX, P.invBucketSide, P.epsilonZero);
			iy = bucket_coord(p.y, P.domainMinY, P.invBucketSide, P.epsilonZero);
		} else {
			ix = (int)__dmul_rn(__dsub_rn(p.x, P.domainMinX), P.invBucketSide);
			iy = (int)__dmul_rn(__dsub_rn(p.y, P.domainMinY), P.invBucketSide);
			iz = (int)__dmul_rn(__dsub_rn(p.z, P.domainMinZ), P.invBucketSide);
		}
		if (ix < 0 || ix >= P.numBucketsX || iy < 0 || iy >= P.numBucketsY || iz < 0 || iz >= P.numBucketsZ) {
			atomicOr(&S.flags[FLAG_GUARD], 2);   // a live particle outside the bucket grid: the reference would corrupt memory here
			ix = min(max(ix, 0), P.numBucketsX - 1); iy = min(max(iy, 0), P.numBucketsY - 1); iz = min(max(iz, 0), P.numBucketsZ - 1);
		}
		if (P.domainTypeBC == 1) {
			// periodic domain: a particle handed over beyond a periodic face (an initial state; the next out-of-domain check wraps it) lies in a
			// border layer, which holds no particles of its own here — it aliases the opposite inner layer (tv_halo_cell).  It is sorted into the
			// inner layer next to it: its search then covers the border layer's alias and that layer, i.e. everything within reach.
			int& ca = P.periodicAxis == 0 ? ix : P.periodicAxis == 1 ? iy : iz;
			const int na = P.periodicAxis == 0 ? P.numBucketsX : P.periodicAxis == 1 ? P.numBucketsY : P.numBucketsZ;
			ca = min(max(ca, 1), na - 2);
		}
		cell = cell_index(K, ix, iy, iz);   // tile-major: the particles of a tile end up in consecutive slots
	}
	newCell[i] = cell;
	// the ghost cell can hold millions of rows (the unused tail of a slab rank's table): one atomic per warp instead of one per row
	const bool ghostRow = cell == S.nC;
	const unsigned act = __activemask();
	const unsigned gm = __ballot_sync(act, ghostRow);
	if (ghostRow) {
		const int lane = threadIdx.x & 31, leader = __ffs(gm) - 1;
		int base = 0;
		if (lane == leader) base = atomicAdd(&cellCount[cell], __popc(gm));
		base = __shfl_sync(gm, base, leader);
		arbRank[i] = base + __popc(gm & ((1u << lane) - 1u));
	} else arbRank[i] = atomicAdd(&cellCount[cell], 1);
}

// ---- 2. exclusive scan of the histogram (3 kernels: per-block scan, scan of block sums, add) ------------------
#define SCAN_T 512
#define SCAN_ITEMS 8
__global__ void __launch_bounds__(SCAN_T) k_scan_block(const int32_t* in, int32_t* out, int32_t* blockSum, int n, int32_t* maxOut) {
	__shared__ int32_t warpSum[SCAN_T / 32];
	const int base = (blockIdx.x * SCAN_T + threadIdx.x) * SCAN_ITEMS;
	int v[SCAN_ITEMS]; int s = 0, mx = 0;
	#pragma unroll
	for (int k = 0; k < SCAN_ITEMS; k++) { const int idx = base + k; v[k] = idx < n ? in[idx] : 0; if (idx < n - 2) mx = max(mx, v[k]); s += v[k]; }
	int incl = s;
	#pragma unroll
	for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if ((threadIdx.x & 31) >= o) incl += t; }
	if ((threadIdx.x & 31) == 31) warpSum[threadIdx.x >> 5] = incl;
	__syncthreads();
	if (threadIdx.x < 32) {
		int w = threadIdx.x < SCAN_T / 32 ? warpSum[threadIdx.x] : 0; int wi = w;
		#pragma unroll
		for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, wi, o); if (threadIdx.x >= o) wi += t; }
		if (threadIdx.x < SCAN_T / 32) warpSum[threadIdx.x] = wi - w;
		if (threadIdx.x == SCAN_T / 32 - 1) blockSum[blockIdx.x] = wi;
	}
	__syncthreads();
	int run = incl - s + warpSum[threadIdx.x >> 5];
	#pragma unroll
	for (int k = 0; k < SCAN_ITEMS; k++) { const int idx = base + k; if (idx < n) out[idx] = run; run += v[k]; }
	mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 16)); mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 8));
	mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 4)); mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 2)); mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
	if ((threadIdx.x & 31) == 0 && mx > 0) atomicMax(maxOut, mx);
}
__global__ void __launch_bounds__(1024) k_scan_sums(int32_t* blockSum, int nBlocks) {   // one block, serial over chunks of 1024
	__shared__ int32_t sh[1024]; __shared__ int32_t carry;
	if (threadIdx.x == 0) carry = 0;
	__syncthreads();
	for (int base = 0; base < nBlocks; base += 1024) {
		const int idx = base + threadIdx.x; const int v = idx < nBlocks ? blockSum[idx] : 0;
		sh[threadIdx.x] = v; __syncthreads();
		for (int o = 1; o < 1024; o <<= 1) { int t = threadIdx.x >= o ? sh[threadIdx.x - o] : 0; __syncthreads(); sh[threadIdx.x] += t; __syncthreads(); }
		if (idx < nBlocks) blockSum[idx] = sh[threadIdx.x] - v + carry;
		__syncthreads();
		if (threadIdx.x == 1023) carry += sh[1023];
		__syncthreads();
	}
}
__global__ void __launch_bounds__(SCAN_T) k_scan_add(int32_t* out, const int32_t* blockSum, int n) {
	const int base = (blockIdx.x * SCAN_T + threadIdx.x) * SCAN_ITEMS; const int add = blockSum[blockIdx.x];
	#pragma unroll
	for (int k = 0; k < SCAN_ITEMS; k++) { const int idx = base + k; if (idx < n) out[idx] += add; }
}

// ---- 3. scatter into bucket ranges (arbitrary order inside a bucket), 4. canonical rank inside the bucket -----
__global__ void k_scatter(const __grid_constant__ DevView S, const int32_t* newCell, const int32_t* arbRank, const int32_t* cellStart, int2* tmp) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= S.n0) return;
	tmp[cellStart[newCell[i]] + arbRank[i]] = make_int2(S.hostIdx[i], i);
}
__global__ void k_rank_fix(const __grid_constant__ DevView S, const int32_t* newCell, const int32_t* cellStart, const int2* tmp, int32_t* perm) {
	const int t = blockIdx.x * blockDim.x + threadIdx.x;
	if (t >= S.n0) return;
	const int2 me = tmp[t];
	const int c = newCell[me.y]; const int s0 = cellStart[c], s1 = cellStart[c + 1];
	if (c == S.nC) { perm[t] = me.y; return; }   // ghosts: no order to keep (and the count-the-smaller ranking is quadratic in the cell size)
	int rank = 0;
	for (int u = s0; u < s1; u++) rank += tmp[u].x < me.x;
	perm[s0 + rank] = me.y;
}

// ---- 5. gather every plane into the new order --------------------------------------------------------------------
#define PERM_MAX 16
struct PermJob8 { const double* src[PERM_MAX]; double* dst[PERM_MAX]; int n; };
__global__ void k_permute8(const int32_t* __restrict__ perm, const __grid_constant__ PermJob8 J, int n0) {
	const int s = blockIdx.x * blockDim.x + threadIdx.x;
	if (s >= n0) return;
	const int o = perm[s];
	#pragma unroll 4
	for (int k = 0; k < J.n; k++) J.dst[k][s] = J.src[k][o];
}
struct PermJob4 { const int32_t* src[PERM_MAX]; int32_t* dst[PERM_MAX]; int n; };
__global__ void k_permute4(const int32_t* __restrict__ perm, const __grid_constant__ PermJob4 J, int n0) {
	const int s = blockIdx.x * blockDim.x + threadIdx.x;
	if (s >= n0) return;
	const int o = perm[s];
	for (int k = 0; k < J.n; k++) J.dst[k][s] = J.src[k][o];
}
// packs + bookkeeping: pos, vel, nearWall flag, cellOf, slotOfHost, float build position
__global__ void k_permute_packs(const int32_t* __restrict__ perm, const __grid_constant__ DevView S, const __grid_constant__ DevView A,
                                const __grid_constant__ KParams K, const int32_t* newCell) {
	const int s = blockIdx.x * blockDim.x + threadIdx.x;
	if (s >= S.n0) return;
	const int o = perm[s];
	const double4 p = S.pos[o];
	A.pos[s] = p; A.vel[s] = S.vel[o];
	A.particleNearWall[s] = S.particleNearWall[o]; A.stepCell[s] = S.stepCell[o];
	if (K.p.domainTypeBC == 1) A.wrapHid[s] = S.wrapHid[o];
	const int h = S.hostIdx[o];
	A.hostIdx[s] = h; S.slotOfHost[h] = s;
	S.cellOf[s] = newCell[o];   // cellOf / bpos are not double-buffered: written at the new slot, read via perm only above
	S.bpos[s] = make_float4((float)(p.x - K.p.domainMinX), (float)(p.y - K.p.domainMinY), (float)(p.z - K.p.domainMinZ), 0.f);
}

// ---- 6. Verlet candidate list -----------------------------------------------------------------------------------
// Global-slot form of the candidate list, built ON DEMAND for the consumers that are not tile passes (PPE assembly,
// debug hooks).  One WARP per slot: the lanes test 32 slots of a bucket at a time and the survivors are compacted in
// order with a ballot; buckets in the reference's order (jz, jy, jx).  Row-major: candidate k of slot i at list[i*cap + k].
__global__ void __launch_bounds__(128) k_build_list(const __grid_constant__ DevView S, const __grid_constant__ KParams K) {
	const int i = (int)(((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
	const int lane = threadIdx.x & 31;
	if (i >= S.n0) return;
	const int c = S.cellOf[i];
	if (c >= S.nC) { if (lane == 0) S.listCnt[i] = 0; return; }   // ghost
	const int nX = K.p.numBucketsX, nY = K.p.numBucketsY, nZ = K.p.numBucketsZ;
	int cx, cy, cz; cell_coords(K, c, cx, cy, cz);
	const float4 bi = S.bpos[i];
	uint32_t* L = S.list + (size_t)i * (size_t)S.listCap;
	int cnt = 0; const int cap = S.listCap;
	const int z0 = K.p.dim == 3 ? max(cz - 1, 0) : 0, z1 = K.p.dim == 3 ? min(cz + 1, nZ - 1) : 0;
	const int y0 = max(cy - 1, 0), y1 = min(cy + 1, nY - 1), x0 = max(cx - 1, 0), x1 = min(cx + 1, nX - 1);
	const unsigned below = (1u << lane) - 1u;
	for (int jz = z0; jz <= z1; jz++) for (int jy = y0; jy <= y1; jy++) for (int jx = x0; jx <= x1; jx++) {
		int ax = jx, ay = jy, az = jz; float sx = 0.f, sy = 0.f, sz = 0.f;
		if (K.p.domainTypeBC == 1) {   // periodic domain: a border layer holds the opposite inner layer, moved by -/+ periodicLength
			const float sh = (float)per_shift_of_coord(K, K.p.periodicAxis == 0 ? jx : K.p.periodicAxis == 1 ? jy : jz);
			per_alias(K, ax, ay, az);
			if (K.p.periodicAxis == 0) sx = sh; else if (K.p.periodicAxis == 1) sy = sh; else sz = sh;
		}
		const int cj = cell_index(K, ax, ay, az);
		const int s0 = S.cellStart[cj], s1 = S.cellStart[cj + 1];
		for (int base = s0; base < s1; base += 32) {
			const int j = base + lane; bool hit = false;
			if (j < s1) {
				float4 bj = S.bpos[j]; bj.x += sx; bj.y += sy; bj.z += sz;
				const float dx = bj.x - bi.x, dy = bj.y - bi.y, dz = bj.z - bi.z;
				hit = dx * dx + dy * dy + dz * dz < K.rl2 && j != i;
			}
			const unsigned b = __ballot_sync(0xffffffffu, hit);
			if (hit) { const int at = cnt + __popc(b & below); if (at < cap) L[at] = (uint32_t)j; }
			cnt += __popc(b);
		}
	}
	if (lane == 0) {
		if (cnt > cap) { atomicMax(&S.flags[FLAG_OVERFLOW], cnt); cnt = cap; }
		S.listCnt[i] = (uint16_t)cnt;
	}
}

__global__ void k_reset_step_scalars(const __grid_constant__ DevView S) {
	S.maxDisp2[0] = 0.f; S.maxDisp2[2] = 0.f; S.flags[FLAG_MAXCELL] = 0;
}

// -------------------------------------------------------------------------------------------------------------------
template <class T> static inline void swp(T& a, T& b) { T t = a; a = b; b = t; }

int pmps_sort_and_list(pmps_handle* h) {
	DevView& S = h->V; DevView& A = h->A; const KParams& K = h->K;
	const int n0 = S.n0, nB = S.nC, T = 256;
	CUDA_TRY(h, cudaMemsetAsync(h->cellCount, 0, sizeof(int32_t) * (size_t)(nB + 2), h->stream));   // nB here = cells of the tile-major index
	LAUNCH(h, "sort.reset", k_reset_step_scalars, 1, 1, S);
	LAUNCH(h, "sort.cell_hist", k_cell_hist, cdiv(n0, T), T, S, K, h->newCell, h->arbRank, h->cellCount);
	const int nScan = nB + 2; const int per = SCAN_T * SCAN_ITEMS; const int nblk = cdiv(nScan, per);
	LAUNCH(h, "sort.scan_block", k_scan_block, nblk, SCAN_T, h->cellCount, S.cellStart, h->scanBlock, nScan, &S.flags[FLAG_MAXCELL]);
	LAUNCH(h, "sort.scan_sums", k_scan_sums, 1, 1024, h->scanBlock, nblk);
	LAUNCH(h, "sort.scan_add", k_scan_add, nblk, SCAN_T, S.cellStart, h->scanBlock, nScan);
	int2* tmp = (int2*)h->tmpHost;
	LAUNCH(h, "sort.scatter", k_scatter, cdiv(n0, T), T, S, h->newCell, h->arbRank, S.cellStart, tmp);
	LAUNCH(h, "sort.rank_fix", k_rank_fix, cdiv(n0, T), T, S, h->newCell, S.cellStart, tmp, h->perm);
	// gather: packs first (needs the old hostIdx), then the planes in groups
	LAUNCH(h, "sort.permute_packs", k_permute_packs, cdiv(n0, T), T, h->perm, S, A, K, h->newCell);
	swp(S.pos, A.pos); swp(S.vel, A.vel); swp(S.particleNearWall, A.particleNearWall); swp(S.hostIdx, A.hostIdx); swp(S.stepCell, A.stepCell); swp(S.wrapHid, A.wrapHid);
	S.pos2 = A.pos; S.vel2 = A.vel;
	{
		std::vector<std::pair<double**, double**>> planes;
#define X(n) planes.push_back({ &S.n, &A.n });
		PMPS_F64_FIELDS(X)
		if (K.p.fluidType == 1) { PMPS_RHEO_FIELDS(X) }
#undef X
#define X(n) planes.push_back({ &S.n.x, &A.n.x }); planes.push_back({ &S.n.y, &A.n.y }); planes.push_back({ &S.n.z, &A.n.z });
		PMPS_V3_FIELDS(X)
#undef X
		for (size_t b = 0; b < planes.size(); b += PERM_MAX) {
			PermJob8 J; J.n = (int)std::min<size_t>(PERM_MAX, planes.size() - b);
			for (int k = 0; k < J.n; k++) { J.src[k] = *planes[b + k].first; J.dst[k] = *planes[b + k].second; }
			LAUNCH(h, "sort.permute8", k_permute8, cdiv(n0, T), T, h->perm, J, n0);
			for (int k = 0; k < J.n; k++) swp(*planes[b + k].first, *planes[b + k].second);
		}
		PermJob4 J4; J4.n = 0;
#define X(f) J4.src[J4.n] = S.f; J4.dst[J4.n] = A.f; J4.n++;
		PMPS_I32_FIELDS(X)
#undef X
		LAUNCH(h, "sort.permute4", k_permute4, cdiv(n0, T), T, h->perm, J4, n0);
#define X(n) swp(S.n, A.n);
		PMPS_I32_FIELDS(X)
#undef X
	}
	CUDA_TRY(h, cudaGetLastError());
	h->haveList = true;      // bucket order valid
	h->listBuilt = false;    // the Verlet list is built on demand (stage API, PPE assembly, debug hooks)
	S.listValid = 0;
	return PMPS_OK;
}

int pmps_ensure_list(pmps_handle* h) {
	if (h->listBuilt) return PMPS_OK;
	DevView& S = h->V;
	if (!S.list) {   // allocated on first use: only the PPE assembly and the debug hooks read global slot ids
		const size_t nWarps = ((size_t)S.n0 + 31) / 32, cnt = nWarps * 32 * (size_t)S.listCap;
		void* a = nullptr; void* b = nullptr;
		if (cudaMalloc(&a, cnt * sizeof(uint32_t)) != cudaSuccess || cudaMalloc(&b, nWarps * 32 * sizeof(uint16_t)) != cudaSuccess)
			return pmps_fail(h, PMPS_E_CUDA, "candidate list allocation (%zu bytes) failed: %s", cnt * 4, cudaGetErrorString(cudaGetLastError()));
		h->allocs.push_back(a); h->allocs.push_back(b);
		S.list = (uint32_t*)a; S.listCnt = (uint16_t*)b;
	}
	LAUNCH(h, "list.build", k_build_list, cdiv((long long)S.n0 * 32, 128), 128, h->V, h->K);
	CUDA_TRY(h, cudaGetLastError());
	h->listBuilt = true; S.listValid = 1;
	return PMPS_OK;
}

// work items + hit list of the new bucket order (every neighbour pass of the step follows it)
static int ilog2_ceil(int n) { int b = 0; while ((1 << b) < n) b++; return b; }
static int launch_hl_build(pmps_handle* h, const char* name, int force) {
	DevView& S = h->V;
	if (h->walkSeq >= PMPS_CURSORS) return pmps_fail(h, PMPS_E_STATE, "more than %d tile passes since the last pmps_update_buckets", PMPS_CURSORS - 1);
	const size_t smem = 16 * (size_t)(TV_MAXM + 4) + (size_t)HL_T * HL_ROW * 2 + 4 * (size_t)(TV_MAXM + 4);   // float positions (+ the dummy record) + the first chunk of one list row per lane + staged index -> slot
	if (h->K.p.domainTypeBC == 1) {
		CUDA_TRY(h, cudaFuncSetAttribute(k_hl_build<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
		ProfScope ps_(h, name); k_hl_build<true><<<h->smCount * 5, HL_T, smem, h->stream>>>(S, h->K, h->walkSeq++, force); h->launches++;
	} else {
		CUDA_TRY(h, cudaFuncSetAttribute(k_hl_build<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
		ProfScope ps_(h, name); k_hl_build<false><<<h->smCount * 5, HL_T, smem, h->stream>>>(S, h->K, h->walkSeq++, force); h->launches++;
	}
	CUDA_TRY(h, cudaGetLastError());
	return PMPS_OK;
}
int pmps_tv_prepare(pmps_handle* h) {
	DevView& S = h->V; const KParams& K = h->K;
	CUDA_TRY(h, cudaMemsetAsync(S.nTiles, 0, sizeof(int32_t), h->stream));
	CUDA_TRY(h, cudaMemsetAsync(S.tileCursor, 0, sizeof(int32_t) * PMPS_CURSORS, h->stream));
	CUDA_TRY(h, cudaMemsetAsync(S.tileWall, 0, (size_t)h->nTilesAll, h->stream));
	h->walkSeq = 0; h->tileWallValid = false;
	const int bx = ilog2_ceil(K.nTX), by = ilog2_ceil(K.nTY), bz = ilog2_ceil(K.nTZ);
	if (bx + by + bz > 30) return pmps_fail(h, PMPS_E_OVERFLOW, "tile grid %d x %d x %d is too large for the Z-order enumeration", K.nTX, K.nTY, K.nTZ);
	const int nCodes = 1 << (bx + by + bz);
	LAUNCH(h, "sort.tile_items", k_tv_items, cdiv(nCodes, 128), 128, S, h->K, nCodes, bx, by, bz);
	{ const int rc = launch_hl_build(h, "list.build", 1); if (rc) return rc; }
	CUDA_TRY(h, cudaMemsetAsync(S.hitMove, 0, sizeof(float), h->stream));   // nobody has moved since this list was made
	h->hitValid = true; h->hitForce = false;
	return PMPS_OK;
}

// Is the hit list still good for the positions of the new epoch?  It reaches hitDelta beyond every cut-off on both sides, so it
// is while nobody has moved farther than hitDelta from where it was when the list was made (hpos; track_hit_disp keeps the
// running maximum over all slots; hitMove[1] = the margin the list was made with).  One thread.
__global__ void k_hits_decide(const __grid_constant__ DevView S, const __grid_constant__ KParams K, int force) {
	const float lim = S.hitMove[1] * 0.98f;   // float positions relative to domainMin: a few 1e-7 of the domain extent of slack
	// (hidden entries — a particle left the domain in this step: which particle finds them depends on the bucket its CURRENT position is
	// in, so the list is re-made at every position epoch until the step ends; pmps_device.cuh, hid_visible)
	if (force || S.hid[0] > 0 || S.flags[FLAG_WRAP_STEP] > 0 || S.hitMove[0] > lim * lim) { S.hitFlag[0] = 1; S.hitFlag[1] += 1; S.hitMove[0] = 0.f; }
	else S.hitFlag[0] = 0;
}
// hit list for the current positions (after the prediction / correction / shifting moved the particles): re-made only when
// k_hits_decide says the old one may miss a pair (k_hl_build returns at once otherwise)
int pmps_tv_hits(pmps_handle* h) {
	DevView& S = h->V;
	if (getenv("PMPS_DEBUG")) {   // what the decision below will see
		float hm[2] = { 0.f, 0.f }; int32_t hc = 0;
		cudaMemcpyAsync(hm, S.hitMove, sizeof hm, cudaMemcpyDeviceToHost, h->stream); cudaMemcpyAsync(&hc, S.hid, sizeof hc, cudaMemcpyDeviceToHost, h->stream); cudaStreamSynchronize(h->stream);
		fprintf(stderr, "[pmps] hit-list check (walk %d): moved %.4g lo since the list was made, margin %.4g lo, forced %d, hidden entries %d\n", h->walkSeq, sqrt((double)hm[0]) / h->K.p.partDist, (double)hm[1] / h->K.p.partDist, h->hitForce ? 1 : 0, hc);
	}
	LAUNCH(h, "list.hits_decide", k_hits_decide, 1, 1, S, h->K, h->hitForce ? 1 : 0);
	h->hitForce = false;
	{ const int rc = launch_hl_build(h, "list.hits", 0); if (rc) return rc; }
	h->hitValid = true;
	return PMPS_OK;
}
