// pmps_debug.cu — parity hooks used by tests/ only (include/pmps.h "debug / parity hooks"): the bucket membership
// and the neighbour set of one particle as the device path sees them, in the reference's (host) indices.
#include "pmps_tv.cuh"
#include <vector>

// The neighbour set of one particle exactly as the tile passes find it: the work item that owns the slot, its halo
// table, the particle's tile-local hit list mapped back to slots, then the reference's predicates (exact).
template <bool POLY>
__global__ void __launch_bounds__(TV_T) k_debug_neighbors(const __grid_constant__ DevView S, const __grid_constant__ KParams K, int hostI, int kind, int32_t* out, int cap, int32_t* count) {
	__shared__ int hBase[TV_HMAX], hOff[TV_HMAX + 1], sItem;
	const int i = S.slotOfHost[hostI];
	const int ci = S.cellOf[i];
	if (threadIdx.x == 0) { sItem = 0x7fffffff; *count = 0; }
	__syncthreads();
	if (ci >= S.nC) return;
	const int nItems = *S.nTiles;
	for (int t = threadIdx.x; t < nItems; t += TV_T) { const int it = S.tileList[t]; if (it == ci / K.TV || it == ~ci) sItem = it; }
	__syncthreads();
	if (sItem == 0x7fffffff) { if (threadIdx.x == 0) *count = -1; return; }
	const TvBox b = tv_box(K, sItem);
	tv_halo_table(S, K, b, hBase, hOff);
	if (threadIdx.x != 0) return;
	const int M = hOff[b.HV];
	const double4 pi = S.pos[i];
	const double mix = POLY ? S.mirrorParticlePos.x[i] : 0.0, miy = POLY ? S.mirrorParticlePos.y[i] : 0.0, miz = POLY ? S.mirrorParticlePos.z[i] : 0.0;
	const uint16_t* L = S.hlist + (size_t)i * (size_t)S.tlCap;
	int n = 0;
	for (int e = 0; e < S.hcnt[i]; e++) {
		const int m = L[tv_list_pos(e)];
		if (m >= M) continue;   // padding
		int hb = 0; while (hOff[hb + 1] <= m) hb++;
		const int j = hBase[hb] + (m - hOff[hb]);
		double4 pj = S.pos[j];
		if (K.p.domainTypeBC == 1) per_move(K, pj, tv_halo_shift(K, b, hb));   // periodic domain: the records of an aliased border layer
		Pair q; q.j = j; q.tj = pk_type(pj); q.ptj = pk_ptype(pj);
		q.rx = pj.x - pi.x; q.ry = pj.y - pi.y; q.rz = pj.z - pi.z; q.d2 = sq3_exact(q.rx, q.ry, q.rz);
		q.mx = pj.x - mix; q.my = pj.y - miy; q.mz = pj.z - miz; q.dm2 = POLY ? sq3_exact(q.mx, q.my, q.mz) : 0.0;
		bool hit;
		if (kind == 0) hit = pred_P<POLY>(q, K.p.reL2); else if (kind == 1) hit = pred_P<POLY>(q, K.p.reS2);
		else if (kind == 2) hit = pred_M(q, K.p.reS2); else if (kind == 3) hit = pred_M(q, K.p.reL2);
		else hit = pred_P<POLY>(q, K.p.distCollisionLimit2);
		if (hit) { if (n < cap) out[n] = S.hostIdx[j]; n++; }
	}
	*count = n;
}

extern "C" int pmps_debug_neighbors(pmps_t* h, int i, int kind, int32_t* out, int cap) {
	if (!h || !out || cap < 0) return PMPS_E_ARG;
	cudaSetDevice(h->device);
	if (!h->haveList) return pmps_fail(h, PMPS_E_STATE, "pmps_debug_neighbors: call pmps_update_buckets first");
	if (i < 0 || i >= h->numParticlesZero) return pmps_fail(h, PMPS_E_ARG, "pmps_debug_neighbors: index out of range");
	if (!h->hitValid) { const int rc = pmps_tv_hits(h); if (rc) return rc; }   // the hit list of the CURRENT positions, as the next pass would use it
	int32_t* d = (int32_t*)h->stage; int32_t* dcount = d + cap;
	if ((size_t)(cap + 1) * 4 > h->stageBytes) return pmps_fail(h, PMPS_E_ARG, "pmps_debug_neighbors: cap too large");
	if (h->K.poly) k_debug_neighbors<true><<<1, TV_T, 0, h->stream>>>(h->V, h->K, i, kind, d, cap, dcount);
	else k_debug_neighbors<false><<<1, TV_T, 0, h->stream>>>(h->V, h->K, i, kind, d, cap, dcount);
	int n = 0;
	CUDA_TRY(h, cudaMemcpyAsync(&n, dcount, 4, cudaMemcpyDeviceToHost, h->stream));
	CUDA_TRY(h, cudaStreamSynchronize(h->stream));
	if (n < 0) return pmps_fail(h, PMPS_E_STATE, "pmps_debug_neighbors: no work item owns the particle's tile");
	if (n > 0) CUDA_TRY(h, cudaMemcpy(out, d, 4 * (size_t)(n < cap ? n : cap), cudaMemcpyDeviceToHost));
	return n;
}

extern "C" int pmps_debug_buckets(pmps_t* h, int32_t* bucketOfParticle, int32_t* bucketStart, int32_t* members) {
	if (!h) return PMPS_E_ARG;
	cudaSetDevice(h->device);
	if (!h->haveList) return pmps_fail(h, PMPS_E_STATE, "pmps_debug_buckets: call pmps_update_buckets first");
	const int n0 = h->numParticlesZero, nB = h->V.nB;
	CUDA_TRY(h, cudaStreamSynchronize(h->stream));
	std::vector<int32_t> cellOf(n0), hostIdx(n0), refB(n0);
	CUDA_TRY(h, cudaMemcpy(cellOf.data(), h->V.cellOf, 4 * (size_t)n0, cudaMemcpyDeviceToHost));
	CUDA_TRY(h, cudaMemcpy(hostIdx.data(), h->V.hostIdx, 4 * (size_t)n0, cudaMemcpyDeviceToHost));
	// device cells are tile-major; the reference numbers buckets iz*nXY + iy*nX + ix (src/MpsBucket.cpp:336)
	const int nX = h->K.p.numBucketsX, nY = h->K.p.numBucketsY;
	std::vector<int32_t> count(nB + 1, 0);
	for (int s = 0; s < n0; s++) {
		int b = -1;
		if (cellOf[s] < h->V.nC) { int ix, iy, iz; cell_coords(h->K, cellOf[s], ix, iy, iz); b = iz * nX * nY + iy * nX + ix; count[b + 1]++; }
		refB[s] = b;
		if (bucketOfParticle) bucketOfParticle[hostIdx[s]] = b;
	}
	for (int b = 0; b < nB; b++) count[b + 1] += count[b];
	if (bucketStart) for (int b = 0; b <= nB; b++) bucketStart[b] = count[b];
	if (members) {   // slot order inside a bucket is ascending host index = the reference's list order
		std::vector<int32_t> fill(count.begin(), count.end() - 1);
		int tail = count[nB];
		for (int s = 0; s < n0; s++) { if (refB[s] >= 0) members[fill[refB[s]]++] = hostIdx[s]; else members[tail++] = hostIdx[s]; }
	}
	return PMPS_OK;
}
