// pmps_debug.cu — parity hooks used by tests/ only (include/pmps.h "debug / parity hooks"): the bucket membership
// and the neighbour set of one particle as the device path sees them, in the reference's (host) indices.
#include "pmps_device.cuh"

template <bool POLY>
__global__ void k_debug_neighbors(const __grid_constant__ DevView S, const __grid_constant__ KParams K, int hostI, int kind, int32_t* out, int cap, int32_t* count) {
	const int i = S.slotOfHost[hostI];
	const double4 pi = S.pos[i];
	const double mix = POLY ? S.mirrorParticlePos.x[i] : 0.0, miy = POLY ? S.mirrorParticlePos.y[i] : 0.0, miz = POLY ? S.mirrorParticlePos.z[i] : 0.0;
	const uint32_t* L = S.list + (size_t)i * (size_t)S.listCap;
	const int cnt = S.listCnt[i]; int n = 0;
	for (int k = 0; k < cnt; k++) {
		Pair q; q.j = (int)L[k]; q.pj = S.pos[q.j];
		q.rx = q.pj.x - pi.x; q.ry = q.pj.y - pi.y; q.rz = q.pj.z - pi.z; q.d2 = sq3_exact(q.rx, q.ry, q.rz);
		q.mx = q.pj.x - mix; q.my = q.pj.y - miy; q.mz = q.pj.z - miz; q.dm2 = POLY ? sq3_exact(q.mx, q.my, q.mz) : 0.0;
		bool hit;
		if (kind == 0) hit = pred_P<POLY>(q, K.p.reL2); else if (kind == 1) hit = pred_P<POLY>(q, K.p.reS2);
		else if (kind == 2) hit = pred_M(q, K.p.reS2); else if (kind == 3) hit = pred_M(q, K.p.reL2);
		else hit = pred_P<POLY>(q, K.p.distCollisionLimit2);
		if (hit) { if (n < cap) out[n] = S.hostIdx[q.j]; n++; }
	}
	*count = n;
}

extern "C" int pmps_debug_neighbors(pmps_t* h, int i, int kind, int32_t* out, int cap) {
	if (!h || !out || cap < 0) return PMPS_E_ARG;
	cudaSetDevice(h->device);
	if (!h->haveList) return pmps_fail(h, PMPS_E_STATE, "pmps_debug_neighbors: call pmps_update_buckets first");
	{ const int rc = pmps_ensure_list(h); if (rc) return rc; }
	if (i < 0 || i >= h->numParticlesZero) return pmps_fail(h, PMPS_E_ARG, "pmps_debug_neighbors: index out of range");
	int32_t* d = (int32_t*)h->stage; int32_t* dcount = d + cap;
	if ((size_t)(cap + 1) * 4 > h->stageBytes) return pmps_fail(h, PMPS_E_ARG, "pmps_debug_neighbors: cap too large");
	if (h->K.poly) k_debug_neighbors<true><<<1, 1, 0, h->stream>>>(h->V, h->K, i, kind, d, cap, dcount);
	else k_debug_neighbors<false><<<1, 1, 0, h->stream>>>(h->V, h->K, i, kind, d, cap, dcount);
	int n = 0;
	CUDA_TRY(h, cudaMemcpyAsync(&n, dcount, 4, cudaMemcpyDeviceToHost, h->stream));
	CUDA_TRY(h, cudaStreamSynchronize(h->stream));
	if (n > 0) CUDA_TRY(h, cudaMemcpy(out, d, 4 * (size_t)(n < cap ? n : cap), cudaMemcpyDeviceToHost));
	return n;
}

extern "C" int pmps_debug_buckets(pmps_t* h, int32_t* bucketOfParticle, int32_t* bucketStart, int32_t* members) {
	if (!h) return PMPS_E_ARG;
	cudaSetDevice(h->device);
	if (!h->haveList) return pmps_fail(h, PMPS_E_STATE, "pmps_debug_buckets: call pmps_update_buckets first");
	const int n0 = h->numParticlesZero, nB = h->V.nB;
	CUDA_TRY(h, cudaStreamSynchronize(h->stream));
	std::vector<int32_t> cellOf(n0), hostIdx(n0);
	CUDA_TRY(h, cudaMemcpy(cellOf.data(), h->V.cellOf, 4 * (size_t)n0, cudaMemcpyDeviceToHost));
	CUDA_TRY(h, cudaMemcpy(hostIdx.data(), h->V.hostIdx, 4 * (size_t)n0, cudaMemcpyDeviceToHost));
	if (bucketStart) CUDA_TRY(h, cudaMemcpy(bucketStart, h->V.cellStart, 4 * (size_t)(nB + 1), cudaMemcpyDeviceToHost));
	for (int s = 0; s < n0; s++) {
		if (bucketOfParticle) bucketOfParticle[hostIdx[s]] = cellOf[s] < nB ? cellOf[s] : -1;
		if (members) members[s] = hostIdx[s];
	}
	return PMPS_OK;
}
