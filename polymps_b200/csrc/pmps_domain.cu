// pmps_domain.cu — MpsParticle::checkParticleOutDomain (reference src/MpsParticle.cpp:234-402), the periodic wrap included.
//
// The reference removes a fluid particle that left [domainMin+bucketSide, domainMax-bucketSide] by copying the LAST
// live particle over it and turning the last array element into a ghost; the loop is serial and order dependent.  On the
// device a flag kernel collects the (normally zero) offenders; a one-thread kernel then replays the reference's
// serial loop over just those host indices in ascending order, re-testing each against the CURRENT content of its
// slot exactly as the reference's loop does.  No host round trip: the fix kernel returns at once when the count is 0.
#include "pmps_device.cuh"

__device__ __forceinline__ bool out_of_domain(const KParams& K, const double4& p) {
	const PmpsParams& P = K.p;
	const double lx = P.domainMinX + P.bucketSide, ly = P.domainMinY + P.bucketSide, lz = P.domainMinZ + P.bucketSide;
	const double hx = P.domainMaxX - P.bucketSide, hy = P.domainMaxY - P.bucketSide, hz = P.domainMaxZ - P.bucketSide;
	return p.x > hx || p.x < lx || p.y > hy || p.y < ly || (P.dim == 3 && (p.z > hz || p.z < lz));
}

// Periodic domain: a particle beyond a face of the physical domain re-enters through the opposite one (src/MpsParticle.cpp:236-269, every
// particle, before the out-of-domain test).  The reference's bucket lists still hold it where it started the step — a domain length
// away from where it now is — so until the lists are re-threaded at the next step start nobody finds it, while its own search runs
// around its new position: the slot gets its hidden flag (DevView::wrapHid, carried through the bucket sort; a particle wrapped back
// within the same step loses it again) and the host re-sorts the buckets before the next pass (pmps_check_out_domain).
__device__ __forceinline__ void periodic_wrap(const DevView& S, const KParams& K, int i, double4& p) {
	double& x = K.p.periodicAxis == 0 ? p.x : K.p.periodicAxis == 1 ? p.y : p.z;
	const double x0 = x;
	if (x > K.p.physDomMax) x = __dsub_rn(x, K.p.periodicLength);
	if (x < K.p.physDomMin) x = __dadd_rn(x, K.p.periodicLength);
	if (x == x0) return;
	S.pos[i] = p;
	S.wrapHid[i] ^= 1;
	atomicAdd(&S.flags[FLAG_WRAPPED], 1);
}

__global__ void k_ood_flag(const __grid_constant__ DevView S, const __grid_constant__ KParams K) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= S.n) return;
	double4 p = S.pos[i];
	if (K.p.domainTypeBC == 1 && pk_type(p) != K.p.ghost && owned_slot(S, i)) periodic_wrap(S, K, i, p);
	if (pk_type(p) != K.p.fluid || !out_of_domain(K, p) || !owned_slot(S, i)) return;
	const int k = atomicAdd(&S.flags[FLAG_OOD_COUNT], 1);
	if (k < PMPS_OOD_CAP) S.oodList[k] = S.hostIdx[i];
}

__global__ void k_ood_fix(const __grid_constant__ DevView S, const __grid_constant__ KParams K, int staleReplay) {
	if (K.p.domainTypeBC == 1) {   // FLAG_WRAP_LAST = particles this check moved through a face (host: re-sort), FLAG_WRAP_STEP = since the step start
		const int w = S.flags[FLAG_WRAPPED];
		if (w > 0) { S.flags[FLAG_WRAPPED] = 0; S.flags[FLAG_WRAP_LAST] += w; S.flags[FLAG_WRAP_STEP] += w; S.hitMove[0] = 1e30f; }
	}
	int cnt = S.flags[FLAG_OOD_COUNT];
	if (cnt == 0) return;
	S.flags[FLAG_OOD_COUNT] = 0;
	if (cnt > PMPS_OOD_CAP) { atomicOr(&S.flags[FLAG_GUARD], 4); cnt = PMPS_OOD_CAP; }
	int32_t* L = S.oodList;
	for (int a = 1; a < cnt; a++) { const int v = L[a]; int b = a; while (b > 0 && L[b - 1] > v) { L[b] = L[b - 1]; b--; } L[b] = v; }
	int live = S.flags[FLAG_LIVE];
	// stale-list visibility (pmps_device.cuh, hid_visible): array element hi keeps its place in the reference's bucket lists — the
	// bucket the particle that left was threaded on at the step start — while its content becomes the moved particle
	int32_t* const H = S.hid; int nh = H[0];
	auto hid_find = [&](int label) { for (int q = 0; q < nh; q++) if (H[1 + q] == label) return q; return -1; };
	for (int e = 0; e < cnt; e++) {
		const int hi = L[e];
		if (!(hi < live)) continue;
		const int si = S.slotOfHost[hi];
		const double4 pcur = S.pos[si];
		if (pk_type(pcur) != K.p.fluid || !out_of_domain(K, pcur)) continue;
		if (S.nOwn != 0x7fffffff) {
			// slab rank (pmps_slab.cu): the last live row of the table is a halo copy, not a particle to move into the hole.  The
			// particle becomes a ghost IN PLACE; the row is reclaimed at the next re-partition (the owner drops ghosts there), the
			// neighbours' copies turn into ghosts with the next position refresh (the type travels with the position).
			double4 g; g.x = K.p.domainMaxX - K.p.partDist; g.y = K.p.domainMaxY - K.p.partDist;
			g.z = K.p.dim == 2 ? 0.0 : K.p.domainMaxZ - K.p.partDist; g.w = pk_make(K.p.ghost, pk_ptype(pcur));
			S.pos[si] = g; double4 vz = S.vel[si]; vz.x = vz.y = vz.z = 0.0; S.vel[si] = vz;
			S.particleBC[si] = K.p.other; S.particleNearWall[si] = 0; S.nearMeshType[si] = 0; S.distParticleWall2[si] = K.big; S.press[si] = 0.0;
			S.flags[FLAG_OOD_TOTAL] += 1;
			continue;
		}
		const int hl = live - 1; const int sl = S.slotOfHost[hl];
		// The reference copies particle hl over particle hi and turns hl into a ghost.  In slot space the DATA stays where
		// it is (slot sl keeps searching around its own position with its own candidate list); the host LABELS are swapped:
		// slot sl now answers to host index hi, slot si — which receives a copy of sl's fields, as array element hl keeps
		// them in the reference — becomes the ghost hl.
		const int cellOfHole = S.stepCell[si];
		{ const int q = hid_find(hl); if (q >= 0) H[1 + q] = -1; }   // the last element becomes a ghost: whatever it was, it is listed nowhere that matters
		if (sl != si) {
			if (staleReplay && hid_find(hi) < 0) {   // (a hole filled twice in one step keeps the bucket of its first occupant: the lists have not changed)
				if (nh < PMPS_HID_CAP) { H[1 + nh] = hi; H[1 + PMPS_HID_CAP + nh] = cellOfHole; nh++; }
				else atomicOr(&S.flags[FLAG_GUARD], 8);
			}
			S.pos[si] = S.pos[sl]; S.vel[si] = S.vel[sl]; S.particleNearWall[si] = S.particleNearWall[sl];
#define X(n) S.n[si] = S.n[sl];
			PMPS_I32_FIELDS(X)
			PMPS_F64_FIELDS(X)
#undef X
			if (K.p.fluidType == 1) {   // src/MpsParticle.cpp:322-337
#define X(n) S.n[si] = S.n[sl];
				PMPS_RHEO_FIELDS(X)
#undef X
			}
#define X(n) S.n.x[si] = S.n.x[sl]; S.n.y[si] = S.n.y[sl]; S.n.z[si] = S.n.z[sl];
			PMPS_V3_FIELDS(X)
#undef X
			S.hostIdx[sl] = hi; S.slotOfHost[hi] = sl; S.hostIdx[si] = hl; S.slotOfHost[hl] = si;
		}
		const int sg = si;   // the slot that becomes the ghost
		double4 g; g.x = K.p.domainMaxX - K.p.partDist; g.y = K.p.domainMaxY - K.p.partDist;
		g.z = K.p.dim == 2 ? 0.0 : K.p.domainMaxZ - K.p.partDist; g.w = pk_make(K.p.ghost, pk_ptype(S.pos[sg]));
		S.pos[sg] = g; double4 vz = S.vel[sg]; vz.x = vz.y = vz.z = 0.0; S.vel[sg] = vz;
		S.particleBC[sg] = K.p.other; S.particleNearWall[sg] = 0; S.nearMeshType[sg] = 0; S.distParticleWall2[sg] = K.big;
		S.press[sg] = 0.0;
		live--; S.flags[FLAG_OOD_TOTAL] += 1;
	}
	S.flags[FLAG_LIVE] = live; H[0] = nh;
	S.hitMove[0] = 1e30f;   // a slot changed its content: the hit lists are rebuilt at the next check
}

extern "C" int pmps_check_out_domain(pmps_t* h) {
	if (!h) return PMPS_E_ARG;
	if (!h->haveList) return pmps_fail(h, PMPS_E_STATE, "pmps_check_out_domain: call pmps_update_buckets first");
	LAUNCH(h, "domain.flag", k_ood_flag, cdiv(h->V.n, 256), 256, h->V, h->K);
	LAUNCH(h, "domain.fix", k_ood_fix, 1, 1, h->V, h->K, h->staleReplay ? 1 : 0);
	CUDA_TRY(h, cudaGetLastError());
	if (h->K.p.domainTypeBC == 1) {
		// periodic domain: a wrapped particle now sits a domain length away from the bucket it is sorted into; the tiles see a particle
		// through its bucket, so the bucket order is re-made at once (one read-back per check; no particle wrapped = nothing to do)
		int32_t w = 0;
		CUDA_TRY(h, cudaMemcpyAsync(&w, h->V.flags + FLAG_WRAP_LAST, sizeof w, cudaMemcpyDeviceToHost, h->stream));
		CUDA_TRY(h, cudaStreamSynchronize(h->stream));
		if (w > 0) {
			CUDA_TRY(h, cudaMemsetAsync(h->V.flags + FLAG_WRAP_LAST, 0, sizeof(int32_t), h->stream));
			h->nWrapped += w;
			const int rc = pmps_rebuild_now(h); if (rc) return rc;
		}
	}
	return PMPS_OK;
}

extern "C" int pmps_set_stale_list_replay(pmps_t* h, int on) { if (!h) return PMPS_E_ARG; h->staleReplay = on != 0; return PMPS_OK; }
extern "C" int pmps_debug_set_stale(pmps_t* h, int n, const int32_t* hostIdx, const int32_t* bucket) {
	if (!h || n < 0 || n > PMPS_HID_CAP || (n > 0 && (!hostIdx || !bucket))) return PMPS_E_ARG;
	cudaSetDevice(h->device);
	if (!h->haveState) return pmps_fail(h, PMPS_E_STATE, "pmps_debug_set_stale: nothing uploaded");
	CUDA_TRY(h, cudaStreamSynchronize(h->stream));
	CUDA_TRY(h, cudaMemcpy(h->V.hid, &n, sizeof(int32_t), cudaMemcpyHostToDevice));
	if (n > 0) {
		CUDA_TRY(h, cudaMemcpy(h->V.hid + 1, hostIdx, sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice));
		CUDA_TRY(h, cudaMemcpy(h->V.hid + 1 + PMPS_HID_CAP, bucket, sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice));
	}
	h->hitValid = false;   // the next pass has the hit list re-made with these entries (k_hits_decide)
	return PMPS_OK;
}
