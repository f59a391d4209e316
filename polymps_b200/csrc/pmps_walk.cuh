// pmps_walk.cuh — walk<B...>(): one neighbour pass of the tile-Verlet walker (pmps_tv.cuh) with the bodies B... fused.
// Included by the translation units that instantiate the walker kernels (pmps_walk_inst.inc), possibly INSIDE the namespace of an
// option set (pmps_optset.h) — hence no include guard games here: system headers and pmps_internal.h come before the namespace.
// The packs the stage calls and the fused step use are listed in pmps_walk_packs.h.
#pragma once
#include "pmps_tv.cuh"


// per tile: does it hold a particle near a polygon wall?  The wall pass sets the flags as it goes (fused step path);
// a stage call that finds them stale recomputes them from particleNearWall.
static __global__ void k_tile_wall_flags(const __grid_constant__ DevView S, const __grid_constant__ KParams K) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= S.n) return;
	if (S.particleNearWall[i] && S.cellOf[i] < S.nC) S.tileWall[S.cellOf[i] / K.TV] = 1;
}

// One neighbour pass over every particle with the bodies B... fused (tile-Verlet walker, pmps_tv.cuh).  polyOverride = 0
// runs the pass with the particle-wall predicate even for polygon walls (the collision pass: mirrors are at "infinity"
// at that point).
template <class... B> int walk(pmps_handle* h, const char* name, int polyOverride) {
	const KParams& K = *reinterpret_cast<const KParams*>(&h->K);   // (an option set's KParams: same layout, pmps_optset.h)
	const bool poly = polyOverride < 0 ? (K.poly != 0) : (polyOverride != 0);
	typedef Fused<B...> F;
	constexpr unsigned FIELDS = PackFields<B...>::value;
	constexpr bool WALLONLY = PackWallOnly<B...>::value;
	constexpr bool SPARSE = !WALLONLY && PackSparse<B...>::value;
	constexpr int rec = tv_rec_bytes(FIELDS);
	static_assert(rec <= TV_REC_MAX, "this pack stages more bytes per record than the tile capacity was sized for");
	if (h->walkSeq >= PMPS_CURSORS) return pmps_fail(h, PMPS_E_STATE, "%s: more than %d neighbour passes since the last pmps_update_buckets", name, PMPS_CURSORS - 1);
	if (WALLONLY && !h->tileWallValid) {
		CUDA_TRY(h, cudaMemsetAsync(h->V.tileWall, 0, (size_t)h->nTilesAll, h->stream));
		LAUNCH(h, "wall.tile_flags", k_tile_wall_flags, cdiv(h->V.n, 256), 256, h->V, K);
		h->tileWallValid = true;
	}
	// candidates: the hit list of the current position epoch; a pass that finds it stale has it checked (and re-made when somebody
	// has moved farther than its margin) first
	if (!h->hitValid) { const int rc = pmps_tv_hits(h); if (rc) return rc; }   // cheap when the old hit list still covers the new positions
	if constexpr (WALLONLY || F::kCloseOnly || SPARSE) {   // passes that act on few tiles: compacted work-item list for k_tv, tails of the rest by k_tails
		constexpr int KIND = WALLONLY ? 0 : SPARSE ? 2 : 1;
		if constexpr (SPARSE) {
			CUDA_TRY(h, cudaMemsetAsync(h->V.tileSparse, 0, (size_t)h->nTilesAll, h->stream));
			ProfScope pm_(h, name);
			k_sparse_mark<F><<<cdiv(h->V.n, 256), 256, 0, h->stream>>>(h->V, K); h->launches++;
		}
		CUDA_TRY(h, cudaMemsetAsync(h->V.actCount, 0, 4 * sizeof(int32_t), h->stream));
		LAUNCH(h, "tiles.item_flags", k_item_flags, cdiv((h->V.nC < h->V.n0 ? h->V.nC : h->V.n0) + 1, 256), 256, h->V, K);   // at most one work item per non-empty cell
		ProfScope pt_(h, name);
		k_tails<KIND, F><<<cdiv(h->V.n, 256), 256, 0, h->stream>>>(h->V, K); h->launches++;
	}
	const size_t smem = (size_t)TV_MAXM * rec + (size_t)TV_FIXED_BYTES;
	const int grid = h->smCount * TV_BLOCKS(FIELDS);
	ProfScope ps_(h, name);
	const int seq = h->walkSeq++;
	// (first launch of an instantiation: opt in to the dynamic shared memory and check that as many blocks fit an SM as the pass was sized for)
#define TV_LAUNCH(P_) do { static bool once_ = false; if (!once_) { once_ = true; \
			CUDA_TRY(h, cudaFuncSetAttribute(k_tv<P_, FIELDS, WALLONLY, SPARSE, F>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
			int nb_ = 0; cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb_, k_tv<P_, FIELDS, WALLONLY, SPARSE, F>, TV_T, smem); \
			if (nb_ < TV_BLOCKS(FIELDS)) fprintf(stderr, "[pmps] %s: %d blocks per SM resident, sized for %d (shared memory / registers)\n", name, nb_, TV_BLOCKS(FIELDS)); } \
		k_tv<P_, FIELDS, WALLONLY, SPARSE, F><<<grid, TV_T, smem, h->stream>>>(h->V, K, seq); } while (0)
	if (poly) TV_LAUNCH(true); else TV_LAUNCH(false);
#undef TV_LAUNCH
	h->launches++;
	if (F::kCopyPos || F::kCopyVel) { k_copy_ghost_tail<<<cdiv(h->V.n0, 256), 256, 0, h->stream>>>(h->V, F::kCopyPos ? 1 : 0, F::kCopyVel ? 1 : 0); h->launches++; }
	CUDA_TRY(h, cudaGetLastError());
	return PMPS_OK;
}

