// pmps_tile.cuh — the tile walker: neighbour passes staged per bucket block in shared memory.
//
// ncu on the list walker (profiles/r01_ncu_walk_v3.txt) showed every neighbour pass latency-bound: each lane chases
// list entry -> position record -> field records through three dependent global loads per candidate.  Here a thread
// block takes one TILE (TILE_CX consecutive buckets of one bucket row, ~32 particles) and first copies every record
// its particles can possibly need — the particles of the (TILE_CX+2) x 3 x 3 surrounding buckets, which are nine (three
// in 2D) CONTIGUOUS slot ranges in the bucket-sorted arrays — into shared memory with fully coalesced, independent
// loads.  Then
//   scan  : the lanes of a particle sweep the staged records with a float distance test (coordinates relative to the
//           tile origin, so float resolves 1e-6 of the particle spacing; the threshold carries that margin) and queue
//           the survivors' indices;
//   pairs : every lane drains its own queue: exact FP64 distance, the reference's predicates, the stage bodies — all
//           operands come from shared memory, all lanes busy;
//   end   : partial sums are combined with warp shuffles, the first lane of the particle runs the body epilogues.
// No candidate list is needed: the reference's own search (27 buckets around the particle) is what gets staged, so the
// neighbour sets are the reference's as long as no particle moved farther than the bucket slack allows — the same
// condition under which the reference's stale-bucket search is complete.
//
// The stage bodies are the ones the list walker uses, untouched: pair() receives a DevView whose pos / vel / press / ...
// pointers address the staged arrays and a Pair whose j is the staged index.
#pragma once
#include "pmps_bodies.cuh"

#define TILE_T 128      // threads per block
#define TILE_CX 2       // buckets per tile along x
#define TILE_QC 32      // queue entries per lane
#define TILE_U 8        // records per lane and scan batch

enum { NF_VEL = 1, NF_PRESS = 2, NF_PNDI = 4, NF_RHO = 8, NF_MEU = 16, NF_CONC = 32, NF_PNDKI = 64 };
// neighbour fields each body reads in pair() (pos and the type word are always staged)
template <class B> struct NbFields { static constexpr unsigned value = 0; };
template <> struct NbFields<BViscosity> { static constexpr unsigned value = NF_VEL | NF_MEU; };
template <> struct NbFields<BWallViscosity> { static constexpr unsigned value = NF_VEL | NF_MEU; };
template <> struct NbFields<BCollision> { static constexpr unsigned value = NF_VEL; };
template <> struct NbFields<BPndDiffusive> { static constexpr unsigned value = NF_VEL | NF_PRESS | NF_PNDI | NF_RHO; };
template <> struct NbFields<BWallPndDiffusive> { static constexpr unsigned value = NF_VEL | NF_PNDI; };
template <> struct NbFields<BMeanPnd<0>> { static constexpr unsigned value = NF_PNDI; };
template <> struct NbFields<BMeanPnd<1>> { static constexpr unsigned value = NF_PNDI; };
template <> struct NbFields<BMeanWallPnd> { static constexpr unsigned value = NF_PNDI; };
template <> struct NbFields<BMeanNeighFluidPnd> { static constexpr unsigned value = NF_PNDKI; };
template <> struct NbFields<BPressMin> { static constexpr unsigned value = NF_PRESS; };
template <> struct NbFields<BPressGradient> { static constexpr unsigned value = NF_PRESS | NF_PNDI; };
template <> struct NbFields<BWallPressGradient> { static constexpr unsigned value = NF_PRESS | NF_PNDI; };
template <> struct NbFields<BGradConcentration> { static constexpr unsigned value = NF_CONC; };
template <> struct NbFields<BWallGradConcentration> { static constexpr unsigned value = NF_CONC; };
template <> struct NbFields<BExtrapolatePress> { static constexpr unsigned value = NF_PRESS | NF_RHO; };
template <> struct NbFields<BWallDummyVel> { static constexpr unsigned value = NF_VEL; };
template <class... B> struct PackFields { static constexpr unsigned value = 0; };
template <class B, class... R> struct PackFields<B, R...> { static constexpr unsigned value = NbFields<B>::value | PackFields<R...>::value; };

__host__ __device__ constexpr int tile_rec_bytes(unsigned f) {
	return 16 + 32 + ((f & NF_VEL) ? 32 : 0) + 8 * (((f & NF_PRESS) ? 1 : 0) + ((f & NF_PNDI) ? 1 : 0) + ((f & NF_RHO) ? 1 : 0) + ((f & NF_MEU) ? 1 : 0) + ((f & NF_CONC) ? 1 : 0) + ((f & NF_PNDKI) ? 1 : 0));
}

// the non-empty tiles of the current bucket order (one thread per tile; warp-aggregated append)
__global__ void k_tile_list(const __grid_constant__ DevView S, const __grid_constant__ KParams K, int nTilesAll, int nXB) {
	const int t = blockIdx.x * blockDim.x + threadIdx.x;
	bool have = false;
	if (t < nTilesAll) {
		const int nX = K.p.numBucketsX;
		const int row = t / nXB, xb = t - row * nXB;
		const int c0 = row * nX + xb * TILE_CX, c1 = min(c0 + TILE_CX, row * nX + nX);
		have = S.cellStart[c1] > S.cellStart[c0];
	}
	const unsigned b = __ballot_sync(0xffffffffu, have);
	if (b == 0) return;
	const int lane = threadIdx.x & 31;
	int base = 0;
	if (lane == 0) base = atomicAdd(S.nTiles, __popc(b));
	base = __shfl_sync(0xffffffffu, base, 0);
	if (have) S.tileList[base + __popc(b & ((1u << lane) - 1u))] = t;
}
__global__ void k_copy_ghost_tail(const __grid_constant__ DevView S, int copyPos, int copyVel) {
	const int i = S.cellStart[S.nB] + blockIdx.x * blockDim.x + threadIdx.x;   // ghosts are parked behind the last bucket
	if (i >= S.n0) return;
	if (copyPos) S.pos2[i] = S.pos[i];
	if (copyVel) S.vel2[i] = S.vel[i];
}

template <bool POLY, unsigned FIELDS, class F>
__global__ void __launch_bounds__(TILE_T, 2) k_tile(const __grid_constant__ DevView S, const __grid_constant__ KParams K, int maxM) {
	extern __shared__ __align__(32) unsigned char smemRaw[];
	__shared__ int sRun[9], sOff[10], sOwn[2];
	// carve the staging arrays
	unsigned char* sp = smemRaw;
	double4* spos = (double4*)sp; sp += 32 * (size_t)maxM;
	double4* svel = nullptr; if (FIELDS & NF_VEL) { svel = (double4*)sp; sp += 32 * (size_t)maxM; }
	float4* fpos = (float4*)sp; sp += 16 * (size_t)maxM;
	double *spress = nullptr, *spndi = nullptr, *srho = nullptr, *smeu = nullptr, *sconc = nullptr, *spndki = nullptr;
	if (FIELDS & NF_PRESS) { spress = (double*)sp; sp += 8 * (size_t)maxM; }
	if (FIELDS & NF_PNDI) { spndi = (double*)sp; sp += 8 * (size_t)maxM; }
	if (FIELDS & NF_RHO) { srho = (double*)sp; sp += 8 * (size_t)maxM; }
	if (FIELDS & NF_MEU) { smeu = (double*)sp; sp += 8 * (size_t)maxM; }
	if (FIELDS & NF_CONC) { sconc = (double*)sp; sp += 8 * (size_t)maxM; }
	if (FIELDS & NF_PNDKI) { spndki = (double*)sp; sp += 8 * (size_t)maxM; }
	uint16_t* q = (uint16_t*)sp + threadIdx.x;   // this thread's queue column
	// the view pair() sees: neighbour fields come from the staged arrays, indexed by the staged index
	DevView T = S;
	T.pos = spos; T.vel = svel; T.press = spress; T.pndi = spndi; T.RHO = srho; T.MEU = smeu; T.concentration = sconc; T.pndki = spndki;

	const int tid = threadIdx.x;
	const int nX = K.p.numBucketsX, nY = K.p.numBucketsY, nZ = K.p.numBucketsZ, nXY = nX * nY;
	const int nXB = (nX + TILE_CX - 1) / TILE_CX;
	const bool three = K.p.dim == 3;
	const int nRuns = three ? 9 : 3;
	const int nTiles = *S.nTiles;
	const float rfilt2 = (float)(K.reMax2 * (1.0 + 1e-5));
	for (int tt = blockIdx.x; tt < nTiles; tt += gridDim.x) {
		const int tile = S.tileList[tt];
		const int row = tile / nXB, xb = tile - row * nXB;
		const int cz = row / nY, cy = row - cz * nY, cx0 = xb * TILE_CX;
		__syncthreads();   // the previous tile is done with the run table and the staged data
		if (tid < nRuns) {   // run r: bucket row (cy + dy, cz + dz), buckets cx0-1 .. cx0+TILE_CX; order jz outer, jy inner
			const int dz = three ? tid / 3 - 1 : 0, dy = three ? tid % 3 - 1 : tid - 1;
			const int jy = cy + dy, jz = cz + dz;
			int s0 = 0, len = 0;
			if (jy >= 0 && jy < nY && jz >= 0 && jz < nZ) {
				const int r0 = jz * nXY + jy * nX;
				const int xa = max(cx0 - 1, 0), xe = min(cx0 + TILE_CX + 1, nX);
				s0 = S.cellStart[r0 + xa]; len = S.cellStart[r0 + xe] - s0;
			}
			sRun[tid] = s0; sOff[tid + 1] = len;
		}
		if (tid == 32) { const int c0 = row * nX + cx0, c1 = min(c0 + TILE_CX, row * nX + nX); sOwn[0] = S.cellStart[c0]; sOwn[1] = S.cellStart[c1]; }
		__syncthreads();
		if (tid == 0) { sOff[0] = 0; for (int r = 0; r < nRuns; r++) sOff[r + 1] += sOff[r]; }
		__syncthreads();
		const int M = sOff[nRuns], own0 = sOwn[0], P = sOwn[1] - own0;
		const int centre = three ? 4 : 1;
		const int lpp = P <= 16 ? 8 : P <= 32 ? 4 : P <= 64 ? 2 : 1;   // lanes per particle
		const int ppp = TILE_T / lpp;                                   // particles per pass
		// tile origin for the float coordinates of the scan
		const double ox = K.p.domainMinX + (double)(cx0 - 1) * K.p.bucketSide, oy = K.p.domainMinY + (double)(cy - 1) * K.p.bucketSide,
		             oz = three ? K.p.domainMinZ + (double)(cz - 1) * K.p.bucketSide : 0.0;
		int stagedChunk = -1;
		for (int p0 = 0; p0 < P; p0 += ppp) {
			const int p = p0 + tid / lpp, l = tid % lpp;
			const int i = own0 + p;
			const unsigned gmask = (lpp == 32 ? 0xffffffffu : ((1u << lpp) - 1u)) << ((tid & 31) & ~(lpp - 1));
			Self s; typename F::Ctx ctx; bool on = false, pairs = false;
			s.i = i; s.lead = l == 0; s.nearWallFluid = false; s.mix = s.miy = s.miz = 0.0; s.type = K.p.ghost; s.ptype = 0;
			if (p < P) {
				s.pi = ld256(S.pos + i); s.type = pk_type(s.pi); s.ptype = pk_ptype(s.pi);
				if (F::kCopyPos && s.lead) S.pos2[i] = s.pi;
				if (F::kCopyVel && s.lead) S.vel2[i] = S.vel[i];
				if (s.type != K.p.ghost) {
					if (POLY) { s.mix = S.mirrorParticlePos.x[i]; s.miy = S.mirrorParticlePos.y[i]; s.miz = S.mirrorParticlePos.z[i]; }
					if (POLY && F::kWall) { s.nearWallFluid = s.type == K.p.fluid && S.particleNearWall[i]; if (s.nearWallFluid) self_wall_frame(s); }
					on = F::begin(ctx, S, K, s);
					pairs = on && F::need_pairs(ctx);
				}
			}
			if (__syncthreads_or(pairs)) {
				const float fx = pairs ? (float)(s.pi.x - ox) : 0.f, fy = pairs ? (float)(s.pi.y - oy) : 0.f, fz = pairs ? (float)(s.pi.z - oz) : 0.f;
				const int mself = sOff[centre] + (i - sRun[centre]);   // own record in the concatenated runs
				for (int c0 = 0; c0 < M; c0 += maxM) {
					const int mc = min(maxM, M - c0);
					if (stagedChunk != c0) {
						__syncthreads();
						for (int idx = tid; idx < mc; idx += TILE_T) {
							const int g = c0 + idx;
							int r = 0;
							#pragma unroll
							for (int u = 1; u < 9; u++) if (u < nRuns && g >= sOff[u]) r = u;
							const int j = sRun[r] + (g - sOff[r]);
							const double4 pj = ld256(S.pos + j);
							spos[idx] = pj;
							fpos[idx] = make_float4((float)(pj.x - ox), (float)(pj.y - oy), (float)(pj.z - oz), 0.f);
							if (FIELDS & NF_VEL) svel[idx] = ld256(S.vel + j);
							if (FIELDS & NF_PRESS) spress[idx] = S.press[j];
							if (FIELDS & NF_PNDI) spndi[idx] = S.pndi[j];
							if (FIELDS & NF_RHO) srho[idx] = S.RHO[j];
							if (FIELDS & NF_MEU) smeu[idx] = S.MEU[j];
							if (FIELDS & NF_CONC) sconc[idx] = S.concentration[j];
							if (FIELDS & NF_PNDKI) spndki[idx] = S.pndki[j];
						}
						__syncthreads();
						stagedChunk = c0;
					}
					// scan + pairs
					// batches of TILE_U records per lane: the loads and the distance arithmetic of a batch are independent, the
					// queue (interleaved over the block: entry k of thread t at q[k*TILE_T + t]) is drained when a lane could overflow
					int qn = 0;
					const int selfRel = mself - c0;
					for (int mm = 0; mm < mc; mm += lpp * TILE_U) {
						unsigned hits = 0;
						if (pairs) {
							float d2[TILE_U];
							#pragma unroll
							for (int u = 0; u < TILE_U; u++) {
								const int m = mm + u * lpp + l;
								const float4 f = fpos[m < mc ? m : 0];
								const float dx = f.x - fx, dy = f.y - fy, dz = f.z - fz;
								d2[u] = dx * dx + dy * dy + dz * dz;
							}
							#pragma unroll
							for (int u = 0; u < TILE_U; u++) {
								const int m = mm + u * lpp + l;
								if (d2[u] < rfilt2 && m < mc && m != selfRel) hits |= 1u << u;
							}
						}
						while (hits) { const int u = __ffs(hits) - 1; hits &= hits - 1; q[(size_t)qn * TILE_T] = (uint16_t)(mm + u * lpp + l); qn++; }
						const bool last = mm + lpp * TILE_U >= mc;
						if (last || __any_sync(0xffffffffu, qn > TILE_QC - TILE_U)) {
							for (int k = 0; k < qn; k++) {
								const int m2 = q[(size_t)k * TILE_T];
								walk_one<POLY, F>(ctx, T, K, s, m2, spos[m2]);
							}
							qn = 0;
						}
					}
				}
			}
			if (on) {
				F::reduce(ctx, gmask);
				if (s.lead) F::end(ctx, S, K, s);
			}
		}
	}
}
