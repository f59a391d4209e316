// pmps_collide.cu — MpsParticleCollision::checkParticleCollisions (src/MpsParticleCollision.cpp:22-119) from the close-pair lists.
//
// The collision model acts on pairs closer than distCollisionLimit (0.85 lo): a handful of the ~46 candidates a particle's hit list
// holds.  Run through the tile walker the pass staged a whole halo (positions + velocities) to reject 40 of them — 2.7 ms of a
// 29 ms flowing step at 10 M particles.  The hit-list build already knows which of a particle's hits are that close when it makes
// the list (within closeR = the limit + twice the list margin, so the set holds as long as the list does) and writes them down as
// GLOBAL slots (DevView::clist); here one thread per fluid particle walks its few entries, reads the partners' position and velocity
// records straight from global memory (neighbours in slot order are neighbours in memory: the sectors are shared by the threads of a
// block) and applies both loops of the reference: dvel from the OLD velocities, then vel2 = vel + dvel (Jacobi: neighbours still read
// vel in this launch).  The predicate ignores the mirror particle (it was reset to "infinity" by the prediction, src/main.cpp:286-290).
#include "pmps_device.cuh"

template <bool FALLBACK, bool PER> __device__ __forceinline__ void collide_pair(const DevView& S, const KParams& K, int i, const double4& pi, const double4& vi, double mi, int j, double& dx, double& dy, double& dz) {
	double4 pj = ld256(S.pos + j);
	if (PER) per_move(K, pj, per_pair_shift(S, K, i, j));   // periodic domain: the partner across the face, as i sees it
	const double rx = pj.x - pi.x, ry = pj.y - pi.y, rz = pj.z - pi.z; const double d2 = sq3_exact(rx, ry, rz);
	if (!(d2 < K.p.distCollisionLimit2)) return;
	const int tj = pk_type(pj);
	if (FALLBACK && tj == K.p.ghost) return;
	const double4 vj = ld256(S.vel + j);
	double f = (vi.x - vj.x) * rx + (vi.y - vj.y) * ry + (vi.z - vj.z) * rz;
	if (f > 0.0) {
		const bool jf = tj == K.p.fluid;
		const double mj = jf ? (pk_ptype(pj) == 1 ? K.p.DNS_FL1 : K.p.DNS_FL2) : K.p.densityWall;
		f *= K.p.restitutionCollision * mj / (mi + mj) / d2;
		if (jf) { dx -= rx * f; dy -= ry * f; dz -= rz * f; }
		else { dx -= 2.0 * rx * f; dy -= 2.0 * ry * f; dz -= 2.0 * rz * f; }
	}
}
template <bool PER>   // PER: periodic domain
__global__ void __launch_bounds__(256) k_collide(const __grid_constant__ DevView S, const __grid_constant__ KParams K) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= S.n0) return;
	const double4 vi = ld256(S.vel + i);
	S.vel2[i] = vi;   // every slot (ghost rows too): the buffers are swapped after the pass
	if (i >= S.n) return;
	const double4 pi = ld256(S.pos + i);
	if (pk_type(pi) != K.p.fluid) return;
	const double mi = pk_ptype(pi) == 1 ? K.p.DNS_FL1 : K.p.DNS_FL2;
	double dx = 0.0, dy = 0.0, dz = 0.0;
	const int cnt = S.ccnt[i];
	if (cnt != 255) {
		const uint32_t* L = S.clist + (size_t)i * PMPS_CLOSE_CAP;
		for (int k = 0; k < cnt; k++) collide_pair<false, PER>(S, K, i, pi, vi, mi, (int)L[k], dx, dy, dz);
	} else {
		// more close partners than a list row holds (a badly compressed spot): the 27 (9) buckets around the particle's own, as at the
		// last rebuild — complete while nobody has moved beyond the search slack
		const int c = S.cellOf[i];
		if (c < S.nC) {
			int cx, cy, cz; cell_coords(K, c, cx, cy, cz);
			const bool three = K.p.dim == 3;
			for (int jz = three ? cz - 1 : 0; jz <= (three ? cz + 1 : 0); jz++) for (int jy = cy - 1; jy <= cy + 1; jy++) for (int jx = cx - 1; jx <= cx + 1; jx++) {
				if (jx < 0 || jx >= K.p.numBucketsX || jy < 0 || jy >= K.p.numBucketsY || jz < 0 || jz >= K.p.numBucketsZ) continue;
				int ax = jx, ay = jy, az = jz; if (PER) per_alias(K, ax, ay, az);
				const int cj = cell_index(K, ax, ay, az);
				for (int j = S.cellStart[cj]; j < S.cellStart[cj + 1]; j++) if (j != i) collide_pair<true, PER>(S, K, i, pi, vi, mi, j, dx, dy, dz);
			}
		}
	}
	S.dvelCollision.x[i] = dx; S.dvelCollision.y[i] = dy; S.dvelCollision.z[i] = dz;
	double4 v = vi; v.x = __dadd_rn(v.x, dx); v.y = __dadd_rn(v.y, dy); v.z = __dadd_rn(v.z, dz);   // second loop (:108-116)
	S.vel2[i] = v;
}

// both loops of checkParticleCollisions in one launch; the caller swaps the velocity buffers
int pmps_collide_lists(pmps_handle* h) {
	if (!h->hitValid) { const int rc = pmps_tv_hits(h); if (rc) return rc; }   // the close-pair lists live and die with the hit list
	if (h->K.p.domainTypeBC == 1) LAUNCH(h, "F2.collision", k_collide<true>, cdiv(h->V.n0, 256), 256, h->V, h->K);
	else LAUNCH(h, "F2.collision", k_collide<false>, cdiv(h->V.n0, 256), 256, h->V, h->K);
	CUDA_TRY(h, cudaGetLastError());
	return PMPS_OK;
}
