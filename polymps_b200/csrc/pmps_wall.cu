// pmps_wall.cu — W1: polygon (STL triangle) wall distance, mirror particle and wall-weight look-up.
// Replaces PolygonMesh::closestPointPNDBoundaryAABB + updateParticlesNearPolygonMesh (reference
// src/PolygonMesh.cpp:585-795) and the libigl calls under them (AABB<...,3>::squared_distance,
// libigl/include/igl/AABB.cpp:363-450,806-830; point_simplex_squared_distance.cpp:44-110).
//
// Everything in this file is predicate-critical: which triangle wins a tie between two equidistant walls decides the
// mirror position of a corner particle, and (int)lerp(...) decides an integer neighbour count.  So all arithmetic here
// is written with non-contracting intrinsics in the reference's evaluation order, and the bounding-box hierarchy is
// built on the host in libigl's order (median split on the rank of the barycentre along the longest box axis) so that
// the traversal visits equidistant triangles in the same order as the reference.
#include "pmps_wall.cuh"
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>

// ------------------------------------------------------------------------------------------------ host: hierarchy
namespace {
struct TreeBuilder {
	int nF; const double* V; const int32_t* F;
	std::vector<double> bbMin, bbMax; std::vector<int32_t> prim; std::vector<int> rankOf;   // rankOf[3*f+axis]
	static int nodes_for(int nF) { int d = 0; while ((1 << d) < nF) d++; return (1 << (d + 1)) - 1 + 2; }
	void build(int nF_, const double* V_, const int32_t* F_) {
		nF = nF_; V = V_; F = F_;
		const int nn = nodes_for(nF);
		bbMin.assign(3 * (size_t)nn, 0.0); bbMax.assign(3 * (size_t)nn, 0.0); prim.assign(nn, -2); rankOf.assign(3 * (size_t)nF, 0);
		std::vector<double> bc(nF); std::vector<int> idx(nF);
		for (int ax = 0; ax < 3; ax++) {
			for (int f = 0; f < nF; f++) {   // barycentre ((v0+v1)+v2)/3 (igl/barycenter.cpp:19-31)
				double s = 0.0; for (int c = 0; c < 3; c++) s += V[3 * F[3 * f + c] + ax];
				bc[f] = s / 3.0; idx[f] = f;
			}
			std::stable_sort(idx.begin(), idx.end(), [&](int a, int b) { return bc[a] < bc[b]; });
			for (int r = 0; r < nF; r++) rankOf[3 * idx[r] + ax] = r;
		}
		std::vector<int> all(nF); for (int f = 0; f < nF; f++) all[f] = f;
		node(0, all);
	}
	void node(int t, const std::vector<int>& I) {
		double lo[3] = { DBL_MAX, DBL_MAX, DBL_MAX }, hi[3] = { -DBL_MAX, -DBL_MAX, -DBL_MAX };
		for (int f : I) for (int c = 0; c < 3; c++) for (int k = 0; k < 3; k++) {
			const double v = V[3 * F[3 * f + c] + k]; lo[k] = std::min(lo[k], v); hi[k] = std::max(hi[k], v);
		}
		for (int k = 0; k < 3; k++) { bbMin[3 * t + k] = lo[k]; bbMax[3 * t + k] = hi[k]; }
		if (I.size() == 1) { prim[t] = I[0]; return; }
		prim[t] = -1;
		int ax = 0; for (int k = 1; k < 3; k++) if (hi[k] - lo[k] > hi[ax] - lo[ax]) ax = k;   // first maximum
		std::vector<int> ranks; ranks.reserve(I.size());
		for (int f : I) ranks.push_back(rankOf[3 * f + ax]);
		std::vector<int> sorted = ranks; std::sort(sorted.begin(), sorted.end());
		const int med = sorted[(I.size() - 1) / 2];
		std::vector<int> L, R;
		for (size_t e = 0; e < I.size(); e++) (ranks[e] <= med ? L : R).push_back(I[e]);
		if (!L.empty()) node(2 * t + 1, L);
		if (!R.empty()) node(2 * t + 2, R);
	}
};
}  // namespace

int pmps_mesh_upload(pmps_handle* h, int mesh_id, int nV, const double* V, int nF, const int32_t* F, const double* N,
                     int nTable, const double* xP, const double* xN, const double* Z, const int32_t* nW) {
	if (mesh_id < 0 || mesh_id >= 3 || nF < 1 || nV < 3 || nTable < 2) return pmps_fail(h, PMPS_E_ARG, "pmps_set_mesh: bad sizes");
	if ((int)h->meshes.size() <= mesh_id) h->meshes.resize(mesh_id + 1, DevMesh{});
	TreeBuilder tb; tb.build(nF, V, F);
	DevMesh& m = h->meshes[mesh_id];
	auto put = [&](const void* src, size_t bytes, void** dst) -> int {
		if (*dst) cudaFree(*dst);
		cudaError_t e = cudaMalloc(dst, bytes); if (e != cudaSuccess) return -1;
		e = cudaMemcpy(*dst, src, bytes, cudaMemcpyHostToDevice); return e == cudaSuccess ? 0 : -1;
	};
	int bad = 0;
	m.nV = nV; m.nF = nF; m.nNodes = (int)tb.prim.size(); m.nTable = nTable;
	bad |= put(V, sizeof(double) * 3 * nV, (void**)&m.V); bad |= put(F, sizeof(int32_t) * 3 * nF, (void**)&m.F);
	bad |= put(N, sizeof(double) * 3 * nF, (void**)&m.N);
	bad |= put(tb.bbMin.data(), sizeof(double) * tb.bbMin.size(), (void**)&m.bbMin);
	bad |= put(tb.bbMax.data(), sizeof(double) * tb.bbMax.size(), (void**)&m.bbMax);
	bad |= put(tb.prim.data(), sizeof(int32_t) * tb.prim.size(), (void**)&m.prim);
	bad |= put(xP, sizeof(double) * nTable, (void**)&m.xDataPND); bad |= put(xN, sizeof(double) * nTable, (void**)&m.xDataNeigh);
	bad |= put(Z, sizeof(double) * nTable, (void**)&m.Zij); bad |= put(nW, sizeof(int32_t) * nTable, (void**)&m.nNeighWall);
	if (bad) return pmps_fail(h, PMPS_E_CUDA, "pmps_set_mesh: device allocation/copy failed");
	if (!h->d_meshes) { if (cudaMalloc((void**)&h->d_meshes, sizeof(DevMesh) * 3) != cudaSuccess) return pmps_fail(h, PMPS_E_CUDA, "pmps_set_mesh: cudaMalloc"); }
	if (cudaMemcpy(h->d_meshes, h->meshes.data(), sizeof(DevMesh) * h->meshes.size(), cudaMemcpyHostToDevice) != cudaSuccess)
		return pmps_fail(h, PMPS_E_CUDA, "pmps_set_mesh: cudaMemcpy");
	h->V.meshes = h->d_meshes; h->V.nMeshes = h->K.p.numOfMeshs;
	return PMPS_OK;
}

__global__ void __launch_bounds__(128) k_wall_closest(const __grid_constant__ DevView S, const __grid_constant__ KParams K) {
	const int i = blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= S.n) return;
	const double4 pi = S.pos[i];
	if (pk_type(pi) == K.p.ghost) return;
	wall_closest_of(S, K, i, pi);
	if (S.particleNearWall[i] && S.cellOf[i] < S.nC) S.tileWall[S.cellOf[i] / K.TV] = 1;   // wall-only passes skip the other tiles
}

extern "C" int pmps_wall_closest_point(pmps_t* h) {
	if (!h) return PMPS_E_ARG;
	cudaSetDevice(h->device);
	if (!h->haveState) return pmps_fail(h, PMPS_E_STATE, "pmps_wall_closest_point: no particle state uploaded");
	if ((int)h->meshes.size() < h->K.p.numOfMeshs || h->K.p.numOfMeshs < 1) return pmps_fail(h, PMPS_E_STATE, "pmps_wall_closest_point: %d meshes expected, %d set", h->K.p.numOfMeshs, (int)h->meshes.size());
	if (h->haveList) CUDA_TRY(h, cudaMemsetAsync(h->V.tileWall, 0, (size_t)h->nTilesAll, h->stream));
	LAUNCH(h, "wall.closest_point", k_wall_closest, cdiv(h->V.n, 128), 128, h->V, h->K);
	h->tileWallValid = h->haveList;
	CUDA_TRY(h, cudaGetLastError());
	return PMPS_OK;
}
