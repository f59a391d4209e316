// pmps_wall.cuh — device side of W1 (polygon-wall distance, mirror particle, wall-weight look-up); see pmps_wall.cu
// for the host side (bounding-box hierarchy in libigl's order) and the rationale for the exact arithmetic.
#pragma once
#include "pmps_device.cuh"

// ------------------------------------------------------------------------------------------------ device: query
// Closest point on a triangle; region tests in the order of Ericson ch. 5 as libigl applies them.
static __device__ __noinline__ void closest_on_triangle(const double* p, const double* a, const double* b, const double* c, double* q) {
	double ab[3], ac[3], ap[3], bp[3], cp[3];
	for (int k = 0; k < 3; k++) { ab[k] = __dsub_rn(b[k], a[k]); ac[k] = __dsub_rn(c[k], a[k]); ap[k] = __dsub_rn(p[k], a[k]); }
	const double d1 = dot3_exact(ab[0], ab[1], ab[2], ap[0], ap[1], ap[2]), d2 = dot3_exact(ac[0], ac[1], ac[2], ap[0], ap[1], ap[2]);
	if (d1 <= 0.0 && d2 <= 0.0) { q[0] = a[0]; q[1] = a[1]; q[2] = a[2]; return; }
	for (int k = 0; k < 3; k++) bp[k] = __dsub_rn(p[k], b[k]);
	const double d3 = dot3_exact(ab[0], ab[1], ab[2], bp[0], bp[1], bp[2]), d4 = dot3_exact(ac[0], ac[1], ac[2], bp[0], bp[1], bp[2]);
	if (d3 >= 0.0 && d4 <= d3) { q[0] = b[0]; q[1] = b[1]; q[2] = b[2]; return; }
	const double vc = __dsub_rn(__dmul_rn(d1, d4), __dmul_rn(d3, d2));
	if (a[0] != b[0] || a[1] != b[1] || a[2] != b[2]) {
		if (vc <= 0.0 && d1 >= 0.0 && d3 <= 0.0) {
			const double v = __ddiv_rn(d1, __dsub_rn(d1, d3));
			for (int k = 0; k < 3; k++) q[k] = __dadd_rn(a[k], __dmul_rn(v, ab[k]));
			return;
		}
	}
	for (int k = 0; k < 3; k++) cp[k] = __dsub_rn(p[k], c[k]);
	const double d5 = dot3_exact(ab[0], ab[1], ab[2], cp[0], cp[1], cp[2]), d6 = dot3_exact(ac[0], ac[1], ac[2], cp[0], cp[1], cp[2]);
	if (d6 >= 0.0 && d5 <= d6) { q[0] = c[0]; q[1] = c[1]; q[2] = c[2]; return; }
	const double vb = __dsub_rn(__dmul_rn(d5, d2), __dmul_rn(d1, d6));
	if (vb <= 0.0 && d2 >= 0.0 && d6 <= 0.0) {
		const double w = __ddiv_rn(d2, __dsub_rn(d2, d6));
		for (int k = 0; k < 3; k++) q[k] = __dadd_rn(a[k], __dmul_rn(w, ac[k]));
		return;
	}
	const double va = __dsub_rn(__dmul_rn(d3, d6), __dmul_rn(d5, d4));
	if (va <= 0.0 && __dsub_rn(d4, d3) >= 0.0 && __dsub_rn(d5, d6) >= 0.0) {
		const double w = __ddiv_rn(__dsub_rn(d4, d3), __dadd_rn(__dsub_rn(d4, d3), __dsub_rn(d5, d6)));
		for (int k = 0; k < 3; k++) q[k] = __dadd_rn(b[k], __dmul_rn(w, __dsub_rn(c[k], b[k])));
		return;
	}
	const double denom = __ddiv_rn(1.0, __dadd_rn(__dadd_rn(va, vb), vc)); const double v = __dmul_rn(vb, denom), w = __dmul_rn(vc, denom);
	for (int k = 0; k < 3; k++) q[k] = __dadd_rn(__dadd_rn(a[k], __dmul_rn(ab[k], v)), __dmul_rn(ac[k], w));
}

__device__ __forceinline__ double box_ext2(const double* lo, const double* hi, const double* p) {   // Eigen AlignedBox::squaredExteriorDistance
	double d = 0.0;
	for (int k = 0; k < 3; k++) {
		if (lo[k] > p[k]) { const double a = __dsub_rn(lo[k], p[k]); d = __dadd_rn(d, __dmul_rn(a, a)); }
		else if (p[k] > hi[k]) { const double a = __dsub_rn(p[k], hi[k]); d = __dadd_rn(d, __dmul_rn(a, a)); }
	}
	return d;
}
__device__ __forceinline__ bool box_has(const double* lo, const double* hi, const double* p) {
	return lo[0] <= p[0] && p[0] <= hi[0] && lo[1] <= p[1] && p[1] <= hi[1] && lo[2] <= p[2] && p[2] <= hi[2];
}

// Depth-first descent in libigl's visiting order, without recursion: every stack entry is a node still to be looked at
// together with the rule under which it may be entered ("always" = its box contains p; otherwise only if its box is
// closer than the best distance found SO FAR, evaluated when the entry is popped — the same moment the recursive
// original evaluates it).
#define TREE_STACK 64
static __device__ __noinline__ void tree_query(const DevMesh& m, const double* p, double* sqOut, int* faceOut, double* cOut) {
	double best = INFINITY; int face = -1; double c[3] = { 0.0, 0.0, 0.0 };
	int stackNode[TREE_STACK]; double stackBound[TREE_STACK]; int sp = 0;   // bound < 0: enter unconditionally
	stackNode[sp] = 0; stackBound[sp] = -1.0; sp++;
	while (sp > 0) {
		sp--; const int t = stackNode[sp]; const double bound = stackBound[sp];
		if (bound >= 0.0 && !(bound < best)) continue;
		const int pr = m.prim[t];
		if (pr >= 0) {
			double q[3];
			closest_on_triangle(p, m.V + 3 * m.F[3 * pr], m.V + 3 * m.F[3 * pr + 1], m.V + 3 * m.F[3 * pr + 2], q);
			const double dd = sq3_exact(__dsub_rn(p[0], q[0]), __dsub_rn(p[1], q[1]), __dsub_rn(p[2], q[2]));
			if (dd < best) { best = dd; face = pr; c[0] = q[0]; c[1] = q[1]; c[2] = q[2]; }
			continue;
		}
		const int L = 2 * t + 1, R = 2 * t + 2;
		const bool inL = box_has(m.bbMin + 3 * L, m.bbMax + 3 * L, p), inR = box_has(m.bbMin + 3 * R, m.bbMax + 3 * R, p);
		const double dl = box_ext2(m.bbMin + 3 * L, m.bbMax + 3 * L, p), dr = box_ext2(m.bbMin + 3 * R, m.bbMax + 3 * R, p);
		// visiting order of the original: [L if inL] [R if inR] then the remaining ones, nearer box first, each only if closer than best
		int ord[2]; double bnd[2]; int n = 0;
		if (inL) { ord[n] = L; bnd[n] = -1.0; n++; }
		if (inR) { ord[n] = R; bnd[n] = -1.0; n++; }
		if (dl < dr) { if (!inL) { ord[n] = L; bnd[n] = dl; n++; } if (!inR) { ord[n] = R; bnd[n] = dr; n++; } }
		else { if (!inR) { ord[n] = R; bnd[n] = dr; n++; } if (!inL) { ord[n] = L; bnd[n] = dl; n++; } }
		for (int k = n - 1; k >= 0; k--) if (sp < TREE_STACK) { stackNode[sp] = ord[k]; stackBound[sp] = bnd[k]; sp++; }
	}
	*sqOut = best; *faceOut = face; cOut[0] = c[0]; cOut[1] = c[1]; cOut[2] = c[2];
}

// table lerp with extrapolation, src/PolygonMesh.cpp:533-581
__device__ __forceinline__ double table_lerp(const double* xs, const double* ys, const int32_t* yi, int n, double x) {
	int k = 0;
	if (x >= xs[n - 2]) k = n - 2; else while (x > xs[k + 1]) k++;
	const double xL = xs[k], xR = xs[k + 1]; const double yL = ys ? ys[k] : (double)yi[k], yR = ys ? ys[k + 1] : (double)yi[k + 1];
	const double dydx = __ddiv_rn(__dsub_rn(yR, yL), __dsub_rn(xR, xL));
	return __dadd_rn(yL, __dmul_rn(dydx, __dsub_rn(x, xL)));
}


// closestPointPNDBoundaryAABB for every mesh in slot order (rigid, deformable, forced; src/PolygonMesh.cpp:585-651), then
// updateParticlesNearPolygonMesh with the tables of mesh 0 (:718-795), for ONE particle
__device__ __forceinline__ void wall_closest_of(const DevView& S, const KParams& K, int i, const double4& pi) {
	const double p[3] = { pi.x, pi.y, pi.z };
	double dist2 = S.distParticleWall2[i];
	for (int me = 0; me < S.nMeshes; me++) {
		double sq, c[3]; int f;
		tree_query(S.meshes[me], p, &sq, &f, c);
		if (sq < dist2) {
			dist2 = sq;
			S.particleAtWallPos.x[i] = c[0]; S.particleAtWallPos.y[i] = c[1]; S.particleAtWallPos.z[i] = c[2];
			S.mirrorParticlePos.x[i] = __dadd_rn(p[0], __dmul_rn(2.0, __dsub_rn(c[0], p[0])));
			S.mirrorParticlePos.y[i] = __dadd_rn(p[1], __dmul_rn(2.0, __dsub_rn(c[1], p[1])));
			S.mirrorParticlePos.z[i] = __dadd_rn(p[2], __dmul_rn(2.0, __dsub_rn(c[2], p[2])));
			const double* Nn = S.meshes[me].N + 3 * f;
			S.polygonNormal.x[i] = Nn[0]; S.polygonNormal.y[i] = Nn[1]; S.polygonNormal.z[i] = Nn[2];
			S.nearMeshType[i] = me;
			if (me == 1) S.elementID[i] = f;
		}
	}
	S.distParticleWall2[i] = dist2;
	double pw = 0.0; int nw = 0;
	if (dist2 < K.p.reL2) {
		S.particleNearWall[i] = 1;
		const DevMesh& m0 = S.meshes[0];
		const double x = sqrt(dist2), reS = sqrt(K.p.reS2), reL = sqrt(K.p.reL2);
		pw = table_lerp(m0.xDataPND, m0.Zij, nullptr, m0.nTable, __ddiv_rn(x, reS));
		nw = (int)table_lerp(m0.xDataNeigh, nullptr, m0.nNeighWall, m0.nTable, __ddiv_rn(x, reL));
	}
	S.pndWallContribution[i] = pw; S.numNeighWallContribution[i] = nw;
}
