// pmps_shim.cpp — the reference-side binding of libpmps: the per-timestep methods of the reference's engine classes
// (called from the UNMODIFIED src/main.cpp:98-456) re-implemented as forwards to the C ABI (include/pmps.h).
//
// Built only where the reference tree is present (polymps_b200/host/Makefile): this file is compiled against the reference's
// own headers and linked with the unmodified reference objects, in whose copies every symbol defined HERE has been made weak
// (objcopy --weaken-symbols), so each hot-path method has exactly one strong definition — this one — and the reference's code
// serves everything else (input parsing, setParameters, bucket geometry, STL reading, wall tables, VTU output).  Nothing of the
// reference is copied or edited.  EVERY engine method main.cpp:98-456 can reach is defined here: forwarded to the C ABI, or —
// where the device path does not have the feature yet — a loud exit(10); tests/test_shim_link.py checks that no per-particle
// method main.o references can fall through to the reference's CPU loop (which would run on stale host arrays).
//
// State policy: the particle arrays live on the device between calls.  The host arrays of MpsParticle are refreshed
//   * before every output iteration (writeOutputFiles reads them, src/main.cpp:234-236): pmps_download of every field;
//   * every step: numParticles, velMax, CFLcurrent (displayInfo, src/MpsInputOutput.cpp:37-55).
// Error behaviour: the reference's methods are void and exit(10) on fatal conditions (src/MpsBucket.cpp:104): same here.
#include <limits>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "MpsParticleSystem.h"
#include "MpsParticle.h"
#include "MpsBucket.h"
#include "MpsPndNeigh.h"
#include "MpsViscosity.h"
#include "MpsPressure.h"
#include "MpsParticleCollision.h"
#include "MpsShifting.h"
#include "MpsVectorMatrix.h"
#include "MpsParticleVelPos.h"
#include "MpsBoundaryCondition.h"
#include "PolygonMesh.h"
#include "MpsInputOutput.h"
#include "pmps.h"

namespace {
pmps_t* g_h = nullptr;
MpsParticleSystem* g_S = nullptr;
MpsParticle* g_P = nullptr;
int g_oodCalls = 0;            // checkParticleOutDomain is called three times per step (src/main.cpp:283,411,441)
bool g_meshSet[3] = { false, false, false };

void die(const char* where, int rc) {
	fprintf(stderr, "libpmps: %s failed (%d): %s\n", where, rc, pmps_last_error(g_h));
	exit(10);
}
#define CALL(fn) do { const int rc_ = (fn); if (rc_ < 0) die(#fn, rc_); } while (0)

PmpsHostArrays host_arrays(MpsParticle* P) {
	PmpsHostArrays a; memset(&a, 0, sizeof a);
	a.numParticles = P->numParticles; a.numParticlesZero = P->numParticlesZero;
#define F(n) a.n = P->n;
	F(particleType) F(particleBC) F(numNeigh) F(PTYPE) F(nearMeshType) F(numNeighWallContribution) F(elementID)
	F(press) F(pndi) F(pndki) F(pndski) F(pndSmall) F(npcdDeviation2) F(concentration) F(velDivergence) F(diffusiveTerm)
	F(pndWallContribution) F(distParticleWall2) F(deviationDotPolygonNormal) F(RHO) F(MEU) F(pndAuxVar1) F(pndAuxVar2)
	F(pos) F(vel) F(acc) F(accStar) F(npcdDeviation) F(gradConcentration) F(correcMatrixRow1) F(correcMatrixRow2) F(correcMatrixRow3)
	F(particleAtWallPos) F(mirrorParticlePos) F(polygonNormal) F(forceWall) F(dvelCollision) F(normal)
#undef F
	if (g_S && g_S->fluidType == viscType::NON_NEWTONIAN) {   // allocated only then (src/MpsInputOutput.cpp:729-744)
#define F(n) a.n = P->n;
		F(Cv) F(II) F(MEU_Y) F(Inertia) F(p_rheo_new) F(p_smooth) F(VF) F(S11) F(S12) F(S13) F(S22) F(S23) F(S33)
#undef F
	}
	a.particleNearWall = reinterpret_cast<uint8_t*>(P->particleNearWall);   // C++ bool, one byte
	if (P->pressurePPE.size() == P->numParticlesZero) { a.pressurePPE = P->pressurePPE.data(); a.sourceTerm = P->sourceTerm.data(); }
	return a;
}

// POD snapshot of MpsParticleSystem after setParameters() / MpsBucket::allocateMemory() — same member names on both sides
PmpsParams params_of(MpsParticleSystem* S) {
	PmpsParams p; memset(&p, 0, sizeof p);
	p.abi_version = PMPS_ABI_VERSION;
	p.shiftingJacobi = 1;   // parallel-safe shifting: the reference's own loop is a race for more than one thread
#define C(n) p.n = (decltype(p.n))S->n;
	C(dim) C(wallType) C(mpsType) C(solverType) C(pndType) C(weightType) C(gradientType) C(gradientCorrection)
	C(slipCondition) C(freeSurfType) C(collisionType) C(shiftingType) C(repulsiveForceType) C(fluidType) C(domainTypeBC)
	C(ghost) C(fluid) C(wall) C(dummyWall) C(surface) C(inner) C(other) C(numBucketsX) C(numBucketsY) C(numBucketsZ)
	C(numOfMeshs) C(domainMinX) C(domainMinY) C(domainMinZ) C(domainMaxX) C(domainMaxY) C(domainMaxZ) C(bucketSide)
	C(invBucketSide) C(partDist) C(timeStep) C(epsilonZero) C(reS) C(reL) C(reS2) C(reL2) C(gravityX) C(gravityY)
	C(gravityZ) C(densityWall) C(DNS_FL1) C(DNS_FL2) C(soundSpeed) C(gamma) C(relaxPress) C(alphaCompressibility)
	C(reRepulsiveForce) C(expectMaxVelocity) C(repForceCoefMitsume) C(repForceCoefLennardJones)
	C(repForceCoefMonaghanKajtar) C(pndSmallZero) C(pndLargeZero) C(pndGradientZero) C(lambdaZero) C(numNeighZero)
	C(coeffViscMultiphase) C(coeffPressGrad) C(coeffPressEMPS) C(coeffPressWCMPS) C(coeffShifting1) C(coeffShifting2)
	C(coeffPPE) C(coeffPPESource) C(distCollisionLimit2) C(restitutionCollision) C(lambdaCollision) C(betaPnd)
	C(betaNeigh) C(delta2) C(velMax)
	C(KNM_VS1) C(KNM_VS2) C(DNS_SDT) C(PHI_1) C(PHI_2) C(cohes) C(MEU0) C(DG) C(I0) C(mm) C(Cd) C(VF_min) C(VF_max) C(relaxPND) C(timeCurrent)
	C(thetaArc) C(hThreshold2) C(dstThreshold2) C(normThreshold2)
#undef C
	p.fractionMethod = S->Fraction_method; p.powerLawN = S->N;
	for (int k = 0; k < 3; k++) p.uniformVelWall[k] = S->uniformVelWall[k];
	if (S->domainTypeBC == 1) {   // periodic domain: the one direction calcDomainLimits handled (src/MpsBucket.cpp:142-289)
		const int a = S->periodicDirectionX ? 0 : S->periodicDirectionY ? 1 : 2;
		p.periodicAxis = a; p.periodicLength = S->periodicLength[a];
		p.physDomMin = a == 0 ? S->physDomMinX : a == 1 ? S->physDomMinY : S->physDomMinZ;
		p.physDomMax = a == 0 ? S->physDomMaxX : a == 1 ? S->physDomMaxY : S->physDomMaxZ;
	}
	return p;
}

// first hot-path call: create the handle from the (by now complete) parameter set and send the particle arrays
pmps_t* handle(MpsParticleSystem* S, MpsParticle* P) {
	if (g_h) return g_h;
	g_S = S; g_P = P;
	const PmpsParams p = params_of(S);
	int dev = 0; if (const char* e = getenv("PMPS_DEVICE")) dev = atoi(e);
	const int rc = pmps_create(&p, dev, &g_h);
	if (rc < 0) { fprintf(stderr, "libpmps: pmps_create failed (%d): %s\n", rc, pmps_last_error(nullptr)); exit(10); }
	const PmpsHostArrays a = host_arrays(P);
	// the reference mallocs its particle arrays once and never frees them (src/MpsInputOutput.cpp:667-750): safe to page-lock for
	// the lifetime of the handle (contract of pmps_register_host); failure to pin is not fatal, the copies then go the pageable way
	{
		const size_t n0 = (size_t)P->numParticlesZero;
		void* v3[] = { P->pos, P->vel, P->acc, P->accStar, P->mirrorParticlePos, P->particleAtWallPos, P->polygonNormal };
		void* f8[] = { P->press, P->pndi, P->pndSmall, P->concentration, P->RHO, P->MEU };
		for (void* q : v3) if (q) pmps_register_host(g_h, q, 24 * n0);
		for (void* q : f8) if (q) pmps_register_host(g_h, q, 8 * n0);
	}
	CALL(pmps_upload(g_h, &a));
	return g_h;
}
pmps_t* handle() {
	if (!g_h) { fprintf(stderr, "libpmps: a mesh / particle method was called before MpsBucket::updateParticlesID\n"); exit(10); }
	return g_h;
}
void pull_scalars() {   // what displayInfo and the host loop read
	PmpsScalars s; CALL(pmps_get_scalars(g_h, &s));
	g_P->numParticles = s.numParticles; g_S->velMax = s.velMax; g_S->CFLcurrent = s.CFLcurrent;
	if (g_S->fluidType == viscType::NON_NEWTONIAN) g_S->CFLvisc = s.CFLvisc;
}
// Before writeOutputFiles: only what the reference's writers read comes back from the device — the union over the output flags
// (private members of MpsInputOutput, not visible from here) of the arrays writeVtuAscii / writeVtuBinary / writeVtuAsciiFreeSurface
// touch (src/MpsInputOutput.cpp:861-1875): positions, velocities, pressure, boundary flags, both number densities, neighbour count,
// NPCD deviation, polygon normal, concentration and its gradient, particle / mesh types, and the rheology fields of a non-Newtonian
// run: 168 of the 585 bytes per particle the full state has.  PMPS_SHIM_PULL_ALL=1 brings everything back (debugging).
void pull_arrays() {
	PmpsHostArrays all = host_arrays(g_P), a;
	if (getenv("PMPS_SHIM_PULL_ALL")) a = all;
	else {
		memset(&a, 0, sizeof a); a.numParticles = all.numParticles; a.numParticlesZero = all.numParticlesZero;
#define F(n) a.n = all.n;
		F(pos) F(vel) F(press) F(particleBC) F(pndi) F(pndSmall) F(numNeigh) F(npcdDeviation) F(polygonNormal) F(concentration) F(gradConcentration)
		F(particleType) F(nearMeshType) F(PTYPE)
		if (g_S->fluidType == viscType::NON_NEWTONIAN) { F(RHO) F(Cv) F(MEU) F(MEU_Y) F(II) F(p_smooth) F(Inertia) F(p_rheo_new) F(VF) }
#undef F
	}
	CALL(pmps_download(g_h, &a));
	g_P->numParticles = a.numParticles;
}
// the iteration about to start writes output (src/main.cpp:234): numOfIterations is incremented right after the call we are in
bool output_next(int inc) { return g_S->iterOutput > 0 && (g_S->numOfIterations + inc) % g_S->iterOutput == 0; }

// row-major copies of an Eigen mesh (column-major) for pmps_set_mesh; the wall tables are private members, passed in by the caller
void unsupported(const char* method, const char* where) {
	fprintf(stderr, "libpmps: %s (%s) is not on the device path of this build; refusing to run the reference's CPU loop on stale host arrays\n", method, where);
	exit(10);
}
void send_mesh(int me, const Eigen::MatrixXd& NV, const Eigen::MatrixXi& NF, const Eigen::MatrixXd& NN, const std::vector<double>& xP,
               const std::vector<double>& xN, const std::vector<double>& Z, const std::vector<int>& nNW) {
	const int nV = (int)NV.rows(), nF = (int)NF.rows();
	std::vector<double> V(3 * (size_t)nV), N(3 * (size_t)nF); std::vector<int32_t> F(3 * (size_t)nF);
	for (int i = 0; i < nV; i++) for (int k = 0; k < 3; k++) V[3 * i + k] = NV(i, k);
	for (int i = 0; i < nF; i++) for (int k = 0; k < 3; k++) { F[3 * i + k] = NF(i, k); N[3 * i + k] = NN(i, k); }
	std::vector<int32_t> nw(nNW.begin(), nNW.end());
	CALL(pmps_set_mesh(handle(), me, nV, V.data(), nF, F.data(), N.data(), (int)xP.size(), xP.data(), xN.data(), Z.data(), nw.data()));
	g_meshSet[me] = true;
}
}   // namespace

#define H(S, P) handle((S), (P))
// ---- MpsBucket (src/MpsBucket.cpp:299-353) -------------------------------------------------------------------------
void MpsBucket::updateParticlesID(MpsParticleSystem *PSystem, MpsParticle *Particles) { CALL(pmps_update_buckets(H(PSystem, Particles))); }

// ---- MpsParticle (src/MpsParticle.cpp:111-130, 234-402) --------------------------------------------------------------
void MpsParticle::addGravity(MpsParticleSystem *PSystem, MpsParticle *Particles) { CALL(pmps_add_gravity(H(PSystem, Particles))); }
void MpsParticle::checkParticleOutDomain(MpsParticleSystem *PSystem) {
	CALL(pmps_check_out_domain(H(PSystem, this)));
	// calls per time step: after the prediction, after the correction, at the end (src/main.cpp:283,411,441); the call in main()
	// before the loop is number 0.  The last one of a step is where the host's view is refreshed.
	if (g_oodCalls > 0 && g_oodCalls % 3 == 0) { pull_scalars(); if (output_next(1)) pull_arrays(); }
	g_oodCalls++;
}

// ---- MpsViscosity (src/MpsViscosity.cpp:20-100, 1527-1910) -----------------------------------------------------------
void MpsViscosity::calcViscosity(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) { CALL(pmps_calc_viscosity(H(PSystem, Particles))); }
// non-Newtonian rheology (src/MpsViscosity.cpp:104-190, :194-700)
void MpsViscosity::calcVolumeFraction(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) { CALL(pmps_calc_volume_fraction(H(PSystem, Particles))); }
void MpsViscosity::calcViscosityInteractionVal(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) {
	CALL(pmps_set_time(H(PSystem, Particles), PSystem->timeCurrent));
	CALL(pmps_calc_viscosity_interaction(H(PSystem, Particles)));
}
// polygon-wall variants of the rheology (src/MpsViscosity.cpp:703-1526; src/main.cpp:391-399): one stage call, the slip condition is in the parameters
void MpsViscosity::calcWallSlipViscosityInteractionVal(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) {
	CALL(pmps_set_time(H(PSystem, Particles), PSystem->timeCurrent)); CALL(pmps_calc_wall_viscosity_interaction(H(PSystem, Particles)));
}
void MpsViscosity::calcWallNoSlipViscosityInteractionVal(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) {
	CALL(pmps_set_time(H(PSystem, Particles), PSystem->timeCurrent)); CALL(pmps_calc_wall_viscosity_interaction(H(PSystem, Particles)));
}
void MpsViscosity::calcWallSlipViscosity(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) { CALL(pmps_wall_viscosity(H(PSystem, Particles))); }
void MpsViscosity::calcWallNoSlipViscosity(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) { CALL(pmps_wall_viscosity(H(PSystem, Particles))); }

// ---- MpsParticleVelPos (src/MpsParticleVelPos.cpp:20-158) ------------------------------------------------------------
void MpsParticleVelPos::predictionVelocityPosition(MpsParticleSystem *PSystem, MpsParticle *Particles) { CALL(pmps_predict(H(PSystem, Particles))); }
void MpsParticleVelPos::correctionVelocityPosition(MpsParticleSystem *PSystem, MpsParticle *Particles) { CALL(pmps_correct(H(PSystem, Particles))); }
void MpsParticleVelPos::updateVelocityParticlesWallDummy(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) { CALL(pmps_update_wall_dummy_vel(H(PSystem, Particles))); }

// ---- MpsParticleCollision (src/MpsParticleCollision.cpp:22-119, dynamic model :124-292) -------------------------------
// pmps_collisions runs the model PmpsParams.collisionType names; main.cpp picks the method by the same flag (src/main.cpp:285-290)
void MpsParticleCollision::checkParticleCollisions(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) { CALL(pmps_collisions(H(PSystem, Particles))); }
void MpsParticleCollision::checkDynamicParticleCollisions(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) { CALL(pmps_collisions(H(PSystem, Particles))); }

// ---- MpsPndNeigh (src/MpsPndNeigh.cpp:21-1049) ------------------------------------------------------------------------
void MpsPndNeigh::setInitialPndNumberOfNeighNPCD(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) { CALL(pmps_initial_pnd(H(PSystem, Particles))); }
void MpsPndNeigh::calcWallNPCD(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) { CALL(pmps_wall_npcd(H(PSystem, Particles))); }
void MpsPndNeigh::calcPndnNeighNPCD(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) { CALL(pmps_pnd_neigh_npcd(H(PSystem, Particles))); }
void MpsPndNeigh::calcPndDiffusiveTerm(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) { CALL(pmps_pnd_diffusive(H(PSystem, Particles))); }
void MpsPndNeigh::calcWallSlipPndDiffusiveTerm(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) { CALL(pmps_wall_pnd_diffusive(H(PSystem, Particles))); }
void MpsPndNeigh::calcWallNoSlipPndDiffusiveTerm(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) { CALL(pmps_wall_pnd_diffusive(H(PSystem, Particles))); }
void MpsPndNeigh::updatePnd(MpsParticle *Particles) { CALL(pmps_update_pnd(handle())); }
void MpsPndNeigh::meanPndParticlesWallDummySurface(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) { CALL(pmps_mean_pnd_wall_dummy_surface(H(PSystem, Particles))); }
void MpsPndNeigh::meanPnd(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) { CALL(pmps_mean_pnd(H(PSystem, Particles))); }
void MpsPndNeigh::meanWallPnd(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) { CALL(pmps_mean_wall_pnd(H(PSystem, Particles))); }
void MpsPndNeigh::meanNeighFluidPnd(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) { CALL(pmps_mean_neigh_fluid_pnd(H(PSystem, Particles))); }

// ---- MpsBoundaryCondition (src/MpsBoundaryCondition.cpp:66-399) --------------------------------------------------------
void MpsBoundaryCondition::updateParticleBC(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) { CALL(pmps_update_bc(H(PSystem, Particles))); }

// ---- MpsVectorMatrix (src/MpsVectorMatrix.cpp:20-93) ---------------------------------------------------------------------
void MpsVectorMatrix::correctionMatrix(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) { CALL(pmps_correction_matrix(H(PSystem, Particles))); }

// ---- MpsPressure (src/MpsPressure.cpp:31-57, 792-846, 1217-1611) -----------------------------------------------------------
void MpsPressure::calcPress(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) {
	CALL(pmps_calc_press(H(PSystem, Particles)));
	if (PSystem->mpsType == calcPressType::IMPLICIT_PND || PSystem->mpsType == calcPressType::IMPLICIT_PND_DIVU) {
		PmpsScalars s; CALL(pmps_get_scalars(g_h, &s)); solverIter = s.solverIter; solverError = s.solverError;
	}
	if (g_oodCalls == 1) { pull_scalars(); pull_arrays(); }   // end of the 0-th step (src/main.cpp:176): iteration 0 writes output
}
void MpsPressure::extrapolatePressParticlesWallDummy(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) { CALL(pmps_extrapolate_press(H(PSystem, Particles))); }
void MpsPressure::predictionPressGradient(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) { CALL(pmps_prediction_press_gradient(H(PSystem, Particles))); }
void MpsPressure::predictionWallPressGradient(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) { CALL(pmps_prediction_wall_press_gradient(H(PSystem, Particles))); }
void MpsPressure::calcPressGradient(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) { CALL(pmps_press_gradient(H(PSystem, Particles))); }
void MpsPressure::calcWallPressGradient(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) { CALL(pmps_wall_press_gradient(H(PSystem, Particles))); }

// ---- MpsShifting (src/MpsShifting.cpp:23-101 type 1, :432-691 type 2, :106-174 / :319-429 normals) ----------------------------
void MpsShifting::calcShifting(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) { CALL(pmps_shifting(H(PSystem, Particles))); }   // pmps_shifting runs the type PmpsParams.shiftingType names
void MpsShifting::calcNormalParticles(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) { CALL(pmps_calc_normals(handle())); }
void MpsShifting::calcWallNormalParticles(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) { CALL(pmps_calc_wall_normals(handle())); }
void MpsShifting::calcConcAndConcGradient(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) { CALL(pmps_shifting(H(PSystem, Particles))); }
void MpsShifting::calcWallConcAndConcGradient(MpsParticleSystem *PSystem, MpsParticle *Particles, MpsBucket *Buckets) { CALL(pmps_wall_shifting(H(PSystem, Particles))); }

// ---- PolygonMesh (src/PolygonMesh.cpp:585-795) ---------------------------------------------------------------------------------
// main.cpp calls closestPointPNDBoundaryAABB once per mesh and then updateParticlesNearPolygonMesh on mesh 0 (src/main.cpp:293-305,
// 529-545); pmps_wall_closest_point does all of it in one pass, so the per-mesh call only makes sure the device has the mesh (again
// every step for the forced mesh, whose vertices the host moves: src/main.cpp:405) and the last call of the group runs the pass.
void PolygonMesh::closestPointPNDBoundaryAABB(double reS2, double reL2, int nP, int wijType, int *Typ, int fld, int msh_id, int sta_id,
	int fem_id, int frw_id, const double *Pos, double *wallPos, double *mirrorPos, double *riw2, int *elementID, int *meshID, double *NormalWall) {
	if (msh_id < 0 || msh_id > 2) { fprintf(stderr, "libpmps: mesh id %d\n", msh_id); exit(10); }
	if (!g_meshSet[msh_id] || msh_id == frw_id) send_mesh(msh_id, NV, NF, NNormals, xDataPND, xDataNeigh, Zij, nNeighWall);
	if (msh_id == 0 && g_S) CALL(pmps_set_uniform_vel_wall(handle(), g_S->uniformVelWall));   // set by updateForcedPolygonMesh (src/main.cpp:405)
}
void PolygonMesh::updateParticlesNearPolygonMesh(double reS2, double reL2, int nP, int wijType, int *Typ, int fld, const double *riw2,
	double *niw, int *numNeighw, std::vector<int>& meshParticlesNeighbors, bool *Nw) {
	CALL(pmps_wall_closest_point(handle()));
}

// ---- output ------------------------------------------------------------------------------------------------------------------
// MpsInputOutput::writeOutputFiles (src/MpsInputOutput.cpp:861-885) picks one of four writers by flags.output_VTU.  The two ASCII ones
// (the only ones the reference's input files call working) are replaced by pmps_write_vtu: same bytes, from the device state, formatted
// by all host cores instead of one fprintf per number.  The binary variants, and PMPS_VTU_HOST=1, run the reference's own writers on the
// host arrays (refreshed by pull_arrays before every output iteration).
void MpsInputOutput::writeOutputFiles(MpsParticleSystem *PSystem, MpsParticle *Particles) {
	if (vtuType == 0 && g_h && !getenv("PMPS_VTU_HOST")) {
		char name[1024]; snprintf(name, sizeof name, "%s/output%05d.vtu", vtuOutputFoldername.c_str(), PSystem->fileNumber);
		const unsigned f = (outputPnd ? PMPS_VTU_PND : 0) | (outputNeigh ? PMPS_VTU_NEIGH : 0) | (outputDeviation ? PMPS_VTU_DEVIATION : 0) |
		                   (outputConcentration ? PMPS_VTU_CONCENTRATION : 0) | (outputNonNewtonian ? PMPS_VTU_NONNEWTONIAN : 0) |
		                   (outputAuxiliar ? PMPS_VTU_AUXILIAR : 0) | (freeSurfWall ? PMPS_VTU_FREE_SURFACE_ONLY : 0);
		CALL(pmps_write_vtu(g_h, name, f));
		return;
	}
	if (vtuType == 0) { if (freeSurfWall) writeVtuAsciiFreeSurface(PSystem, Particles); else writeVtuAscii(PSystem, Particles); }
	else { if (freeSurfWall) writeVtuBinaryFreeSurface(PSystem, Particles); else writeVtuBinary(PSystem, Particles); }
}
