// oracle/ref_driver.cpp — TEST INFRASTRUCTURE. Not part of the product; never linked into libpmps.
//
// Drives the UNMODIFIED reference engine objects (every /root/reference/src/*.cpp except main.cpp, compiled
// where they lie by oracle/Makefile) through the call order of the reference driver
// (start-up: src/main.cpp:98-186, time step: src/main.cpp:230-468) and dumps raw double/int arrays after
// each stage, so that (a) our C restatement (oracle/mps_oracle.c) and (b) the CUDA path can be checked
// against the reference's own arithmetic.  VTU output cannot be used for this: it prints "%lf" (6 decimals).
//
// usage: ref_driver <json base name> <nsteps> <dump dir> <comma list of steps to dump stage by stage | none>
//        [time]            -> with "time": no dumps, prints wall time of the step loop (CPU baseline)
// cwd must contain input/<name>.json (hard-coded prefix, src/MpsInputOutput.cpp:63,99).
// Always run the parity dumps with OMP_NUM_THREADS=1 (the reference's shifting loop races otherwise,
// src/MpsShifting.cpp:478-536).
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <set>
#include <sys/time.h>
#include <sys/stat.h>
#include <unistd.h>

#include <Eigen/Core>
#include <Eigen/Sparse>
#include "MpsParticleSystem.h"
#include "MpsParticle.h"
#include "MpsBucket.h"
#include "MpsBoundaryCondition.h"
#include "MpsInputOutput.h"
#include "MpsParticleCollision.h"
#include "MpsParticleVelPos.h"
#include "MpsPndNeigh.h"
#include "MpsPressure.h"
#include "MpsShifting.h"
#include "MpsVectorMatrix.h"
#include "MpsViscosity.h"
#define private public   // the wall tables (xDataPND, Zij, ...) are private members; the harness only reads them
#include "PolygonMesh.h"
#undef private

static double now() { timeval tv; gettimeofday(&tv, 0); return tv.tv_sec + 1e-6 * tv.tv_usec; }

// ---- tiny record container: [u32 name_len][name][u8 dtype 0=f64 1=i32 2=u8][u64 count][payload] ----
struct Dump {
	FILE* f = nullptr;
	void open(const std::string& path) { f = fopen(path.c_str(), "wb"); if (!f) { perror(path.c_str()); exit(2); } }
	void close() { if (f) fclose(f); f = nullptr; }
	void rec(const char* name, unsigned char dt, const void* p, unsigned long long n, size_t esz) {
		unsigned int l = (unsigned int)strlen(name);
		fwrite(&l, 4, 1, f); fwrite(name, 1, l, f); fwrite(&dt, 1, 1, f); fwrite(&n, 8, 1, f);
		if (n) fwrite(p, esz, n, f);
	}
	void d(const char* name, const double* p, size_t n) { rec(name, 0, p, n, 8); }
	void i(const char* name, const int* p, size_t n) { rec(name, 1, p, n, 4); }
	void b(const char* name, const bool* p, size_t n) { rec(name, 2, p, n, 1); }
};

struct Sim {
	MpsParticleSystem* ps; MpsParticle* pa; MpsBucket* bk; MpsBoundaryCondition* bc; MpsInputOutput* io;
	MpsPndNeigh* pn; MpsViscosity* vi; MpsParticleCollision* co; MpsShifting* sh; MpsPressure* pr;
	MpsVectorMatrix* vm; MpsParticleVelPos* vp; PolygonMesh* mesh;
	std::vector<int> nearMesh; double *frwX = nullptr, *frwY = nullptr, *frwZ = nullptr;
};

static void dump_params(const Sim& s, const std::string& path) {
	FILE* f = fopen(path.c_str(), "w");
	const MpsParticleSystem* P = s.ps;
	fprintf(f, "{\n");
#define PD(x) fprintf(f, "  \"%s\": %.17g,\n", #x, (double)P->x)
#define PI(x) fprintf(f, "  \"%s\": %d,\n", #x, (int)P->x)
	PI(wallType); PI(femOn); PI(forcedOn);
	PD(domainMinX); PD(domainMinY); PD(domainMinZ); PD(domainMaxX); PD(domainMaxY); PD(domainMaxZ);
	PI(domainTypeBC); PI(numBC); PI(limitTypeBC); PI(periodicDirectionX); PI(periodicDirectionY); PI(periodicDirectionZ);
	PD(densityFluid); PD(densityWall); PD(KNM_VS1); PI(fluidType); PD(gravityX); PD(gravityY); PD(gravityZ);
	PD(KNM_VS2); PD(DNS_FL1); PD(DNS_FL2); PD(DNS_SDT); PD(PHI_1); PD(PHI_2); PD(cohes); PD(MEU0); PD(DG); PD(I0); PD(mm); PD(Cd); PD(VF_min); PD(VF_max); PI(Fraction_method);
	PD(dim); PD(partDist); PD(timeStep); PD(timeSimulation); PI(iterOutput); PD(cflNumber); PI(weightType); PI(slipCondition);
	PD(reS); PD(reL); PI(gradientType); PI(gradientCorrection); PD(relaxPress); PI(mpsType); PD(soundSpeed); PD(gamma);
	PI(solverType); PD(alphaCompressibility); PD(relaxPND); PD(shiftingType); PD(dri); PD(coefA); PD(machNumber);
	PI(pndType); PD(diffusiveCoef); PI(repulsiveForceType); PD(reRepulsiveForce); PD(expectMaxVelocity);
	PD(repForceCoefMitsume); PD(repForceCoefLennardJones); PD(repForceCoefMonaghanKajtar); PD(EPS_RE);
	PI(freeSurfType); PD(pndThreshold); PD(neighThreshold); PD(npcdThreshold); PD(thetaThreshold); PD(normThreshold);
	PI(collisionType); PD(collisionRatio); PD(distLimitRatio); PD(lambdaCollision);
	PI(ghost); PI(fluid); PI(wall); PI(dummyWall); PI(surface); PI(inner); PI(other);
	PD(epsilonZero); PI(numOfMeshs);
	PD(physDomMinX); PD(physDomMinY); PD(physDomMinZ); PD(physDomMaxX); PD(physDomMaxY); PD(physDomMaxZ);
	PI(numBucketsX); PI(numBucketsY); PI(numBucketsZ); PI(numBucketsXY); PI(numBucketsXYZ);
	PD(bucketSide); PD(invBucketSide);
	fprintf(f, "  \"periodicLength\": [%.17g, %.17g, %.17g],\n", P->periodicLength[0], P->periodicLength[1], P->periodicLength[2]);
	PD(coeffViscosity); PD(coeffViscMultiphase); PD(coeffPressGrad); PD(coeffPressEMPS); PD(coeffPressWCMPS);
	PD(coeffShifting1); PD(coeffShifting2); PD(coeffPPE); PD(coeffPPESource);
	PD(pndGradientZero); PD(pndSmallZero); PD(pndLargeZero); PD(lambdaZero); PD(numNeighZero);
	PD(distCollisionLimit); PD(distCollisionLimit2); PD(restitutionCollision);
	PD(betaPnd); PD(betaNeigh); PD(delta2); PD(thetaArc); PD(hThreshold2); PD(dstThreshold2); PD(normThreshold2);
	PD(invPartDist); PD(reS2); PD(reL2);
	fprintf(f, "  \"uniformVelWall\": [%.17g, %.17g, %.17g],\n", P->uniformVelWall[0], P->uniformVelWall[1], P->uniformVelWall[2]);
	fprintf(f, "  \"numParticles\": %d, \"numParticlesZero\": %d\n}\n", s.pa->numParticles, s.pa->numParticlesZero);
#undef PD
#undef PI
	fclose(f);
}

static void dump_mesh(const Sim& s, const std::string& path) {
	Dump D; D.open(path);
	for (int me = 0; me < s.ps->numOfMeshs; me++) {
		const PolygonMesh& M = s.mesh[me];
		std::vector<double> V(M.NV.rows() * 3), Nn(M.NF.rows() * 3); std::vector<int> F(M.NF.rows() * 3);
		for (int r = 0; r < M.NV.rows(); r++) for (int c = 0; c < 3; c++) V[r * 3 + c] = M.NV(r, c);
		for (int r = 0; r < M.NF.rows(); r++) for (int c = 0; c < 3; c++) { F[r * 3 + c] = M.NF(r, c); Nn[r * 3 + c] = M.NNormals(r, c); }
		char nm[64];
		snprintf(nm, 64, "mesh%d_V", me); D.d(nm, V.data(), V.size());
		snprintf(nm, 64, "mesh%d_F", me); D.i(nm, F.data(), F.size());
		snprintf(nm, 64, "mesh%d_N", me); D.d(nm, Nn.data(), Nn.size());
		snprintf(nm, 64, "mesh%d_xDataPND", me); D.d(nm, M.xDataPND.data(), M.xDataPND.size());
		snprintf(nm, 64, "mesh%d_xDataNeigh", me); D.d(nm, M.xDataNeigh.data(), M.xDataNeigh.size());
		snprintf(nm, 64, "mesh%d_Zij", me); D.d(nm, M.Zij.data(), M.Zij.size());
		snprintf(nm, 64, "mesh%d_nNeighWall", me); D.i(nm, M.nNeighWall.data(), M.nNeighWall.size());
	}
	D.close();
}

// every array that is live on the hot path (include/MpsParticle.h:139-241)
static void dump_state(const Sim& s, const std::string& path) {
	Dump D; D.open(path);
	const MpsParticle* p = s.pa; size_t n = p->numParticlesZero;
	int np[2] = { p->numParticles, p->numParticlesZero }; D.i("numParticles", np, 2);
	double sc[4] = { s.ps->velMax, s.ps->CFLcurrent, s.ps->timeCurrent, (double)s.ps->numOfIterations }; D.d("scalars", sc, 4);
	D.i("particleType", p->particleType, n); D.i("particleBC", p->particleBC, n); D.i("numNeigh", p->numNeigh, n);
	D.i("PTYPE", p->PTYPE, n); D.i("nearMeshType", p->nearMeshType, n); D.b("particleNearWall", p->particleNearWall, n);
	D.i("numNeighWallContribution", p->numNeighWallContribution, n); D.i("elementID", p->elementID, n);
	D.d("press", p->press, n); D.d("pndi", p->pndi, n); D.d("pndki", p->pndki, n); D.d("pndski", p->pndski, n);
	D.d("pndSmall", p->pndSmall, n); D.d("npcdDeviation2", p->npcdDeviation2, n); D.d("concentration", p->concentration, n);
	D.d("velDivergence", p->velDivergence, n); D.d("diffusiveTerm", p->diffusiveTerm, n);
	D.d("pndWallContribution", p->pndWallContribution, n); D.d("distParticleWall2", p->distParticleWall2, n);
	D.d("deviationDotPolygonNormal", p->deviationDotPolygonNormal, n);
	D.d("RHO", p->RHO, n); D.d("MEU", p->MEU, n); D.d("pndAuxVar1", p->pndAuxVar1, n); D.d("pndAuxVar2", p->pndAuxVar2, n);
	D.d("pos", p->pos, 3 * n); D.d("vel", p->vel, 3 * n); D.d("acc", p->acc, 3 * n); D.d("accStar", p->accStar, 3 * n);
	D.d("npcdDeviation", p->npcdDeviation, 3 * n); D.d("gradConcentration", p->gradConcentration, 3 * n);
	D.d("correcMatrixRow1", p->correcMatrixRow1, 3 * n); D.d("correcMatrixRow2", p->correcMatrixRow2, 3 * n);
	D.d("correcMatrixRow3", p->correcMatrixRow3, 3 * n);
	D.d("particleAtWallPos", p->particleAtWallPos, 3 * n); D.d("mirrorParticlePos", p->mirrorParticlePos, 3 * n);
	D.d("polygonNormal", p->polygonNormal, 3 * n); D.d("forceWall", p->forceWall, 3 * n);
	if (s.ps->freeSurfType == calcBCType::PND_ARC) D.d("normal", p->normal, 3 * n);   // (only then: the committed goldens keep their layout)
	if (s.ps->mpsType == calcPressType::IMPLICIT_PND || s.ps->mpsType == calcPressType::IMPLICIT_PND_DIVU) {
		D.d("pressurePPE", p->pressurePPE.data(), n); D.d("sourceTerm", p->sourceTerm.data(), n);
		double it[2] = { (double)s.pr->getSolverIter(), s.pr->getSolverError() }; D.d("solver", it, 2);
	}
	if (s.ps->fluidType == viscType::NON_NEWTONIAN) {
		D.d("Cv", p->Cv, n); D.d("II", p->II, n); D.d("MEU_Y", p->MEU_Y, n); D.d("Inertia", p->Inertia, n); D.d("p_rheo_new", p->p_rheo_new, n);
		D.d("p_smooth", p->p_smooth, n); D.d("VF", p->VF, n); D.d("S11", p->S11, n); D.d("S12", p->S12, n); D.d("S13", p->S13, n);
		D.d("S22", p->S22, n); D.d("S23", p->S23, n); D.d("S33", p->S33, n);
		double cv[1] = { s.ps->CFLvisc }; D.d("CFLvisc", cv, 1);
	}
	D.i("firstParticleInBucket", p->firstParticleInBucket, s.ps->numBucketsXYZ);
	D.i("lastParticleInBucket", p->lastParticleInBucket, s.ps->numBucketsXYZ);
	D.i("nextParticleInSameBucket", p->nextParticleInSameBucket, n);
	D.i("bucketPeriodicBC", p->bucketPeriodicBC, s.ps->numBucketsXYZ);
	D.close();
}

static void polygon_wall_pass(Sim& s) {   // src/main.cpp:293-314 (same three calls are made at start-up, :533-554)
	MpsParticleSystem* P = s.ps; MpsParticle* A = s.pa;
	for (int me = 0; me < P->numOfMeshs; me++)
		s.mesh[me].closestPointPNDBoundaryAABB(P->reS2, P->reL2, A->numParticles, P->weightType, A->particleType, P->fluid, me,
			meshType::FIXED, meshType::DEFORMABLE, meshType::FORCED, A->pos, A->particleAtWallPos, A->mirrorParticlePos,
			A->distParticleWall2, A->elementID, A->nearMeshType, A->polygonNormal);
	s.nearMesh.clear();
	s.mesh[0].updateParticlesNearPolygonMesh(P->reS2, P->reL2, A->numParticles, P->weightType, A->particleType, P->fluid,
		A->distParticleWall2, A->pndWallContribution, A->numNeighWallContribution, s.nearMesh, A->particleNearWall);
}

int main(int argc, char** argv) {
	if (argc < 5) { fprintf(stderr, "usage: ref_driver <json base> <nsteps> <dump dir> <steps csv|none> [time]\n"); return 2; }
	std::string name = argv[1]; int nsteps = atoi(argv[2]); std::string dir = argv[3];
	std::set<int> dumpSteps; bool timing = argc > 5 && !strcmp(argv[5], "time");
	if (strcmp(argv[4], "none")) { char* t = strtok(argv[4], ","); while (t) { dumpSteps.insert(atoi(t)); t = strtok(0, ","); } }
	mkdir(dir.c_str(), 0777);
	{ // the reference asks for the case name on stdin (src/MpsInputOutput.cpp:91-93)
		std::string tmp = dir + "/.stdin"; FILE* f = fopen(tmp.c_str(), "w"); fprintf(f, "%s\n", name.c_str()); fclose(f);
		if (!freopen(tmp.c_str(), "r", stdin)) { perror("freopen"); return 2; }
	}
	Eigen::initParallel();
	Sim s; memset(&s.ps, 0, sizeof(void*));
	s.ps = new MpsParticleSystem(); s.pa = new MpsParticle(); s.io = new MpsInputOutput();
	MpsParticleSystem* P = s.ps; MpsParticle* A = s.pa;
	s.io->readInputFile(P, A);
	A->setParameters(P);
	s.bk = new MpsBucket(); s.bk->allocateMemory(P, A);
	s.bc = new MpsBoundaryCondition(); s.bc->setBucketBC(P, A, s.bk);
	s.bk->updateParticlesID(P, A);
	A->checkParticleOutDomain(P);
	s.pn = new MpsPndNeigh(); s.mesh = nullptr;
	const bool poly = P->wallType == boundaryWallType::POLYGON;
	if (poly) {
		s.mesh = new PolygonMesh[P->numOfMeshs];
		for (int me = 0; me < P->numOfMeshs; me++) {
			s.mesh[me].initPolygonMesh(A->numParticles);
			if (me == 0) s.mesh[me].readPolygonMeshFile(s.io->meshRigidFilename);
			if (P->femOn && me == 1) s.mesh[me].readPolygonMeshFile(s.io->meshDeformableFilename);
			if (P->forcedOn && me == 2) s.mesh[me].readPolygonMeshFile(s.io->meshForcedFilename);
		}
		for (int me = 0; me < P->numOfMeshs; me++) s.mesh[me].initWijnNeigh((int)P->dim, P->weightType, P->partDist, P->reL, P->reS);
		if (P->forcedOn) {
			int nv = s.mesh[meshType::FORCED].NV.rows();
			s.frwX = (double*)malloc(8 * nv); s.frwY = (double*)malloc(8 * nv); s.frwZ = (double*)malloc(8 * nv);
			for (int k = 0; k < nv; k++) { s.frwX[k] = s.mesh[meshType::FORCED].NV(k, 0); s.frwY[k] = s.mesh[meshType::FORCED].NV(k, 1); s.frwZ[k] = s.mesh[meshType::FORCED].NV(k, 2); }
			P->uniformVelWall[0] = P->uniformVelWall[1] = P->uniformVelWall[2] = 0.0;
		}
		polygon_wall_pass(s);
		s.pn->calcWallNPCD(P, A, s.bk);
	}
	s.vi = new MpsViscosity(); s.co = new MpsParticleCollision(); s.sh = new MpsShifting(); s.pr = new MpsPressure();
	const bool implicit = P->mpsType == calcPressType::IMPLICIT_PND || P->mpsType == calcPressType::IMPLICIT_PND_DIVU;
	if (implicit) s.pr->allocateMemory(A);
	s.vm = new MpsVectorMatrix(); s.vp = new MpsParticleVelPos();

	if (!timing) { dump_params(s, dir + "/params.json"); if (poly) dump_mesh(s, dir + "/mesh.bin"); dump_state(s, dir + "/pre_init.bin"); }
	// 0-th step (src/main.cpp:157-180)
	s.pn->setInitialPndNumberOfNeighNPCD(P, A, s.bk);
	if (poly) s.pn->meanWallPnd(P, A, s.bk);
	s.pn->meanPnd(P, A, s.bk);
	s.pn->meanNeighFluidPnd(P, A, s.bk);
	if (P->freeSurfType == calcBCType::PND_ARC) { s.sh->calcNormalParticles(P, A, s.bk); if (poly) s.sh->calcWallNormalParticles(P, A, s.bk); }
	s.bc->updateParticleBC(P, A, s.bk);
	s.pr->calcPress(P, A, s.bk);
	if (!timing) dump_state(s, dir + "/init.bin");

	double t0 = now();
	for (int step = 1; step <= nsteps; step++) {
		const bool dd = !timing && dumpSteps.count(step);
		int stage = 0; char tag[256];
#define STAGE(nm) do { if (dd) { snprintf(tag, 256, "%s/s%04d_%02d_%s.bin", dir.c_str(), step, stage, nm); dump_state(s, tag); } stage++; } while (0)
		if (dd) { snprintf(tag, 256, "%s/s%04d_pre.bin", dir.c_str(), step); dump_state(s, tag); }
		s.bk->updateParticlesID(P, A); STAGE("updateParticlesID");
		if (P->fluidType == viscType::NON_NEWTONIAN) { s.vi->calcVolumeFraction(P, A, s.bk); s.vi->calcViscosityInteractionVal(P, A, s.bk); STAGE("nonNewtonian1"); }
		s.vi->calcViscosity(P, A, s.bk); STAGE("calcViscosity");
		A->addGravity(P, A); STAGE("addGravity");
		if (P->relaxPress < 1.0) { s.pr->predictionPressGradient(P, A, s.bk); if (poly) s.pr->predictionWallPressGradient(P, A, s.bk); STAGE("predictionPressGradient"); }
		s.vp->predictionVelocityPosition(P, A); STAGE("predictionVelocityPosition");
		A->checkParticleOutDomain(P); STAGE("checkParticleOutDomain1");
		if (P->collisionType == colType::PC) s.co->checkParticleCollisions(P, A, s.bk); else s.co->checkDynamicParticleCollisions(P, A, s.bk);
		STAGE("collisions");
		if (poly) { polygon_wall_pass(s); STAGE("wallClosestPoint"); s.pn->calcWallNPCD(P, A, s.bk); STAGE("calcWallNPCD"); }
		if (P->gradientCorrection) { s.vm->correctionMatrix(P, A, s.bk); STAGE("correctionMatrix1"); }
		s.pn->calcPndnNeighNPCD(P, A, s.bk); STAGE("calcPndnNeighNPCD");
		if (P->pndType == calcPNDType::DIFFUSIVE) {
			s.pn->calcPndDiffusiveTerm(P, A, s.bk); STAGE("calcPndDiffusiveTerm");
			if (poly) {
				if (P->slipCondition == slipBC::FREE_SLIP) s.pn->calcWallSlipPndDiffusiveTerm(P, A, s.bk);
				else if (P->slipCondition == slipBC::NO_SLIP) s.pn->calcWallNoSlipPndDiffusiveTerm(P, A, s.bk);
				STAGE("calcWallPndDiffusiveTerm");
			}
			s.pn->updatePnd(A); STAGE("updatePnd");
			s.pn->meanPndParticlesWallDummySurface(P, A, s.bk); STAGE("meanPndParticlesWallDummySurface");
		} else if (P->pndType == calcPNDType::MEAN_SUM_WIJ) {
			if (poly) s.pn->meanWallPnd(P, A, s.bk);
			s.pn->meanPnd(P, A, s.bk); STAGE("meanPnd");
		}
		if (P->freeSurfType == calcBCType::PND_ARC) { s.sh->calcNormalParticles(P, A, s.bk); if (poly) s.sh->calcWallNormalParticles(P, A, s.bk); STAGE("normals"); }
		s.bc->updateParticleBC(P, A, s.bk); STAGE("updateParticleBC");
		s.pr->calcPress(P, A, s.bk); STAGE("calcPress");
		if (!poly) { s.pr->extrapolatePressParticlesWallDummy(P, A, s.bk); STAGE("extrapolatePress"); }
		if (P->gradientCorrection) { s.vm->correctionMatrix(P, A, s.bk); STAGE("correctionMatrix2"); }
		s.pr->calcPressGradient(P, A, s.bk); STAGE("calcPressGradient");
		if (poly) { s.pr->calcWallPressGradient(P, A, s.bk); STAGE("calcWallPressGradient"); }
		if (P->fluidType == viscType::NON_NEWTONIAN) {
			s.vi->calcVolumeFraction(P, A, s.bk); s.vi->calcViscosityInteractionVal(P, A, s.bk);
			if (poly) { if (P->slipCondition == slipBC::FREE_SLIP) s.vi->calcWallSlipViscosityInteractionVal(P, A, s.bk); else if (P->slipCondition == slipBC::NO_SLIP) s.vi->calcWallNoSlipViscosityInteractionVal(P, A, s.bk); }
			STAGE("nonNewtonian2");
		}
		if (poly) {
			if (P->slipCondition == slipBC::FREE_SLIP) s.vi->calcWallSlipViscosity(P, A, s.bk);
			else if (P->slipCondition == slipBC::NO_SLIP) s.vi->calcWallNoSlipViscosity(P, A, s.bk);
			STAGE("calcWallViscosity");
		}
		if (P->forcedOn) s.mesh[meshType::FORCED].updateForcedPolygonMesh(s.frwX, s.frwY, s.frwZ, P->uniformVelWall, P->timeStep, P->timeCurrent);
		s.vp->correctionVelocityPosition(P, A); STAGE("correctionVelocityPosition");
		A->checkParticleOutDomain(P); STAGE("checkParticleOutDomain2");
		if (P->shiftingType == 1) { s.sh->calcShifting(P, A, s.bk); STAGE("calcShifting"); }
		else if (P->shiftingType == 2) {
			s.sh->calcConcAndConcGradient(P, A, s.bk); STAGE("calcConcAndConcGradient");
			if (poly) { s.sh->calcWallConcAndConcGradient(P, A, s.bk); STAGE("calcWallConcAndConcGradient"); }
		}
		if (!poly) { s.vp->updateVelocityParticlesWallDummy(P, A, s.bk); STAGE("updateVelocityParticlesWallDummy"); }
		A->checkParticleOutDomain(P); STAGE("checkParticleOutDomain3");
		P->numOfIterations++; P->timeCurrent += P->timeStep;
		if (dd) { snprintf(tag, 256, "%s/s%04d_end.bin", dir.c_str(), step); dump_state(s, tag); }
#undef STAGE
	}
	double t1 = now();
	if (!timing) dump_state(s, dir + "/final.bin");
	printf("REF_DRIVER steps=%d particles=%d loop_seconds=%.6f particle_steps_per_s=%.6e velMax=%.17g\n",
		nsteps, A->numParticles, t1 - t0, nsteps > 0 ? (double)A->numParticles * nsteps / (t1 - t0) : 0.0, P->velMax);
	return 0;
}
