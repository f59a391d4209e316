/* pmps.h — C ABI of the B200-native PolyMPS particle hot path (libpmps.so).
 *
 * Drop-in boundary: the reference has no plugin/FFI layer; its seam is the set of
 * engine-class methods  void f(MpsParticleSystem*, MpsParticle*, MpsBucket*)  called once per time step from
 * mainLoopOfSimulation (reference src/main.cpp:230-468).  Every pmps_* stage call below replaces the body of exactly
 * one of those methods (cited per function); the C++ shims in polymps_b200/host/ keep the reference class names and
 * signatures and forward here.  Plain pointers and sizes only: no C++/torch types cross this boundary.
 *
 * Conventions
 *  - every function returns 0 on success, <0 on error (PMPS_E_*); pmps_last_error() gives the message.
 *    The reference itself is void/exit()/throw (src/MpsBucket.cpp:104, src/MpsInputOutput.cpp:521); the shims map
 *    a non-zero return to the reference behaviour (message on stderr + exit(10)).
 *  - host arrays use the reference layout: scalars a[i], vectors a[3*i+k], int32 ids, `bool` flags as 1 byte
 *    (include/MpsParticle.h:139-241).  The library never frees or reallocates a host pointer.
 *  - a handle owns all device memory; one host thread per handle (the reference is single-driver too).
 *  - there is NO CPU fallback: if no CUDA device is usable pmps_create fails with PMPS_E_CUDA.
 */
#ifndef PMPS_H_
#define PMPS_H_
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define PMPS_ABI_VERSION 5

enum { PMPS_OK = 0, PMPS_E_ARG = -1, PMPS_E_CUDA = -2, PMPS_E_STATE = -3, PMPS_E_UNSUPPORTED = -4, PMPS_E_OVERFLOW = -5 };

/* POD snapshot of the reference's MpsParticleSystem after setParameters()/MpsBucket::allocateMemory()
 * (include/MpsParticleSystem.h:133-309, src/MpsParticle.cpp:136-227, src/MpsBucket.cpp:20-67).  Same names. */
typedef struct PmpsParams {
	int32_t abi_version;            /* = PMPS_ABI_VERSION */
	/* options (enums as in include/MpsParticleSystem.h:20-124) */
	int32_t dim;                    /* 2 or 3 ((int)PSystem->dim) */
	int32_t wallType;               /* 0 particle walls, 1 polygon walls */
	int32_t mpsType;                /* 0 E-MPS, 1 WC-MPS, 2 implicit PND, 3 implicit PND + div(u) */
	int32_t solverType;             /* 0 CG, 1 BiCGSTAB */
	int32_t pndType;                /* 0 sum wij, 1 mean, 2 diffusive */
	int32_t weightType;             /* 0..4 (src/MpsParticle.cpp:57-108) */
	int32_t gradientType;           /* 0..3 (src/MpsPressure.cpp:1288-1300) */
	int32_t gradientCorrection;     /* bool */
	int32_t slipCondition;          /* 0 no-slip, 1 free-slip */
	int32_t freeSurfType;           /* 0 PND+neigh, 1 PND+NPCD, 2 PND+ARC (particle normals + arc test, src/MpsBoundaryCondition.cpp:205-390) */
	int32_t collisionType;          /* 0 PC, 1 DPC */
	int32_t shiftingType;           /* 0 off, 1 DW*Uij, 2 grad C */
	int32_t repulsiveForceType;     /* 0 Harada 1 Mitsume 2 Lennard-Jones 3 Monaghan-Kajtar */
	int32_t fluidType;              /* 0 Newtonian, 1 non-Newtonian (two-phase granular rheology; with polygon walls in 3D only, as the reference's wall variants are) */
	int32_t domainTypeBC;           /* 0 none, 1 periodic in ONE direction (periodicAxis; src/MpsBucket.cpp:142-289, src/MpsParticle.cpp:238-269) */
	int32_t ghost, fluid, wall, dummyWall;      /* particle type ids */
	int32_t surface, inner, other;              /* particleBC ids */
	int32_t numBucketsX, numBucketsY, numBucketsZ;
	int32_t numOfMeshs;
	int32_t shiftingJacobi;         /* 1: shifting reads a position snapshot (default, parallel-safe);
	                                   the reference's 1-thread behaviour is a Gauss-Seidel sweep in index order */
	int32_t periodicAxis;           /* domainTypeBC 1: 0 x, 1 y, 2 z (the reference's domain limits accept exactly one periodic direction) */
	int32_t fractionMethod;         /* PSystem->Fraction_method: 1 neighbour count, 2 weighted (src/MpsViscosity.cpp:106,145) */
	/* geometry */
	double domainMinX, domainMinY, domainMinZ, domainMaxX, domainMaxY, domainMaxZ;   /* AFTER the +-3 lo extension */
	double bucketSide, invBucketSide;
	double partDist, timeStep, epsilonZero;
	double reS, reL, reS2, reL2;
	/* physics */
	double gravityX, gravityY, gravityZ;
	double densityWall, DNS_FL1, DNS_FL2;
	double soundSpeed, gamma;
	double relaxPress, alphaCompressibility;
	double reRepulsiveForce, expectMaxVelocity;
	double repForceCoefMitsume, repForceCoefLennardJones, repForceCoefMonaghanKajtar;
	double uniformVelWall[3];
	/* derived coefficients (src/MpsParticle.cpp:196-222) */
	double pndSmallZero, pndLargeZero, pndGradientZero, lambdaZero, numNeighZero;
	double coeffViscMultiphase, coeffPressGrad, coeffPressEMPS, coeffPressWCMPS;
	double coeffShifting1, coeffShifting2, coeffPPE, coeffPPESource;
	double distCollisionLimit2, restitutionCollision, lambdaCollision;
	double betaPnd, betaNeigh, delta2;
	double velMax;                  /* velMax carried in from the host (0 at start, src/MpsParticle.cpp:212) */
	/* non-Newtonian rheology (fluidType 1), names as in include/MpsParticleSystem.h; src/MpsInputOutput.cpp:250-316 */
	double KNM_VS1, KNM_VS2, DNS_SDT, PHI_1, PHI_2, cohes, MEU0, DG, I0, mm, Cd, VF_min, VF_max;
	double relaxPND;                /* semi_implicit_mps.source_term.relax_pnd: the divergence part of the IMPLICIT_PND_DIVU source (src/MpsPressure.cpp:373) */
	double timeCurrent;             /* time at pmps_create; the fused step advances its own copy by timeStep (src/main.cpp:463) */
	/* ARC free-surface detection (freeSurfType 2), src/MpsParticle.cpp:219-222 */
	double thetaArc, hThreshold2, dstThreshold2, normThreshold2;
	/* periodic domain (domainTypeBC 1), along periodicAxis: PSystem->periodicLength[axis], physDomMin / physDomMax of that axis
	 * (src/MpsBucket.cpp:150-156).  The first and the last bucket layer of the axis alias the last and the first inner layer, shifted
	 * by -/+ periodicLength (copyDataBetweenBuckets :356-417, getPeriodicLengths src/MpsParticle.cpp:48-54) */
	double periodicLength, physDomMin, physDomMax;
	double powerLawN;               /* PSystem->N, physical.rheological.power_law_index: read only by the polygon-wall variants of the rheology (src/MpsViscosity.cpp:1090,1505) */
} PmpsParams;

/* Pointers into the host's MpsParticle arrays (include/MpsParticle.h:139-241); NULL = not transferred. */
typedef struct PmpsHostArrays {
	int32_t numParticles;           /* live particles (MpsParticle::numParticles) */
	int32_t numParticlesZero;       /* allocated length (MpsParticle::numParticlesZero) */
	int32_t *particleType, *particleBC, *numNeigh, *PTYPE, *nearMeshType, *numNeighWallContribution, *elementID;
	uint8_t *particleNearWall;      /* C++ bool, 1 byte */
	double *press, *pndi, *pndki, *pndski, *pndSmall, *npcdDeviation2, *concentration, *velDivergence, *diffusiveTerm;
	double *pndWallContribution, *distParticleWall2, *deviationDotPolygonNormal, *RHO, *MEU, *pndAuxVar1, *pndAuxVar2;
	double *pressurePPE, *sourceTerm;                      /* Eigen::VectorXd::data() */
	double *pos, *vel, *acc, *accStar, *npcdDeviation, *gradConcentration;
	double *correcMatrixRow1, *correcMatrixRow2, *correcMatrixRow3;
	double *particleAtWallPos, *mirrorParticlePos, *polygonNormal, *forceWall, *dvelCollision;
	double *normal;                 /* particle normal of the ARC free-surface detection (src/MpsShifting.cpp:106-174,319-429) */
	/* non-Newtonian fields (allocated by the reference only when fluidType == 1, src/MpsInputOutput.cpp:729-744) */
	double *Cv, *II, *MEU_Y, *Inertia, *p_rheo_new, *p_smooth, *VF, *S11, *S12, *S13, *S22, *S23, *S33;
} PmpsHostArrays;

typedef struct PmpsScalars {      /* what displayInfo prints (src/MpsInputOutput.cpp:45-53) + health counters */
	int32_t numParticles;
	int32_t solverIter;
	double velMax, CFLcurrent, solverError;
	double maxDisplacement;       /* max |x - x(at bucket rebuild)| seen this step; > (bucketSide-reL)/2 means the
	                                 reference's own stale-bucket search would start missing neighbours */
	int32_t outOfDomainEvents;    /* particles turned into ghosts so far (src/MpsParticle.cpp:286-396) */
	int32_t maxBucketCount;       /* fullest bucket at the last rebuild */
	double CFLvisc;               /* fluidType 1: timeStep*max(MEU)/(rho_mix lo^2), src/MpsViscosity.cpp:690 */
} PmpsScalars;

typedef struct pmps_handle pmps_t;

/* ---- lifetime ---------------------------------------------------------------------------------------------- */
int pmps_create(const PmpsParams* params, int device, pmps_t** out);
int pmps_destroy(pmps_t* h);
const char* pmps_last_error(const pmps_t* h);       /* h may be NULL: last error of a failed pmps_create */
int pmps_abi_version(void);

/* polygon mesh after the reference's initMesh (src/main.cpp:471-555): vertices/faces/normals as produced by
 * igl::remove_duplicate_vertices + per_face_normals (src/PolygonMesh.cpp:98-106) and the four 50-entry wall
 * tables of initWijnNeigh (src/PolygonMesh.cpp:377-527).  Call again for a forced mesh after it moved (:413). */
int pmps_set_mesh(pmps_t* h, int mesh_id, int nV, const double* V, int nF, const int32_t* F, const double* faceNormals,
                  int nTable, const double* xDataPND, const double* xDataNeigh, const double* Zij, const int32_t* nNeighWall);

/* ---- transfers --------------------------------------------------------------------------------------------- */
int pmps_upload(pmps_t* h, const PmpsHostArrays* a);           /* all non-NULL arrays, host (reference) order */
int pmps_download(pmps_t* h, PmpsHostArrays* a);               /* scatter back to host order; non-NULL arrays only */
int pmps_get_scalars(pmps_t* h, PmpsScalars* out);
/* Optional: page-lock a host array the caller will pass to pmps_upload / pmps_download / pmps_step_host again and again, so the
 * copies run as asynchronous DMA (about 2x the pageable rate).  The library never pins caller memory on its own.  CONTRACT: the
 * range must stay allocated, at this address, until pmps_unregister_host(ptr) or pmps_destroy; freeing or reallocating a
 * registered array first is undefined behaviour (the driver keeps a stale registration).  Arrays the caller already pinned
 * (cudaHostAlloc / cudaHostRegister) are accepted and left alone.  The reference's MpsParticle arrays are malloc'd once and
 * never freed (src/MpsInputOutput.cpp:667-750, empty destructors src/MpsParticle.cpp:20-22): the shim registers them. */
int pmps_register_host(pmps_t* h, void* ptr, size_t bytes);
int pmps_unregister_host(pmps_t* h, void* ptr);
int pmps_set_velmax(pmps_t* h, double velMax);
int pmps_set_uniform_vel_wall(pmps_t* h, const double v[3]);
/* one host-driven step: upload `in`, step once, download `out` (non-NULL arrays), the copies the step does not wait for
 * overlapped with it on a second stream.  Same results as pmps_upload + pmps_step(h, 1) + pmps_download. */
int pmps_step_host(pmps_t* h, const PmpsHostArrays* in, PmpsHostArrays* out);
/* device-resident inputs (bench `value` leg, multi-GPU slabs): same layout as the host arrays, device pointers */
/* The reference's particle file straight into the device layout (replaces MpsInputOutput::readMpsParticleFile,
 * src/MpsInputOutput.cpp:645-857, for a fresh handle): "0", N, N rows "type x y z vx vy vz press pressAverage".  Parsed in chunks by a few
 * host threads; the other live fields get the reference's initial values.  The state is as after pmps_upload of the arrays the
 * reference's reader builds; no host array of N entries is ever allocated. */
int pmps_load_grid(pmps_t* h, const char* path, int* numParticles);
int pmps_upload_device(pmps_t* h, const PmpsHostArrays* device_arrays);
/* The same rows with new values — a host-driven loop that hands back the particle table it received (a slab rank between two
 * re-partitions): bucket order, hit lists and the partition survive, the displacement trackers decide about rebuilds as inside a step. */
int pmps_upload_values_device(pmps_t* h, const PmpsHostArrays* device_arrays);
/* Row sets: gather / scatter the particles with host indices d_rows[0..n) from / into compact device arrays of n entries (the non-NULL
 * members of the struct).  The upload makes a NEW particle table of the handle (a slab re-partition that rewrites only the rows that
 * changed hands or places): bucket order and lists are re-made before the next pass; numParticles = the new live count. */
int pmps_download_rows_device(pmps_t* h, PmpsHostArrays* device_arrays, int n, const int32_t* d_rows);
int pmps_upload_rows_device(pmps_t* h, const PmpsHostArrays* device_arrays, int n, const int32_t* d_rows, int numParticles);
int pmps_download_device(pmps_t* h, PmpsHostArrays* device_arrays);   /* gather to host order into device arrays */

/* ---- stage calls: one per reference method, in mainLoopOfSimulation order (src/main.cpp:256-456) ----------- */
int pmps_update_buckets(pmps_t* h);              /* MpsBucket::updateParticlesID            src/MpsBucket.cpp:299-353 (+ copyDataBetweenBuckets :356-417 for a periodic domain) */
int pmps_calc_volume_fraction(pmps_t* h);        /* MpsViscosity::calcVolumeFraction        src/MpsViscosity.cpp:104-190 (fluidType 1) */
int pmps_calc_viscosity_interaction(pmps_t* h);  /* MpsViscosity::calcViscosityInteractionVal src/MpsViscosity.cpp:194-700 (fluidType 1) */
int pmps_calc_wall_viscosity_interaction(pmps_t* h);  /* MpsViscosity::calcWallSlipViscosityInteractionVal / calcWallNoSlipViscosityInteractionVal src/MpsViscosity.cpp:703-1526
                                                       * (fluidType 1, polygon walls, by slipCondition; called after the one above at src/main.cpp:391-399 only) */
int pmps_set_time(pmps_t* h, double timeCurrent);/* PSystem->timeCurrent as the rheology reads it (3D branch, src/MpsViscosity.cpp:529) */
int pmps_calc_viscosity(pmps_t* h);              /* MpsViscosity::calcViscosity             src/MpsViscosity.cpp:20-100 */
int pmps_add_gravity(pmps_t* h);                 /* MpsParticle::addGravity                 src/MpsParticle.cpp:111-130 */
int pmps_prediction_press_gradient(pmps_t* h);      /* MpsPressure::predictionPressGradient     src/MpsPressure.cpp:932-1028 (relax_fact < 1) */
int pmps_prediction_wall_press_gradient(pmps_t* h); /* MpsPressure::predictionWallPressGradient src/MpsPressure.cpp:1031-1214 */
int pmps_predict(pmps_t* h);                     /* MpsParticleVelPos::predictionVelocityPosition  src/MpsParticleVelPos.cpp:20-50 */
int pmps_check_out_domain(pmps_t* h);            /* MpsParticle::checkParticleOutDomain     src/MpsParticle.cpp:234-402; periodic domain: the wrap (:236-269) first,
                                                  * and when it moved anybody (one 4-byte read-back) the bucket order is re-made before the call returns */
int pmps_collisions(pmps_t* h);                  /* MpsParticleCollision::checkParticleCollisions / ...Dynamic...  src/MpsParticleCollision.cpp:22-292 */
int pmps_wall_closest_point(pmps_t* h);          /* PolygonMesh::closestPointPNDBoundaryAABB (all meshes) + updateParticlesNearPolygonMesh  src/PolygonMesh.cpp:585-795 */
int pmps_wall_npcd(pmps_t* h);                   /* MpsPndNeigh::calcWallNPCD               src/MpsPndNeigh.cpp:105-195 */
int pmps_correction_matrix(pmps_t* h);           /* MpsVectorMatrix::correctionMatrix       src/MpsVectorMatrix.cpp:20-93 */
int pmps_pnd_neigh_npcd(pmps_t* h);              /* MpsPndNeigh::calcPndnNeighNPCD          src/MpsPndNeigh.cpp:199-331 */
int pmps_pnd_diffusive(pmps_t* h);               /* MpsPndNeigh::calcPndDiffusiveTerm       src/MpsPndNeigh.cpp:334-489 */
int pmps_wall_pnd_diffusive(pmps_t* h);          /* MpsPndNeigh::calcWall{Slip,NoSlip}PndDiffusiveTerm  src/MpsPndNeigh.cpp:492-744 */
int pmps_update_pnd(pmps_t* h);                  /* MpsPndNeigh::updatePnd                  src/MpsPndNeigh.cpp:747-774 */
int pmps_mean_pnd_wall_dummy_surface(pmps_t* h); /* MpsPndNeigh::meanPndParticlesWallDummySurface  src/MpsPndNeigh.cpp:777-852 */
int pmps_update_bc(pmps_t* h);                   /* MpsBoundaryCondition::updateParticleBC  src/MpsBoundaryCondition.cpp:66-399 */
int pmps_calc_normals(pmps_t* h);                   /* MpsShifting::calcNormalParticles, src/MpsShifting.cpp:106-174 (ARC free surface, src/main.cpp:351-357) */
int pmps_calc_wall_normals(pmps_t* h);              /* MpsShifting::calcWallNormalParticles, src/MpsShifting.cpp:319-429 */
int pmps_calc_press(pmps_t* h);                  /* MpsPressure::calcPress (EOS or PPE+Krylov)  src/MpsPressure.cpp:31-57 */
int pmps_extrapolate_press(pmps_t* h);           /* MpsPressure::extrapolatePressParticlesWallDummy  src/MpsPressure.cpp:792-846 */
int pmps_press_gradient(pmps_t* h);              /* MpsPressure::calcPressGradient          src/MpsPressure.cpp:1217-1338 */
int pmps_wall_press_gradient(pmps_t* h);         /* MpsPressure::calcWallPressGradient      src/MpsPressure.cpp:1341-1550 */
int pmps_wall_viscosity(pmps_t* h);              /* MpsViscosity::calcWall{Slip,NoSlip}Viscosity  src/MpsViscosity.cpp:1527-1910 */
int pmps_correct(pmps_t* h);                     /* MpsParticleVelPos::correctionVelocityPosition  src/MpsParticleVelPos.cpp:53-97 */
int pmps_shifting(pmps_t* h);                    /* MpsShifting::calcConcAndConcGradient (+ calcShifting for type 1)  src/MpsShifting.cpp:23-101,432-575 */
int pmps_wall_shifting(pmps_t* h);               /* MpsShifting::calcWallConcAndConcGradient  src/MpsShifting.cpp:578-691 */
int pmps_update_wall_dummy_vel(pmps_t* h);       /* MpsParticleVelPos::updateVelocityParticlesWallDummy  src/MpsParticleVelPos.cpp:100-158 */
/* 0-th step passes (src/main.cpp:157-180) */
int pmps_initial_pnd(pmps_t* h);                 /* MpsPndNeigh::setInitialPndNumberOfNeighNPCD  src/MpsPndNeigh.cpp:21-101 */
int pmps_mean_wall_pnd(pmps_t* h);               /* MpsPndNeigh::meanWallPnd                src/MpsPndNeigh.cpp:923-994 */
int pmps_mean_pnd(pmps_t* h);                    /* MpsPndNeigh::meanPnd                    src/MpsPndNeigh.cpp:855-920 */
int pmps_mean_neigh_fluid_pnd(pmps_t* h);        /* MpsPndNeigh::meanNeighFluidPnd          src/MpsPndNeigh.cpp:997-1049 */

/* ---- fused fast path ---------------------------------------------------------------------------------------- */
int pmps_init_step0(pmps_t* h);                  /* the whole 0-th step, src/main.cpp:121-180 (buckets, wall pass, initial PND, BC, pressure) */
int pmps_step(pmps_t* h, int nsteps);            /* nsteps x the loop body of src/main.cpp:256-460, state stays on device */

/* ---- slab decomposition over the GPUs of one box ----------------------------------------------
 * The reference is a single-process OpenMP code (no domain decomposition to mirror); these calls are what a
 * distributed host (polymps_b200/slab.py: one process per GPU, NCCL send/recv between neighbouring slabs) needs on
 * top of the calls above.  A rank uploads a particle table whose host indices [0, nOwned) are the particles inside its
 * slab and [nOwned, numParticles) are halo copies of the neighbours' particles within bucketSide of its cuts, then runs
 * the fused step phase by phase: after each phase it refreshes the halo copies of the fields that phase produced.
 * velMax (src/MpsParticleVelPos.cpp:91), the displacement since the bucket rebuild and the out-of-domain test
 * (src/MpsParticle.cpp:234-402) count owned particles only; pmps_slab_collect / pmps_slab_apply let the caller
 * reduce them over the ranks so that every rank takes the same rebuild decision and uses the same velMax. */
enum { PMPS_HF_POS = 1, PMPS_HF_VEL = 2, PMPS_HF_PRESS = 4, PMPS_HF_PNDI = 8, PMPS_HF_CONC = 16, PMPS_HF_PNDKI = 32,
       PMPS_HF_BC = 64, PMPS_HF_PPE_X = 128 /* pressurePPE: the CG iterate / warm start */, PMPS_HF_PPE_P = 256 /* the CG search direction */ };
int pmps_halo_planes(unsigned fields);           /* doubles per particle of a packed buffer (pos carries {x,y,z,type bits}) */
int pmps_slab_begin(pmps_t* h, int nOwned);      /* after each (re-)upload of the rank's particle table */
int pmps_halo_pack(pmps_t* h, unsigned fields, const int32_t* d_hostIndex, int n, double* d_out);   /* device pointers; plane-major */
int pmps_halo_unpack(pmps_t* h, unsigned fields, int firstHostIndex, int n, const double* d_in);    /* onto host indices [first, first+n) */
int pmps_step_phases(void);                      /* number of phases of the fused step */
int pmps_step_phase(pmps_t* h, int phase, unsigned* haloFieldsAfter);   /* one phase of the loop body of src/main.cpp:256-460 */
int pmps_slab_collect(pmps_t* h, double* d_out4);        /* device, FOUR doubles: {max displacement^2 since rebuild, max shift^2, velMax^2, displacement^2 the step before ended with} of this rank */
int pmps_slab_apply(pmps_t* h, const double* d_in3);     /* the first three after the MAX all-reduce: ends the step */
/* Rebuild decision of a slab step, taken where a single GPU takes it: right after the prediction (phase 0 and its halo refresh), from
 * the four scalars of pmps_slab_collect MAX-all-reduced over the ranks — so every rank answers alike.  1: re-partition now (migrants,
 * new halo copies, pmps_upload_device of the new table), then pmps_slab_mid_rebuild (bucket order + hit lists at the predicted
 * positions), then go on with phase 1.  A driver that never calls it gets the step-start predictor of pmps_slab_wants_rebuild. */
int pmps_slab_mid_check(pmps_t* h, const double* d_in4);
int pmps_slab_mid_rebuild(pmps_t* h);
int pmps_slab_wants_rebuild(pmps_t* h);
/* Incompressible scheme on slabs: the conjugate-gradient solve (src/MpsPressure.cpp:424-445; Eigen ConjugateGradient.h:27-91) sums
 * its dot products over OWNED rows and calls back into the host at the two points where ranks must talk: `sum(user, n)` must
 * SUM-all-reduce, in place, the first n doubles of the device workspace given here; `refresh(user, fields)` must refresh the halo
 * copies of `fields` (PMPS_HF_PPE_P before every matrix-vector product).  Both run on pmps_stream(h). */
typedef void (*pmps_sum_fn)(void* user, int n);
typedef void (*pmps_refresh_fn)(void* user, unsigned fields);
int pmps_slab_set_hooks(pmps_t* h, double* d_workspace8, pmps_sum_fn sum, pmps_refresh_fn refresh, void* user);                  /* 1: re-partition (migrate + new halo) and re-upload before the next step */

/* ---- integrated observables on the device (north_star: surge-front position and pressure-probe history of a dam break) -----------
 * The reference samples on the host: its pressure sensors average the pressure of the particles inside a window
 * (src/MpsInputOutput.cpp:1937-2229) and every output step writes all arrays (:861-1875).  These two calls reduce on the device —
 * one pass over the position records, two-stage reductions in a fixed order — so a run records its history without a particle
 * array leaving the GPU.  On a slab rank they cover the OWNED particles; the caller combines the ranks (MIN / MAX / SUM).
 * pmps_observe_extent: bounding box and count of the particles of one type (surge front = mx[0] of the fluid);
 * pmps_observe_box: count, pressure sum and pressure maximum of the particles of one type inside [lo, hi] (probe = sum / count). */
int pmps_observe_extent(pmps_t* h, int particleType, double mn[3], double mx[3], long long* count);
int pmps_observe_box(pmps_t* h, int particleType, const double lo[3], const double hi[3], long long* count, double* pressSum, double* pressMax);
/* pmps_write_vtu: the particle file of an output step, written straight from the device state — the same bytes MpsInputOutput::
 * writeVtuAscii (flags without PMPS_VTU_FREE_SURFACE_ONLY, src/MpsInputOutput.cpp:1313-1580) or writeVtuAsciiFreeSurface (with it,
 * :1582-1875: free-surface and wall particles only) puts out for the same state.  flags = the reference's output switches
 * (flags.output_VTU of the input file, src/MpsInputOutput.cpp:150-170).  Only the arrays the file holds come back from the device;
 * the text is formatted by all host cores (PMPS_VTU_THREADS overrides the count).  The binary variants of the reference
 * ("ONLY ASCII WORKING", its own input files say) are not provided. */
enum { PMPS_VTU_PND = 1, PMPS_VTU_NEIGH = 2, PMPS_VTU_DEVIATION = 4, PMPS_VTU_CONCENTRATION = 8, PMPS_VTU_NONNEWTONIAN = 16,
       PMPS_VTU_AUXILIAR = 32, PMPS_VTU_FREE_SURFACE_ONLY = 64 };
int pmps_write_vtu(pmps_t* h, const char* filename, unsigned flags);
/* pmps_observe_extent_static: same arguments and results as pmps_observe_extent.  No counterpart in the reference; added
 * for callers that may not allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  The call allocates nothing: the per-block partial records of the reduction live in a file-scope device array of
 * 1024 records (80 KiB, held by every GPU that loads the library until exit) instead of the handle's staging area.  One record per
 * block, four blocks per SM: a device that needs more than 1024 records (> 256 SMs) is refused with PMPS_E_ARG and a message in
 * pmps_last_error.  Every call on one GPU shares the array (the previous call's contents are overwritten before they are read,
 * nothing has to be cleared), so calls that overlap on the same GPU — other handles, other streams — race: serialise them. */
int pmps_observe_extent_static(pmps_t* h, int particleType, double mn[3], double mx[3], long long* count);

/* ---- debug / parity hooks (tests only) ---------------------------------------------------------------------- */
/* "%f " of every value with the formatter pmps_write_vtu uses (host code, no device needed); returns the length written */
int pmps_debug_format_f6(const double* v, int n, char* out, size_t cap);
/* bucket id of every host-order particle as the reference computes it at rebuild time, and the members of every
 * bucket in list order (= ascending host index): flattened CSR. members may be NULL to query sizes. */
int pmps_debug_buckets(pmps_t* h, int32_t* bucketOfParticle, int32_t* bucketStart /* nB+1 */, int32_t* members /* N */);
/* neighbour set of host-order particle i under predicate kind: 0 = P(reL), 1 = P(reS), 2 = M(reS), 3 = M(reL),
 * 4 = collision; returns count, fills out[0..cap) with host-order ids in traversal order. */
int pmps_debug_neighbors(pmps_t* h, int i, int kind, int32_t* out, int cap);
/* Stale bucket lists after an out-of-domain swap.  checkParticleOutDomain copies the last live particle over the one that left
 * (src/MpsParticle.cpp:298-396) but the reference re-threads its bucket lists only at the next step start (src/main.cpp:256): for the
 * rest of the step the moved particle is found by a neighbour loop only from the buckets around the one the particle that left was
 * listed in.  A single-GPU handle replays that (default on) so that results equal the reference's; off = every particle is found
 * from where it is (what a slab rank does, whose table order means nothing across ranks). */
int pmps_set_stale_list_replay(pmps_t* h, int on);
/* tests: declare array elements (host indices) as listed in the given reference buckets (iz*nXY + iy*nX + ix) until the next
 * pmps_update_buckets / step start, as if an out-of-domain swap of this step had moved them there */
int pmps_debug_set_stale(pmps_t* h, int n, const int32_t* hostIdx, const int32_t* bucket);
/* per-kernel device time of the last pmps_step (CUDA events on the step stream), name list is static */
int pmps_profile_enable(pmps_t* h, int on);
int pmps_profile_get(pmps_t* h, int* n, const char*** names, const float** ms_total, const int** launches);
int pmps_launch_count(pmps_t* h, long long* launches); /* kernels launched by this handle so far */
/* work counters of the temporal caches (bench.py reports their deltas over the timed region): out[0] bucket-order rebuilds (sort +
 * Verlet list; the reference does src/MpsBucket.cpp:299-353 every step), out[1] hit-list refilters, out[2] fused steps, out[3] 1 if
 * a step ever ended with a particle beyond the search slack (sticky until the next pmps_upload) */
int pmps_get_counters(pmps_t* h, long long out[4]);
void* pmps_stream(pmps_t* h);                    /* cudaStream_t the stages run on */
int pmps_synchronize(pmps_t* h);                 /* wait for everything queued on the handle's stream */

#ifdef __cplusplus
}
#endif
#endif /* PMPS_H_ */
