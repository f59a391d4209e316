/* oracle/mps_oracle.h — CPU restatement of the PolyMPS per-timestep hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load this library; the product
 * (libpmps.so, polymps_b200/) never links, imports or executes anything under oracle/.
 *
 * Parity status: PINNED.  Every stage function below is checked bit-for-bit (tests/test_oracle_vs_reference.py,
 * tests/test_oracle_golden.py) against raw array dumps produced by the UNMODIFIED reference objects
 * (oracle/ref_driver.cpp linked with /root/reference/src/[all except main].cpp, OMP_NUM_THREADS=1); the dumps for the small cases
 * are committed under tests/golden/ together with the script that made them.  The reference owns no tests or
 * golden vectors of its own, so "outputs of the reference itself run here" is the pin.
 *
 * The structs are the product's C-ABI PODs (include/pmps.h) so that the same host arrays feed both sides.
 */
#ifndef MPS_ORACLE_H_
#define MPS_ORACLE_H_
#include "../include/pmps.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct OBuckets {          /* MpsParticle bucket arrays, include/MpsParticle.h:139-143 */
	int32_t* first; int32_t* last; int32_t* next;   /* nB, nB, numParticlesZero */
} OBuckets;

typedef struct OMesh {             /* one polygon mesh + the libigl AABB hierarchy in heap order */
	int32_t nV, nF, nNodes, nTable;
	const double* V; const int32_t* F; const double* N;
	double* bbMin; double* bbMax; int32_t* prim;   /* nNodes*3, nNodes*3, nNodes (-1 inner, -2 absent) — filled by oracle_build_tree */
	const double* xDataPND; const double* xDataNeigh; const double* Zij; const int32_t* nNeighWall;
} OMesh;

int  oracle_tree_nodes(int nF);                                   /* capacity needed for bbMin/bbMax/prim */
void oracle_build_tree(OMesh* m);                                 /* igl::AABB::init, libigl AABB.cpp:29-196 */
void oracle_closest_point(const OMesh* m, const double p[3], double* sqrD, int* face, double c[3]);

void oracle_update_buckets(const PmpsParams* P, PmpsHostArrays* A, OBuckets* B);          /* src/MpsBucket.cpp:299-353 */
void oracle_calc_viscosity(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B);     /* src/MpsViscosity.cpp:20-100 */
void oracle_calc_volume_fraction(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B);   /* src/MpsViscosity.cpp:104-190 */
double oracle_calc_viscosity_interaction(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B); /* src/MpsViscosity.cpp:194-700, returns CFLvisc */
void oracle_calc_wall_viscosity_interaction(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B); /* src/MpsViscosity.cpp:703-1526 (by slipCondition) */
void oracle_add_gravity(const PmpsParams* P, PmpsHostArrays* A);                           /* src/MpsParticle.cpp:111-130 */
void oracle_predict(const PmpsParams* P, PmpsHostArrays* A);                               /* src/MpsParticleVelPos.cpp:20-50 */
int  oracle_check_out_domain(const PmpsParams* P, PmpsHostArrays* A);                      /* src/MpsParticle.cpp:234-402 */
void oracle_collisions(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B);         /* src/MpsParticleCollision.cpp:22-119 */
void oracle_wall_closest_point(const PmpsParams* P, PmpsHostArrays* A, const OMesh* meshes); /* src/PolygonMesh.cpp:585-795 */
void oracle_wall_npcd(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B);          /* src/MpsPndNeigh.cpp:105-195 */
void oracle_correction_matrix(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B);  /* src/MpsVectorMatrix.cpp:20-166 */
void oracle_pnd_neigh_npcd(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B);     /* src/MpsPndNeigh.cpp:199-331 */
void oracle_pnd_diffusive(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B);      /* src/MpsPndNeigh.cpp:334-489 */
void oracle_wall_pnd_diffusive(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B); /* src/MpsPndNeigh.cpp:492-744 */
void oracle_vel_divergence(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B);       /* src/MpsPressure.cpp:479-551 */
void oracle_wall_vel_divergence(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B);  /* src/MpsPressure.cpp:554-788 */
void oracle_update_pnd(const PmpsParams* P, PmpsHostArrays* A);                            /* src/MpsPndNeigh.cpp:747-774 */
void oracle_mean_pnd_wall_dummy_surface(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B); /* src/MpsPndNeigh.cpp:777-852 */
void oracle_update_bc(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B);          /* src/MpsBoundaryCondition.cpp:66-399 */
void oracle_calc_normals(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B);       /* src/MpsShifting.cpp:106-174 */
void oracle_calc_wall_normals(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B);  /* src/MpsShifting.cpp:319-429 */
void oracle_calc_press_eos(const PmpsParams* P, PmpsHostArrays* A);                        /* src/MpsPressure.cpp:62-163 */
int  oracle_solve_ppe(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B, double* err); /* src/MpsPressure.cpp:169-286,424-476 */
void oracle_extrapolate_press(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B);  /* src/MpsPressure.cpp:792-846 */
void oracle_press_gradient(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B);     /* src/MpsPressure.cpp:1217-1338 */
void oracle_wall_press_gradient(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B, double velMax); /* src/MpsPressure.cpp:1341-1611 */
void oracle_prediction_press_gradient(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B);          /* src/MpsPressure.cpp:932-1028 */
void oracle_prediction_wall_press_gradient(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B, double velMax); /* src/MpsPressure.cpp:1031-1214 */
void oracle_wall_viscosity(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B);     /* src/MpsViscosity.cpp:1527-1910 */
double oracle_correct(const PmpsParams* P, PmpsHostArrays* A);                             /* src/MpsParticleVelPos.cpp:53-97, returns velMax */
void oracle_shifting(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B);           /* src/MpsShifting.cpp:432-575 */
void oracle_wall_shifting(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B);      /* src/MpsShifting.cpp:578-691 */
void oracle_update_wall_dummy_vel(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B); /* src/MpsParticleVelPos.cpp:100-158 */
void oracle_initial_pnd(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B);        /* src/MpsPndNeigh.cpp:21-101 */
void oracle_mean_wall_pnd(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B);      /* src/MpsPndNeigh.cpp:923-994 */
void oracle_mean_pnd(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B);           /* src/MpsPndNeigh.cpp:855-920 */
void oracle_mean_neigh_fluid_pnd(const PmpsParams* P, PmpsHostArrays* A, const OBuckets* B); /* src/MpsPndNeigh.cpp:997-1049 */
/* neighbour set of i under predicate kind (same numbering as pmps_debug_neighbors), traversal order */
int  oracle_neighbors(const PmpsParams* P, const PmpsHostArrays* A, const OBuckets* B, int i, int kind, int32_t* out, int cap);
/* whole 0-th step / whole time step in the reference's order (src/main.cpp:121-180, :256-460) */
void oracle_init_step0(PmpsParams* P, PmpsHostArrays* A, OBuckets* B, const OMesh* meshes);
void oracle_step(PmpsParams* P, PmpsHostArrays* A, OBuckets* B, const OMesh* meshes, int* solverIter, double* solverErr);
void oracle_set_threads(int n);

#ifdef __cplusplus
}
#endif
#endif
