// pmps_walk_g.cu — one pack group of walker kernels for one option set: compiled 4 x (option sets) times with -DPMPS_WALK_GROUP=k
// [-DPMPS_OPTSET_USE_WC3D | -DPMPS_OPTSET_USE_INC3D] (Makefile)
#include "pmps_walk_inst.inc"
