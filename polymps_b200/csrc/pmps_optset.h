// pmps_optset.h — option sets of the walker kernels.
//
// The stage bodies branch on a dozen option enums of the input file (weight function, gradient form, correction matrix on/off,
// dimension, slip condition, ...).  Read from the parameter block they cost a constant load, a compare and a branch each, per pair,
// and they cut the unrolled pair loop into small basic blocks.  For the option sets of the named configurations the walker
// translation units are compiled a second time inside a namespace in which `KParams::p` is a struct DERIVED from PmpsParams whose
// static constexpr members HIDE the run-time fields of the same name: the bodies' `K.p.weightType` then is a literal, the dead
// branches vanish, and nothing in the bodies changes.  Same size and layout (static members occupy no storage), so the parameter
// block of a handle is reinterpreted; pmps_create picks the set whose constants equal the handle's parameters, the generic kernels
// otherwise (measured, 10 M particles at rest: 24.95 -> 22.55 ms per step, F3 -18 %, F5 -20 %).
#pragma once
#include "pmps_internal.h"

// X(field, value): InputDamWC / dam1610 WC option set in 3D (weight re/r - 1, gradient type 3 without correction, no-slip, diffusive PND,
// PC collisions, Lennard-Jones wall repulsion) with the reference's default particle / boundary ids
#define PMPS_OPTSET_WC3D(X) X(int32_t, weightType, 0) X(int32_t, gradientCorrection, 0) X(int32_t, gradientType, 3) X(int32_t, dim, 3) X(int32_t, slipCondition, 0) \
	X(int32_t, pndType, 2) X(int32_t, collisionType, 0) X(int32_t, repulsiveForceType, 2) X(int32_t, mpsType, 1) X(int32_t, fluidType, 0) X(int32_t, domainTypeBC, 0) \
	X(int32_t, ghost, -1) X(int32_t, fluid, 0) X(int32_t, wall, 2) X(int32_t, dummyWall, 3) X(int32_t, surface, 1) X(int32_t, inner, 0) X(int32_t, other, -1)
// InputDamINC option set in 3D (gradient type 0 with the correction matrix, PND by sum, implicit pressure)
#define PMPS_OPTSET_INC3D(X) X(int32_t, weightType, 0) X(int32_t, gradientCorrection, 1) X(int32_t, gradientType, 0) X(int32_t, dim, 3) X(int32_t, slipCondition, 0) \
	X(int32_t, pndType, 0) X(int32_t, collisionType, 0) X(int32_t, repulsiveForceType, 2) X(int32_t, mpsType, 2) X(int32_t, fluidType, 0) X(int32_t, domainTypeBC, 0) \
	X(int32_t, ghost, -1) X(int32_t, fluid, 0) X(int32_t, wall, 2) X(int32_t, dummyWall, 3) X(int32_t, surface, 1) X(int32_t, inner, 0) X(int32_t, other, -1)

enum { PMPS_OPT_GENERIC = 0, PMPS_OPT_WC3D = 1, PMPS_OPT_INC3D = 2, PMPS_OPT_SETS = 3 };
#define PMPS_OPT_MATCH(T, n, v) && P.n == (v)
static inline int pmps_optset_of(const PmpsParams& P) {
	if (true PMPS_OPTSET_WC3D(PMPS_OPT_MATCH)) return PMPS_OPT_WC3D;
	if (true PMPS_OPTSET_INC3D(PMPS_OPT_MATCH)) return PMPS_OPT_INC3D;
	return PMPS_OPT_GENERIC;
}
// one neighbour pass by pack id (pmps_walk.cuh, PMPS_WALK_PACKS) with the kernels of option set `optset`
int pmps_walk_dispatch(int optset, int packId, int group, pmps_handle* h, const char* name, int polyOverride);
