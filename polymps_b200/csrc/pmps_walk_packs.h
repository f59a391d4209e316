// pmps_walk_packs.h — the packs of stage bodies that run as one walker launch: X(id, group, bodies...).  Every pack is two kernels
// (polygon / particle walls) per option set and the front end spends most of the build on them, so the packs are spread over four
// translation units per option set (group = which one; balanced by compile time) that build side by side.
#pragma once
#define PMPS_WALK_PACKS(X) \
	X(0, 0, BViscosity, BPredictTail) X(1, 0, BViscosity) X(2, 0, BCollision) X(3, 0, BCollision, BCollisionTail) X(4, 0, BCollisionDPC) X(5, 0, BCollisionDPC, BCollisionDPCTail) \
	X(6, 0, BWallNpcd) X(7, 0, BInitialPnd) X(8, 0, BShiftVel) \
	X(9, 1, BPndNeighNpcd, BPndDiffusive) X(10, 1, BPndNeighNpcd) X(11, 1, BPndDiffusive) X(12, 1, BWallPndDiffusive) X(13, 1, BWallVelDivergence) X(14, 1, BVelDivergence) \
	X(15, 1, BMeanPnd<0>) X(16, 1, BMeanPnd<1>) X(17, 1, BMeanWallPnd) \
	X(18, 2, BPressGradient) X(19, 2, BPressGradient, BCorrectTail) X(20, 2, BPredPressGradient) X(21, 2, BPredWallPressGradient) X(22, 2, BWallPressGradient) \
	X(23, 2, BWallPressGradient, BWallViscosity, BCorrectTail) X(24, 2, BWallViscosity) X(25, 2, BPressMin) X(26, 2, BMeanNeighFluidPnd) \
	X(27, 3, BCorrectionMatrix) X(28, 3, BCorrectionMatrix, BPressMin) X(29, 3, BExtrapolatePress) X(30, 3, BConcentration) X(31, 3, BGradConcentration) X(32, 3, BWallGradConcentration) \
	X(33, 3, BWallDummyVel) X(34, 3, BVolumeFraction) X(35, 3, BStrainRate) \
	X(36, 0, BNormal) X(37, 1, BWallNormal) X(38, 2, BArcBC) X(39, 3, BWallStrainRate)
#define PMPS_WALK_GROUPS 4
