"""Host-side mirror of the reference's MpsParticleSystem (include/MpsParticleSystem.h:133-309).

`ParticleSystem.from_json` follows the key names/defaults of MpsInputOutput::readInputFile
(src/MpsInputOutput.cpp:134-505); `set_parameters` restates MpsParticle::setParameters
(src/MpsParticle.cpp:136-227) and `allocate_buckets` restates MpsBucket::allocateMemory /
calcDomainLimits (src/MpsBucket.cpp:20-140).  All arithmetic is plain Python float (IEEE double) in the
reference's evaluation order so the derived constants are bit-identical (tests/test_params.py pins them
against values dumped from the reference itself).  Field names are the reference's.
"""
from __future__ import annotations

import ctypes
import json
import math
import re
import sys
from dataclasses import dataclass, field

EPS_ZERO = 10.0 * sys.float_info.epsilon  # include/MpsParticleSystem.h:241


def _weight(dst: float, re_: float, wtype: int) -> float:
    """MpsParticle::weight / weightGradient (src/MpsParticle.cpp:57-91; both tables are identical)."""
    if wtype == 1:
        return re_ / dst + dst / re_ - 2.0
    if wtype == 2:
        return re_ / dst - dst / re_
    if wtype == 3:
        return (1.0 - dst / re_) * (1.0 - dst / re_) * (1.0 - dst / re_)
    if wtype == 4:
        return (1.0 - dst / re_) * (1.0 - dst / re_)
    return re_ / dst - 1.0


def _strip_json_comments(text: str) -> str:
    # nlohmann is called with skip_comments=true (src/MpsInputOutput.cpp:126-129)
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return re.sub(r"(?m)^\s*//.*$", "", text)


@dataclass
class ParticleSystem:
    # ---- JSON inputs (defaults = the .value(key, default) defaults of the reference) ----
    wallType: int = 1
    femOn: bool = False
    forcedOn: bool = False
    domainMinX: float = 0.0
    domainMinY: float = 0.0
    domainMinZ: float = 0.0
    domainMaxX: float = 0.0
    domainMaxY: float = 0.0
    domainMaxZ: float = 0.0
    domainTypeBC: int = 0
    limitTypeBC: int = 0
    numBC: int = 1
    periodicDirectionX: bool = False
    periodicDirectionY: bool = False
    periodicDirectionZ: bool = False
    densityFluid: float = 1000.0
    densityWall: float = 1000.0
    KNM_VS1: float = 0.000001
    fluidType: int = 0
    gravityX: float = 0.0
    gravityY: float = 0.0
    gravityZ: float = -9.81
    KNM_VS2: float = 0.000001
    DNS_SDT: float = 1540.0
    MEU0: float = 0.03
    N: float = 1.2                # power_law_index (src/MpsInputOutput.cpp:272)
    PHI_1: float = 0.541
    PHI_2: float = 0.6
    cohes: float = 0.0
    fractionMethod: int = 2
    DG: float = 0.0035
    I0: float = 0.75
    mm: float = 100.0
    Cd: float = 0.47
    VF_min: float = 0.25
    VF_max: float = 0.65
    DNS_FL1: float = 1000.0
    DNS_FL2: float = 1540.0
    dim: float = 3.0
    partDist: float = 0.01
    timeStep: float = 0.0005
    timeSimulation: float = 1.0
    iterOutput: int = 80
    cflNumber: float = 0.2
    weightType: int = 0
    slipCondition: int = 0
    reS: float = 2.1
    reL: float = 2.1
    gradientType: int = 0
    gradientCorrection: bool = False
    relaxPress: float = 1.0
    mpsType: int = 1
    soundSpeed: float = 15.0
    gamma: float = 7.0
    solverType: int = 0
    alphaCompressibility: float = 0.000001
    relaxPND: float = 0.001
    shiftingType: float = 2
    dri: float = 0.01
    coefA: float = 2.0
    machNumber: float = 0.1
    pndType: int = 0
    diffusiveCoef: float = 0.35
    repulsiveForceType: int = 2
    reRepulsiveForce: float = 0.5
    expectMaxVelocity: float = 6.0
    repForceCoefMitsume: float = 40000000.0
    repForceCoefLennardJones: float = 2.0
    repForceCoefMonaghanKajtar: float = 1.0
    EPS_RE: float = 0.01
    freeSurfType: int = 0
    pndThreshold: float = 0.98
    neighThreshold: float = 0.85
    npcdThreshold: float = 0.20
    thetaThreshold: float = 45.0
    normThreshold: float = 0.1
    collisionType: int = 0
    collisionRatio: float = 0.20
    distLimitRatio: float = 0.85
    lambdaCollision: float = 0.20
    ghost: int = -1
    fluid: int = 0
    wall: int = 2
    dummyWall: int = 3
    surface: int = 1
    inner: int = 0
    other: int = -1
    # paths (MpsInputOutput members)
    gridFilename: str = "oops"
    meshRigidFilename: str = "oops"
    meshDeformableFilename: str = "oops"
    meshForcedFilename: str = "oops"
    vtuOutputFoldername: str = "oops"
    # ---- derived (set_parameters / allocate_buckets) ----
    epsilonZero: float = EPS_ZERO
    numOfMeshs: int = 0
    reS2: float = 0.0
    reL2: float = 0.0
    pndSmallZero: float = 0.0
    pndLargeZero: float = 0.0
    pndGradientZero: float = 0.0
    lambdaZero: float = 0.0
    numNeighZero: float = 0.0
    coeffViscosity: float = 0.0
    coeffViscMultiphase: float = 0.0
    coeffPressEMPS: float = 0.0
    coeffPressGrad: float = 0.0
    coeffPressWCMPS: float = 0.0
    coeffShifting1: float = 0.0
    coeffShifting2: float = 0.0
    coeffPPE: float = 0.0
    coeffPPESource: float = 0.0
    invPartDist: float = 0.0
    distCollisionLimit: float = 0.0
    distCollisionLimit2: float = 0.0
    restitutionCollision: float = 0.0
    betaPnd: float = 0.0
    betaNeigh: float = 0.0
    delta2: float = 0.0
    thetaArc: float = 0.0
    hThreshold2: float = 0.0
    dstThreshold2: float = 0.0
    normThreshold2: float = 0.0
    velMax: float = 0.0
    CFLcurrent: float = 0.0
    timeCurrent: float = 0.0
    numOfIterations: int = 0
    bucketSide: float = 0.0
    invBucketSide: float = 0.0
    numBucketsX: int = 0
    numBucketsY: int = 0
    numBucketsZ: int = 0
    numBucketsXY: int = 0
    numBucketsXYZ: int = 0
    periodicLength: list = field(default_factory=lambda: [0.0, 0.0, 0.0])
    uniformVelWall: list = field(default_factory=lambda: [0.0, 0.0, 0.0])
    _parameters_set: bool = False

    # ------------------------------------------------------------------ JSON
    @classmethod
    def from_json(cls, path_or_dict) -> "ParticleSystem":
        """Key names and defaults of MpsInputOutput::readInputFile (src/MpsInputOutput.cpp:134-505)."""
        if isinstance(path_or_dict, dict):
            je = path_or_dict
        else:
            with open(path_or_dict, "r") as fh:
                je = json.loads(_strip_json_comments(fh.read()))
        ps = cls()

        def g(d, key, default):
            return d.get(key, default) if isinstance(d, dict) else default

        fl = je.get("flags", {})
        ps.wallType = int(g(fl, "wall_type", 1))
        ps.femOn = bool(g(fl, "fem_MESH", False))
        ps.forcedOn = bool(g(fl, "forced_MESH", False))
        pn = je.get("pathNames", {})
        ps.gridFilename = g(pn, "particle_grid_file", "oops")
        ps.meshRigidFilename = g(pn, "mesh_rigid_file", "oops")
        ps.meshDeformableFilename = g(pn, "mesh_deformable_file", "oops")
        ps.meshForcedFilename = g(pn, "mesh_forced_file", "oops")
        ps.vtuOutputFoldername = g(pn, "vtu_output_folder", "oops")
        dm = je.get("domain", {})
        ps.domainMinX = float(g(dm.get("min", {}), "x", 0.0))
        ps.domainMinY = float(g(dm.get("min", {}), "y", 0.0))
        ps.domainMinZ = float(g(dm.get("min", {}), "z", 0.0))
        ps.domainMaxX = float(g(dm.get("max", {}), "x", 0.0))
        ps.domainMaxY = float(g(dm.get("max", {}), "y", 0.0))
        ps.domainMaxZ = float(g(dm.get("max", {}), "z", 0.0))
        bd = dm.get("boundary", {})
        ps.domainTypeBC = int(g(bd, "type", 0))
        ps.limitTypeBC = int(g(bd, "limit", 0))
        ps.numBC = 1
        di = bd.get("direction", {})
        ps.periodicDirectionX = bool(g(di, "x", False))
        ps.periodicDirectionY = bool(g(di, "y", False))
        ps.periodicDirectionZ = bool(g(di, "z", False))
        ph = je.get("physical", {})
        ps.densityFluid = float(g(ph, "fluid_density", 1000.0))
        ps.densityWall = float(g(ph, "wall_density", 1000.0))
        ps.KNM_VS1 = float(g(ph, "kinematic_visc", 0.000001))
        ps.fluidType = int(g(ph, "fluid_type", 0))
        gr = ph.get("gravity", {})
        ps.gravityX = float(g(gr, "x", 0.0))
        ps.gravityY = float(g(gr, "y", 0.0))
        ps.gravityZ = float(g(gr, "z", -9.81))
        rh = ph.get("rheological", {})
        ps.KNM_VS2 = float(g(rh, "kinematic_visc_phase_2", 0.000001))
        ps.DNS_FL1 = float(g(rh, "fluid_density_phase_1", 1000.0))
        ps.DNS_FL2 = float(g(rh, "fluid_density_phase_2", 1540.0))
        ps.DNS_SDT = float(g(rh, "sediment_density", 1540.0))
        ps.MEU0 = float(g(rh, "consistency_index", 0.03))
        ps.N = float(g(rh, "power_law_index", 1.2))
        phi = rh.get("phi", {})
        ps.PHI_1 = float(g(phi, "lower", 0.541)); ps.PHI_2 = float(g(phi, "second", 0.6))
        ps.cohes = float(g(rh, "cohes_coeff", 0.0))
        ps.fractionMethod = int(g(rh, "fraction_method", 2))
        ps.DG = float(g(rh, "grain_size", 0.0035)); ps.I0 = float(g(rh, "I0", 0.75)); ps.mm = float(g(rh, "mm", 100.0))
        ps.Cd = float(g(rh, "drag_coeff", 0.47))
        vf = rh.get("volume_fraction", {})
        ps.VF_min = float(g(vf, "min", 0.25)); ps.VF_max = float(g(vf, "max", 0.65))
        nu = je.get("numerical", {})
        ps.dim = float(g(nu, "dimension", 3.0))
        ps.partDist = float(g(nu, "particle_dist", 0.01))
        ps.timeStep = float(g(nu, "time_step", 0.0005))
        ps.timeSimulation = float(g(nu, "final_time", 1.0))
        ps.iterOutput = int(g(nu, "iter_output", 80))
        ps.cflNumber = float(g(nu, "CFL_number", 0.2))
        ps.weightType = int(g(nu, "weight_type", 0))
        ps.slipCondition = int(g(nu, "slip_condition", 0))
        er = nu.get("effective_radius", {})
        ps.reS = float(g(er, "small", 2.1))
        ps.reL = float(g(er, "large", 2.1))
        gd = nu.get("gradient", {})
        ps.gradientType = int(g(gd, "type", 0))
        ps.gradientCorrection = bool(g(gd, "correction", False))
        ps.relaxPress = float(g(gd, "relax_fact", 1.0))
        ps.mpsType = int(g(nu, "mps_type", 1))
        eq = nu.get("explicit_mps", {}).get("equation_state", {})
        ps.soundSpeed = float(g(eq, "speed_sound", 15.0))
        ps.gamma = float(g(eq, "gamma", 7.0))
        si = nu.get("semi_implicit_mps", {})
        ps.solverType = int(g(si, "solver_type", 0))
        ps.alphaCompressibility = float(g(si.get("weak_compressibility", {}), "alpha", 0.000001))
        ps.relaxPND = float(g(si.get("source_term", {}), "relax_pnd", 0.001))
        sh = nu.get("particle_shifting", {})
        ps.shiftingType = float(g(sh, "type", 2))
        ps.dri = float(g(sh, "DRI", 0.01))
        ps.coefA = float(g(sh, "coef_A", 2.0))
        ps.machNumber = float(g(sh, "mach_number", 0.1))
        pd = nu.get("pnd", {})
        ps.pndType = int(g(pd, "type", 0))
        ps.diffusiveCoef = float(g(pd, "diffusive_coeff", 0.35))
        wf = nu.get("wall_repulsive_force", {})
        ps.repulsiveForceType = int(g(wf, "type", 2))
        ps.reRepulsiveForce = float(g(wf, "re", 0.5))
        ps.expectMaxVelocity = float(g(wf, "maxVel", 6.0))
        co = wf.get("coefficient", {})
        ps.repForceCoefMitsume = float(g(co, "Mitsume", 40000000.0))
        ps.repForceCoefLennardJones = float(g(co, "Lennard-Jones", 2.0))
        ps.repForceCoefMonaghanKajtar = float(g(co, "Monaghan-Kajtar", 1.0))
        ps.EPS_RE = float(g(wf, "eps_re", 0.01))
        fs = nu.get("free_surface_threshold", {})
        ps.freeSurfType = int(g(fs, "type", 0))
        ps.pndThreshold = float(g(fs, "pnd", 0.98))
        ps.neighThreshold = float(g(fs, "neigh", 0.85))
        ps.npcdThreshold = float(g(fs, "NPCD", 0.20))
        ps.thetaThreshold = float(g(fs, "ARC", 45.0))
        ps.normThreshold = float(g(fs, "normal", 0.1))
        pc = nu.get("particle_collision", {})
        ps.collisionType = int(g(pc, "type", 0))
        ps.collisionRatio = float(g(pc, "ratio", 0.20))
        ps.distLimitRatio = float(g(pc, "dist_limit_ratio", 0.85))
        ps.lambdaCollision = float(g(pc, "lambda", 0.20))
        pt = nu.get("particle_type", {})
        ps.ghost = int(g(pt, "ghost", -1))
        ps.fluid = int(g(pt, "fluid", 0))
        ps.wall = int(g(pt, "wall", 2))
        ps.dummyWall = int(g(pt, "dummyWall", 3))
        bt = nu.get("boundary_type", {})
        ps.surface = int(g(bt, "free_surface", 1))
        ps.inner = int(g(bt, "inner", 0))
        ps.other = int(g(bt, "other", -1))
        # guards of the reference (src/MpsInputOutput.cpp:531-543)
        if ps.partDist * ps.partDist <= ps.epsilonZero:
            raise ValueError("squared particle distance below machine precision")
        if ps.mpsType in (2, 3) and ps.pndType in (1, 2):
            raise ValueError("set 'pnd-type to 0' in the json file for incompressible MPS")
        return ps

    # ------------------------------------------------------------------ setParameters
    def set_parameters(self) -> "ParticleSystem":
        """MpsParticle::setParameters (src/MpsParticle.cpp:136-227), same evaluation order."""
        assert not self._parameters_set
        lo = self.partDist
        self.domainMinX -= lo * 3.0
        self.domainMinY -= lo * 3.0
        self.domainMinZ -= lo * 3.0
        self.domainMaxX += lo * 3.0
        self.domainMaxY += lo * 3.0
        self.domainMaxZ += lo * 3.0
        if int(self.dim) == 2:
            self.domainMinZ = 0.0
            self.domainMaxZ = 0.0
        self.numOfMeshs = (1 if self.wallType == 1 else 0) + (1 if self.femOn else 0) + (1 if self.forcedOn else 0)
        self.reS = lo * self.reS
        self.reL = lo * self.reL
        self.reRepulsiveForce = lo * self.reRepulsiveForce
        self.reS2 = self.reS * self.reS
        self.reL2 = self.reL * self.reL
        self.EPS_RE = self.EPS_RE * self.reS2 / 4.0
        n0S = n0L = n0G = lam = nN = 0.0
        lmin = math.ceil(self.reL / lo) + 1
        lmax = math.ceil(self.reL / lo) + 2
        two_d = int(self.dim) == 2
        zr = range(0, 1) if two_d else range(-lmin, lmax)
        for ix in range(-lmin, lmax):
            for iy in range(-lmin, lmax):
                for iz in zr:
                    x = lo * float(ix)
                    y = lo * float(iy)
                    z = lo * float(iz)
                    dst2 = x * x + y * y + z * z
                    if dst2 <= self.reL2:
                        if dst2 <= self.epsilonZero:
                            continue
                        dst = math.sqrt(dst2)
                        n0L += _weight(dst, self.reL, self.weightType)
                        lam += dst2 * _weight(dst, self.reL, self.weightType)
                        nN += 1
                        if dst2 <= self.reS2:
                            n0S += _weight(dst, self.reS, self.weightType)
                            n0G += _weight(dst, self.reS, self.weightType)
        self.pndSmallZero, self.pndLargeZero, self.pndGradientZero, self.numNeighZero = n0S, n0L, n0G, nN
        self.lambdaZero = lam / n0L
        self.coeffViscosity = 2.0 * self.KNM_VS1 * self.dim / (n0L * self.lambdaZero)
        self.coeffViscMultiphase = 2.0 * self.dim / (n0L * self.lambdaZero)
        self.coeffPressEMPS = self.soundSpeed * self.soundSpeed / n0S
        self.coeffPressGrad = -self.dim / n0G
        self.coeffPressWCMPS = self.soundSpeed * self.soundSpeed
        self.coeffShifting1 = self.dri * lo / n0S
        self.coeffShifting2 = self.coefA * lo * lo * self.cflNumber * self.machNumber
        self.coeffPPE = 2.0 * self.dim / (n0L * self.lambdaZero)
        self.coeffPPESource = self.relaxPND / (self.timeStep * self.timeStep * n0S)
        self.invPartDist = 1.0 / lo
        self.distCollisionLimit = lo * self.distLimitRatio
        self.distCollisionLimit2 = self.distCollisionLimit * self.distCollisionLimit
        self.restitutionCollision = 1.0 + self.collisionRatio
        self.numOfIterations = 0
        self.timeCurrent = 0.0
        self.velMax = 0.0
        self.CFLcurrent = self.cflNumber
        self.betaPnd = self.pndThreshold * n0S
        self.betaNeigh = self.neighThreshold * nN
        self.delta2 = self.npcdThreshold * self.npcdThreshold * lo * lo
        # ARC free-surface detection (src/MpsParticle.cpp:219-222)
        self.thetaArc = self.thetaThreshold / 180.0 * 3.14159265
        self.hThreshold2 = 1.33 * 1.33 * lo * lo
        self.dstThreshold2 = 2.0 * self.hThreshold2
        self.normThreshold2 = self.normThreshold * self.normThreshold
        self._parameters_set = True
        return self

    # ------------------------------------------------------------------ MpsBucket::allocateMemory
    def allocate_buckets(self, pos=None) -> "ParticleSystem":
        """MpsBucket::allocateMemory (src/MpsBucket.cpp:20-67) and, for particle walls, calcDomainLimits
        (src/MpsBucket.cpp:71-140, non-periodic branch).  `pos` = (N,3) initial positions (particle walls only)."""
        assert self._parameters_set
        self.bucketSide = self.reL * (1.0 + self.cflNumber)
        self.invBucketSide = 1.0 / self.bucketSide
        self.numBucketsX = int((self.domainMaxX - self.domainMinX) * self.invBucketSide) + 3
        self.numBucketsY = int((self.domainMaxY - self.domainMinY) * self.invBucketSide) + 3
        self.numBucketsZ = int((self.domainMaxZ - self.domainMinZ) * self.invBucketSide) + 3
        if int(self.dim) == 2:
            self.numBucketsZ = 1
        self.periodicLength = [0.0, 0.0, 0.0]
        self.periodicAxis = 0; self.periodicLen = 0.0; self.physDomMin = 0.0; self.physDomMax = 0.0
        if self.domainTypeBC == 1:
            self._periodic_limits(pos)
        elif self.wallType == 0:
            if self.limitTypeBC == 0:
                if pos is None:
                    raise ValueError("particle walls: initial positions are needed for the domain limits")
                lim = [[float(pos[:, k].min()), float(pos[:, k].max())] for k in range(3)]
                testdim = sum(1 for k in range(3) if lim[k][0] != lim[k][1])
                if testdim != self.dim:
                    raise ValueError("Dimensions in json [%d] and grid file [%d] do not match!" % (int(self.dim), testdim))
            else:
                lim = [[self.domainMinX, self.domainMaxX], [self.domainMinY, self.domainMaxY], [self.domainMinZ, self.domainMaxZ]]
            lo = self.partDist
            for k in range(3):
                if k < 2 or self.dim == 3:
                    lim[k][0] -= 0.5 * lo
                    lim[k][1] += 0.5 * lo
            nb = [self.numBucketsX, self.numBucketsY, self.numBucketsZ]
            for k in range(int(self.dim)):
                lim[k][0] -= self.bucketSide
                lim[k][1] += self.bucketSide
                nb[k] = int((lim[k][1] - lim[k][0]) / self.bucketSide + 1)
                lim[k][1] = lim[k][0] + nb[k] * self.bucketSide
            self.numBucketsX, self.numBucketsY, self.numBucketsZ = nb
            self.domainMinX, self.domainMaxX = lim[0]
            self.domainMinY, self.domainMaxY = lim[1]
            self.domainMinZ, self.domainMaxZ = lim[2]
            self.invBucketSide = 1.0 / self.bucketSide
        self.numBucketsXY = self.numBucketsX * self.numBucketsY
        self.numBucketsXYZ = self.numBucketsXY * self.numBucketsZ
        return self

    def _periodic_limits(self, pos):
        """calcDomainLimits, periodic branch (src/MpsBucket.cpp:142-289): ONE periodic direction.  Along it the physical domain is
        the particle extent (or the json limits) widened by half a particle distance each side, bucketSide is shrunk so that a whole
        number of buckets covers the periodic length, and one more bucket layer is added on each side — the layers that alias the
        opposite inner layers (copyDataBetweenBuckets, :356-417).  The other directions get one bucket of margin."""
        dirs = [self.periodicDirectionX, self.periodicDirectionY, self.periodicDirectionZ]
        if sum(1 for d in dirs if d) != 1:
            raise ValueError("periodic domain: exactly one of direction.x / .y / .z must be true (the reference's limits handle one)")
        a = dirs.index(True)
        if a == 2 and int(self.dim) == 2:
            raise ValueError("periodic domain: direction z in a 2D case")
        json_lim = [[self.domainMinX, self.domainMaxX], [self.domainMinY, self.domainMaxY], [self.domainMinZ, self.domainMaxZ]]
        if self.limitTypeBC == 0:
            if pos is None:
                raise ValueError("periodic domain with limit 0: initial positions are needed for the domain limits")
            lim = [[float(pos[:, k].min()), float(pos[:, k].max())] for k in range(3)]
            testdim = sum(1 for k in range(3) if lim[k][0] != lim[k][1])
            if testdim != self.dim:
                raise ValueError("Dimensions in json [%d] and grid file [%d] do not match!" % (int(self.dim), testdim))
        else:
            lim = [list(v) for v in json_lim]
        nb = [self.numBucketsX, self.numBucketsY, self.numBucketsZ]
        L = lim[a][1] - lim[a][0] + self.partDist
        lim[a][0] -= self.partDist * 0.5
        lim[a][1] += self.partDist * 0.5
        phys = (lim[a][0], lim[a][1])
        nb[a] = int((lim[a][1] - lim[a][0]) / self.bucketSide)
        self.bucketSide = L / int(L / self.bucketSide)
        lim[a][0] -= self.bucketSide
        lim[a][1] += self.bucketSide
        nb[a] += 2
        others = [k for k in range(3) if k != a and (k < 2 or int(self.dim) == 3 or a == 2)]
        for k in others:
            if self.wallType == 0:
                lim[k][0] -= self.bucketSide
                lim[k][1] += self.bucketSide
            else:
                lim[k] = [json_lim[k][0] - self.bucketSide, json_lim[k][1] + self.bucketSide]
            nb[k] = int((lim[k][1] - lim[k][0]) / self.bucketSide + 1)
            lim[k][1] = lim[k][0] + nb[k] * self.bucketSide
        self.periodicLength = [0.0, 0.0, 0.0]; self.periodicLength[a] = L
        self.periodicAxis = a; self.periodicLen = L; self.physDomMin, self.physDomMax = phys
        self.numBucketsX, self.numBucketsY, self.numBucketsZ = nb
        self.domainMinX, self.domainMaxX = lim[0]
        self.domainMinY, self.domainMaxY = lim[1]
        self.domainMinZ, self.domainMaxZ = lim[2]
        self.invBucketSide = 1.0 / self.bucketSide


# ---------------------------------------------------------------------- ctypes mirror of PmpsParams (include/pmps.h)
class PmpsParams(ctypes.Structure):
    _I = ["abi_version", "dim", "wallType", "mpsType", "solverType", "pndType", "weightType", "gradientType",
          "gradientCorrection", "slipCondition", "freeSurfType", "collisionType", "shiftingType", "repulsiveForceType",
          "fluidType", "domainTypeBC", "ghost", "fluid", "wall", "dummyWall", "surface", "inner", "other",
          "numBucketsX", "numBucketsY", "numBucketsZ", "numOfMeshs", "shiftingJacobi", "periodicAxis", "fractionMethod"]
    _D = ["domainMinX", "domainMinY", "domainMinZ", "domainMaxX", "domainMaxY", "domainMaxZ", "bucketSide", "invBucketSide",
          "partDist", "timeStep", "epsilonZero", "reS", "reL", "reS2", "reL2", "gravityX", "gravityY", "gravityZ",
          "densityWall", "DNS_FL1", "DNS_FL2", "soundSpeed", "gamma", "relaxPress", "alphaCompressibility",
          "reRepulsiveForce", "expectMaxVelocity", "repForceCoefMitsume", "repForceCoefLennardJones",
          "repForceCoefMonaghanKajtar"]
    _D2 = ["pndSmallZero", "pndLargeZero", "pndGradientZero", "lambdaZero", "numNeighZero", "coeffViscMultiphase",
           "coeffPressGrad", "coeffPressEMPS", "coeffPressWCMPS", "coeffShifting1", "coeffShifting2", "coeffPPE",
           "coeffPPESource", "distCollisionLimit2", "restitutionCollision", "lambdaCollision", "betaPnd", "betaNeigh",
           "delta2", "velMax"]
    _D3 = ["KNM_VS1", "KNM_VS2", "DNS_SDT", "PHI_1", "PHI_2", "cohes", "MEU0", "DG", "I0", "mm", "Cd", "VF_min", "VF_max", "relaxPND", "timeCurrent",
           "thetaArc", "hThreshold2", "dstThreshold2", "normThreshold2"]
    _D4 = ["periodicLength", "physDomMin", "physDomMax", "powerLawN"]   # of the periodic axis (ParticleSystem.periodicLen / physDomMin / physDomMax)
    _fields_ = ([(n, ctypes.c_int32) for n in _I] + [(n, ctypes.c_double) for n in _D]
                + [("uniformVelWall", ctypes.c_double * 3)] + [(n, ctypes.c_double) for n in _D2 + _D3 + _D4])


def to_c_params(ps: ParticleSystem, shifting_jacobi: bool = True) -> PmpsParams:
    c = PmpsParams()
    for n in PmpsParams._I:
        if n == "abi_version":
            c.abi_version = 5
        elif n == "shiftingJacobi":
            c.shiftingJacobi = 1 if shifting_jacobi else 0
        elif n == "periodicAxis":
            c.periodicAxis = int(getattr(ps, "periodicAxis", 0))
        else:
            setattr(c, n, int(getattr(ps, n)))
    for n in PmpsParams._D + PmpsParams._D2 + PmpsParams._D3:
        setattr(c, n, float(getattr(ps, n)))
    for k in range(3):
        c.uniformVelWall[k] = float(ps.uniformVelWall[k])
    c.periodicLength = float(getattr(ps, "periodicLen", 0.0)); c.physDomMin = float(getattr(ps, "physDomMin", 0.0)); c.physDomMax = float(getattr(ps, "physDomMax", 0.0)); c.powerLawN = float(ps.N)
    return c
