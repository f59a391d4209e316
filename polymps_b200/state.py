"""Host-side particle arrays in the reference's layout (include/MpsParticle.h:139-241): scalars a[i], vectors
a[3*i+k], int32 ids, 1-byte bool flags — and the ctypes mirror of PmpsHostArrays (include/pmps.h).

`HostState.from_case` follows the allocation + initial values of MpsInputOutput::readMpsParticleFile
(src/MpsInputOutput.cpp:645-857): everything zero, distParticleWall2 = 10e8*lo, RHO/PTYPE/MEU from the fluid type.
"""
from __future__ import annotations

import ctypes

import numpy as np

I32 = ["particleType", "particleBC", "numNeigh", "PTYPE", "nearMeshType", "numNeighWallContribution", "elementID"]
U8 = ["particleNearWall"]
F64 = ["press", "pndi", "pndki", "pndski", "pndSmall", "npcdDeviation2", "concentration", "velDivergence", "diffusiveTerm",
       "pndWallContribution", "distParticleWall2", "deviationDotPolygonNormal", "RHO", "MEU", "pndAuxVar1", "pndAuxVar2",
       "pressurePPE", "sourceTerm"]
VEC = ["pos", "vel", "acc", "accStar", "npcdDeviation", "gradConcentration", "correcMatrixRow1", "correcMatrixRow2",
       "correcMatrixRow3", "particleAtWallPos", "mirrorParticlePos", "polygonNormal", "forceWall", "dvelCollision", "normal"]
# non-Newtonian fields (src/MpsInputOutput.cpp:729-744); transferred only by handles created with fluidType 1
RHEO = ["Cv", "II", "MEU_Y", "Inertia", "p_rheo_new", "p_smooth", "VF", "S11", "S12", "S13", "S22", "S23", "S33"]
ALL_FIELDS = I32 + U8 + F64 + VEC + RHEO


class PmpsHostArrays(ctypes.Structure):
    _fields_ = ([("numParticles", ctypes.c_int32), ("numParticlesZero", ctypes.c_int32)]
                + [(n, ctypes.POINTER(ctypes.c_int32)) for n in I32]
                + [(n, ctypes.POINTER(ctypes.c_uint8)) for n in U8]
                + [(n, ctypes.POINTER(ctypes.c_double)) for n in F64]
                + [(n, ctypes.POINTER(ctypes.c_double)) for n in VEC]
                + [(n, ctypes.POINTER(ctypes.c_double)) for n in RHEO])


class PmpsScalars(ctypes.Structure):
    _fields_ = [("numParticles", ctypes.c_int32), ("solverIter", ctypes.c_int32), ("velMax", ctypes.c_double),
                ("CFLcurrent", ctypes.c_double), ("solverError", ctypes.c_double), ("maxDisplacement", ctypes.c_double),
                ("outOfDomainEvents", ctypes.c_int32), ("maxBucketCount", ctypes.c_int32), ("CFLvisc", ctypes.c_double)]


class HostState:
    """All per-particle host arrays (numpy, C-contiguous, reference layout)."""

    def __init__(self, n: int):
        self.numParticles = n
        self.numParticlesZero = n
        for f in I32:
            setattr(self, f, np.zeros(n, dtype=np.int32))
        for f in U8:
            setattr(self, f, np.zeros(n, dtype=np.uint8))
        for f in F64 + RHEO:
            setattr(self, f, np.zeros(n, dtype=np.float64))
        for f in VEC:
            setattr(self, f, np.zeros(3 * n, dtype=np.float64))

    @classmethod
    def from_case(cls, case, ps) -> "HostState":
        s = cls(case.n)
        s.particleType[:] = case.ptype
        s.pos[:] = np.asarray(case.pos, dtype=np.float64).reshape(-1)
        s.vel[:] = np.asarray(case.vel, dtype=np.float64).reshape(-1)
        s.press[:] = case.press
        s.distParticleWall2[:] = 10e8 * ps.partDist
        s.RHO[:] = ps.DNS_FL1
        s.PTYPE[:] = 1
        s.MEU[:] = ps.KNM_VS1 * ps.DNS_FL1
        if ps.fluidType == 1:   # grid id 1 = sediment: fluid particle of phase 2 (src/MpsInputOutput.cpp:839-856)
            sed = s.particleType == 1
            s.particleType[sed] = ps.fluid; s.PTYPE[sed] = 2
            s.RHO[sed] = ps.DNS_FL2; s.MEU[sed] = ps.KNM_VS2 * ps.DNS_FL2
        return s

    @classmethod
    def from_dump(cls, d: dict) -> "HostState":
        """From a raw array dump of the reference (dict name -> array, see tests/golden/make_golden.py)."""
        n = int(d["numParticles"][1])
        s = cls(n)
        s.numParticles = int(d["numParticles"][0])
        for f in ALL_FIELDS:
            if f in d:
                getattr(s, f)[:] = d[f]
        return s

    def copy(self) -> "HostState":
        o = HostState(self.numParticlesZero)
        o.numParticles = self.numParticles
        for f in ALL_FIELDS:
            getattr(o, f)[:] = getattr(self, f)
        return o

    def c_struct(self, fields=None) -> PmpsHostArrays:
        """Pointers to the arrays (all, or only `fields`; the rest NULL)."""
        c = PmpsHostArrays()
        c.numParticles = self.numParticles
        c.numParticlesZero = self.numParticlesZero
        for f in ALL_FIELDS:
            if fields is not None and f not in fields:
                continue
            a = getattr(self, f)
            assert a.flags["C_CONTIGUOUS"]
            t = ctypes.c_int32 if f in I32 else ctypes.c_uint8 if f in U8 else ctypes.c_double
            setattr(c, f, a.ctypes.data_as(ctypes.POINTER(t)))
        return c

    def vec(self, name: str) -> np.ndarray:
        return getattr(self, name).reshape(-1, 3)
