"""Polygon (STL triangle) wall, host side.

Host-side mirror of the parts of the reference's PolygonMesh that are *not* on the per-step hot path and therefore
stay on the host: binary STL read, duplicate-vertex merge, face normals
(src/PolygonMesh.cpp:42-108 via igl::readSTL / remove_duplicate_vertices / per_face_normals) and the wall
weight tables Z(r), nNeigh(r) of initWijnNeigh (src/PolygonMesh.cpp:377-527).  The per-step closest-point /
mirror-particle evaluation is the CUDA kernel `wall_closest_point` (pmps_wall_closest_point).
"""
from __future__ import annotations

import math
import struct
from dataclasses import dataclass

import numpy as np

_STL_DT = np.dtype([("n", "<f4", 3), ("v", "<f4", (3, 3)), ("a", "<u2")])


def write_binary_stl(path: str, tris: np.ndarray, normals: np.ndarray | None = None) -> None:
    """tris: (nF,3,3) float; stored as float32 like any binary STL."""
    tris = np.asarray(tris, dtype=np.float64)
    rec = np.zeros(len(tris), dtype=_STL_DT)
    rec["v"] = tris.astype(np.float32)
    if normals is None:
        e1 = tris[:, 1] - tris[:, 0]
        e2 = tris[:, 2] - tris[:, 0]
        nn = np.cross(e1, e2)
        nn /= np.maximum(np.linalg.norm(nn, axis=1, keepdims=True), 1e-300)
        normals = nn
    rec["n"] = np.asarray(normals, dtype=np.float32)
    with open(path, "wb") as fh:
        fh.write(b"polymps_b200 synthetic tank".ljust(80, b" "))
        fh.write(struct.pack("<I", len(tris)))
        fh.write(rec.tobytes())


def read_binary_stl(path: str) -> np.ndarray:
    with open(path, "rb") as fh:
        b = fh.read()
    n = struct.unpack("<I", b[80:84])[0]
    if len(b) < 84 + 50 * n:
        raise IOError("STL mesh %s could not be read (ASCII STL is not supported here)" % path)
    rec = np.frombuffer(b[84:84 + 50 * n], dtype=_STL_DT)
    return rec["v"].astype(np.float64)  # (nF,3,3)


def box_triangles(lx: float, ly: float, lz: float, open_top: bool = True) -> np.ndarray:
    """Axis-aligned tank [0,lx]x[0,ly]x[0,lz], two triangles per face, open top (z = lz) by default:
    10 triangles with the face order of the reference's dam-break tank (bottom, x=0, y=ly, x=lx, y=0)."""
    A = np.array
    t = [
        [[lx, ly, 0], [0, 0, 0], [0, ly, 0]], [[0, 0, 0], [lx, ly, 0], [lx, 0, 0]],              # bottom z=0
        [[0, ly, lz], [0, 0, 0], [0, 0, lz]], [[0, 0, 0], [0, ly, lz], [0, ly, 0]],              # x=0
        [[0, ly, lz], [lx, ly, 0], [0, ly, 0]], [[lx, ly, 0], [0, ly, lz], [lx, ly, lz]],        # y=ly
        [[lx, ly, 0], [lx, 0, lz], [lx, 0, 0]], [[lx, 0, lz], [lx, ly, 0], [lx, ly, lz]],        # x=lx
        [[lx, 0, lz], [0, 0, 0], [lx, 0, 0]], [[0, 0, 0], [lx, 0, lz], [0, 0, lz]],              # y=0
    ]
    if not open_top:
        t += [[[0, 0, lz], [lx, 0, lz], [0, ly, lz]], [[lx, ly, lz], [0, ly, lz], [lx, 0, lz]]]
    return A(t, dtype=np.float64)


@dataclass
class PolygonMesh:
    NV: np.ndarray          # (nV,3) merged vertices
    NF: np.ndarray          # (nF,3) int32
    NNormals: np.ndarray    # (nF,3) unit face normals
    xDataPND: np.ndarray = None
    xDataNeigh: np.ndarray = None
    Zij: np.ndarray = None
    nNeighWall: np.ndarray = None

    @classmethod
    def from_triangles(cls, tris: np.ndarray) -> "PolygonMesh":
        """Vertex merge as igl::remove_duplicate_vertices(V,F,2e-15,...): unique rows of round(V/eps) in
        lexicographic order, first occurrence kept; then per-face unit normals (v1-v0)x(v2-v0)."""
        tris = np.asarray(tris, dtype=np.float64)
        V = tris.reshape(-1, 3)
        F = np.arange(len(V), dtype=np.int64).reshape(-1, 3)
        key = np.round(V / 2.0e-15)
        _, first, inv = np.unique(key, axis=0, return_index=True, return_inverse=True)
        NV = V[first]
        NF = inv.reshape(-1)[F].astype(np.int32)
        v1 = NV[NF[:, 1]] - NV[NF[:, 0]]
        v2 = NV[NF[:, 2]] - NV[NF[:, 0]]
        N = np.cross(v1, v2)
        r = np.sqrt(N[:, 0] * N[:, 0] + N[:, 1] * N[:, 1] + N[:, 2] * N[:, 2])
        N = np.where(r[:, None] == 0.0, 0.0, N / np.where(r == 0.0, 1.0, r)[:, None])
        return cls(NV=np.ascontiguousarray(NV), NF=np.ascontiguousarray(NF), NNormals=np.ascontiguousarray(N))

    @classmethod
    def from_stl(cls, path: str) -> "PolygonMesh":
        return cls.from_triangles(read_binary_stl(path))

    def init_wall_tables(self, dim: int, wtype: int, lo: float, reL: float, reS: float, n: int = 50) -> "PolygonMesh":
        """PolygonMesh::initWijnNeigh (src/PolygonMesh.cpp:377-527): half-space lattice of virtual wall particles at
        y = (iy-0.5)*lo, iy<=0, sampled at wall distances r_k = k*reS/n, k=1..n.  Scalar Python floats, reference
        loop order, so the tables are bit-identical (pinned by tests/test_mesh.py)."""
        reS2, reL2 = reS * reS, reL * reL
        limS, limL = math.ceil(reS / lo), math.ceil(reL / lo)
        xP, xN, Z, NN = [], [], [], []

        def w(rij):
            if wtype == 1:
                return reS / rij + rij / reS - 2.0
            if wtype == 2:
                return reS / rij - rij / reS
            if wtype == 3:
                return math.pow(1.0 - rij / reS, 3.0)
            if wtype == 4:
                return math.pow(1.0 - rij / reS, 2.0)
            return reS / rij - 1.0

        for i in range(1, n + 1):
            xi, yi, zi = 0.0, i * reS / n, 0.0
            riw = (0.0 - xi) * (0.0 - xi) + (0.0 - yi) * (0.0 - yi)
            if dim == 3:
                riw = riw + (0.0 - zi) * (0.0 - zi)
            riw = math.sqrt(riw)
            xP.append(riw / reS)
            xN.append(riw / reL)
            n0 = 0.0
            zr = range(-limS, limS + 1) if dim == 3 else (0,)
            for ix in range(-limS, limS + 1):
                for iy in range(-limS, 1):
                    for iz in zr:
                        xj, yj, zj = lo * float(ix), lo * float(iy - 0.5), lo * float(iz)
                        rij2 = (xj - xi) * (xj - xi) + (yj - yi) * (yj - yi)
                        if dim == 3:
                            rij2 = rij2 + (zj - zi) * (zj - zi)
                        if rij2 < reS2:
                            if rij2 == 0.0:
                                continue
                            n0 += w(math.sqrt(rij2))
            Z.append(n0)
            cnt = 0
            zr = range(-limL, limL + 1) if dim == 3 else (0,)
            for ix in range(-limL, limL + 1):
                for iy in range(-limL, 1):
                    for iz in zr:
                        xj, yj, zj = lo * float(ix), lo * float(iy - 0.5), lo * float(iz)
                        rij2 = (xj - xi) * (xj - xi) + (yj - yi) * (yj - yi)
                        if dim == 3:
                            rij2 = rij2 + (zj - zi) * (zj - zi)
                        if rij2 < reL2 and rij2 != 0.0:
                            cnt += 1
            NN.append(cnt)
        self.xDataPND = np.array(xP)
        self.xDataNeigh = np.array(xN)
        self.Zij = np.array(Z)
        self.nNeighWall = np.array(NN, dtype=np.int32)
        return self
