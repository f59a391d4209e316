"""Deterministic synthetic cases in the reference's own input formats (.json / .grid / .stl).

The reference ships no .grid files (its input/grid/grid.zip is a stripped LFS blob), so every
run — the reference's and ours — starts from a grid synthesized here: a cubic lattice at ((i+1/2) lo, ...)
(README.md:65-74 of the reference for the format; parser src/MpsInputOutput.cpp:659-776).  The JSON written by
`write_case` uses the reference's key names so that the unmodified reference binary reads it.

Generators:
  dam3d(...)           3D dam break, polygon (STL) tank, the InputDamWC / InputDamINC option sets  
  dam2d_particle(...)  2D dam break with particle walls (1 wall layer id 2 + 2 dummy layers id 3), the
                       dam1610_h300_lo0p0050_{WC,INC} option sets
"""
from __future__ import annotations

import json
import math
import os
from dataclasses import dataclass, field

import numpy as np

from .mesh import box_triangles, write_binary_stl


@dataclass
class Case:
    name: str
    options: dict                      # JSON document (reference key names)
    ptype: np.ndarray                  # (N,) int32 material id
    pos: np.ndarray                    # (N,3)
    vel: np.ndarray                    # (N,3)
    press: np.ndarray                  # (N,)
    tris: np.ndarray | None = None     # (nF,3,3) rigid mesh triangles (polygon walls)
    meta: dict = field(default_factory=dict)
    deformable_tris: np.ndarray | None = None   # mesh 1 (fem_MESH) and mesh 2 (forced_MESH, moved by the host every step):
    forced_tris: np.ndarray | None = None       # src/main.cpp:471-493 reads them in this order

    @property
    def n(self) -> int:
        return len(self.ptype)


def base_options(*, dim, lo, dt, wall_type, mps_type, c0=15.0, re_small=2.1, re_large=2.1, grad_type=3,
                 grad_corr=False, shifting=2, pnd_type=2, fs_type=1, fs_pnd=0.98, max_vel=4.0, gravity=(0.0, 0.0, -9.81),
                 dom_min=(0.0, 0.0, 0.0), dom_max=(1.0, 1.0, 1.0), final_time=1.0, solver=0, alpha=1e-6, relax_pnd=0.01,
                 collision=0, slip=0, rep_type=2, weight_type=0, fs_neigh=0.85, fluid_type=0, rheo=None, relax=1.0) -> dict:
    """A JSON document with the reference's key names (src/MpsInputOutput.cpp:134-505)."""
    return {
        "flags": {"wall_type": wall_type, "fem_MESH": False, "forced_MESH": False,
                  "output_VTU": {"type": 0, "freeSurfWall": False, "pnd": True, "neigh": False, "deviation": False,
                                 "concentration": False, "auxiliar": False, "non_newtonian": False},
                  "output_TXT": {"press": False, "force": False}},
        "pathNames": {"particle_grid_file": "input/grid/case.grid", "mesh_rigid_file": "input/stl/case.stl",
                      "mesh_deformable_file": "input/stl/none.stl", "mesh_forced_file": "input/stl/none.stl",
                      "vtu_output_folder": "output/case", "forceTxt_file": "output/force.txt",
                      "pressTxt_file": "output/press.txt"},
        "domain": {"min": dict(zip("xyz", map(float, dom_min))), "max": dict(zip("xyz", map(float, dom_max))),
                   "boundary": {"type": 0, "limit": 0, "direction": {"x": False, "y": False, "z": False}}},
        "physical": {"fluid_density": 1000.0, "wall_density": 1000.0, "kinematic_visc": 1e-6, "fluid_type": fluid_type,
                     "gravity": dict(zip("xyz", map(float, gravity))),
                     "rheological": dict({"kinematic_visc_phase_2": 1e-6, "fluid_density_phase_1": 1000.0,
                                          "fluid_density_phase_2": 1540.0, "sediment_density": 1540.0}, **(rheo or {}))},
        "numerical": {
            "dimension": dim, "particle_dist": lo, "time_step": dt, "final_time": final_time, "iter_output": 1000000,
            "CFL_number": 0.2, "weight_type": weight_type, "slip_condition": slip,
            "effective_radius": {"small": re_small, "large": re_large},
            "gradient": {"type": grad_type, "correction": grad_corr, "relax_fact": relax},
            "mps_type": mps_type,
            "explicit_mps": {"equation_state": {"speed_sound": c0, "gamma": 7.0}},
            "semi_implicit_mps": {"solver_type": solver, "weak_compressibility": {"alpha": alpha},
                                  "source_term": {"relax_pnd": relax_pnd}},
            "particle_shifting": {"type": shifting, "DRI": 0.01, "coef_A": 2.0, "mach_number": 0.1, "adj_vel_A": 0.9},
            "pnd": {"type": pnd_type, "diffusive_coeff": 0.35},
            "wall_repulsive_force": {"type": rep_type, "re": 0.5, "maxVel": max_vel,
                                     "coefficient": {"Mitsume": 40000000.0, "Lennard-Jones": 2.0, "Monaghan-Kajtar": 1.0},
                                     "eps_re": 0.01},
            "free_surface_threshold": {"type": fs_type, "pnd": fs_pnd, "neigh": fs_neigh, "NPCD": 0.2, "ARC": 45.0,
                                       "normal": 0.2},
            "particle_collision": {"type": collision, "ratio": 0.2, "dist_limit_ratio": 0.85, "lambda": 0.2},
            "particle_type": {"ghost": -1, "fluid": 0, "wall": 2, "dummyWall": 3},
            "boundary_type": {"free_surface": 1, "inner": 0, "other": -1},
        },
    }


def dam3d(nx: int, ny: int, nz: int, *, lo: float = 0.01, incompressible: bool = False, tank=None,
          jitter: float = 0.0, seed: int = 20261104, name: str | None = None, dt: float | None = None,
          offset=(0.0, 0.0, 0.0), **overrides) -> Case:
    """3D dam break in an open-top polygon tank (InputDamWC / InputDamINC option sets).

    Fluid column of nx*ny*nz particles on the lattice ((i+1/2)lo, (j+1/2)lo, (m+1/2)lo) (+offset, +optional jitter);
    tank = (Lx, Ly, Lz), default the Lobovsky proportions scaled to the column: Lx = (1.61/0.6)*nx*lo,
    Ly = ny*lo, Lz = 2*nz*lo.  c0 = 15*sqrt(H/0.3), dt = 0.375*lo/c0 (the acoustic Courant number of InputDamWC),
    maxVel = 2*sqrt(g H) — the synthetic-block recipe of SURVEY.md §8d.
    """
    H = nz * lo
    if tank is None:
        tank = ((1.61 / 0.6) * nx * lo, ny * lo, 2.0 * nz * lo)
    c0 = 15.0 * math.sqrt(H / 0.3)
    if dt is None:
        dt = 0.375 * lo / c0
    i, j, m = np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij")
    # x fastest in memory order (like a grid file written x-inner): order by (m, j, i)
    order = np.lexsort((i.ravel(), j.ravel(), m.ravel()))
    pos = np.stack([(i.ravel()[order] + 0.5) * lo + offset[0], (j.ravel()[order] + 0.5) * lo + offset[1],
                    (m.ravel()[order] + 0.5) * lo + offset[2]], axis=1)
    if jitter > 0.0:
        rng = np.random.default_rng(seed)
        pos = pos + rng.uniform(-jitter * lo, jitter * lo, size=pos.shape)
    n = len(pos)
    kw = dict(dim=3, lo=lo, dt=dt, wall_type=1, c0=c0, max_vel=2.0 * math.sqrt(9.81 * H),
              dom_min=(0.0, 0.0, 0.0), dom_max=(tank[0], tank[1], tank[2]))
    if incompressible:   # InputDamINC: gradient 0 + correction, shifting off, pnd 0, CG
        kw.update(mps_type=2, grad_type=0, grad_corr=True, shifting=0, pnd_type=0, dt=overrides.pop("dt_inc", dt * 2.0))
    else:                # InputDamWC
        kw.update(mps_type=1, grad_type=3, grad_corr=False, shifting=2, pnd_type=2)
    kw.update(overrides)
    opts = base_options(**kw)
    return Case(name=name or ("dam3d_%s_%dx%dx%d" % ("inc" if incompressible else "wc", nx, ny, nz)), options=opts,
                ptype=np.zeros(n, dtype=np.int32), pos=np.ascontiguousarray(pos), vel=np.zeros((n, 3)),
                press=np.zeros(n),
                tris=box_triangles(*tank).astype(np.float32).astype(np.float64),  # what a binary STL can hold
                meta={"tank": tank, "H": H, "c0": c0})


def flow_field(pos: np.ndarray, *, Lc: float, H: float, W: float, vmax: float, lo: float, cell: float = 40.0) -> np.ndarray:
    """The MOVING initial state of the benchmark block (both arms start from it): counter-rotating vortex cells of about
    `cell` particle spacings in the x-z plane, peak speed vmax (= expectMaxVelocity of the case, 2 sqrt(g H): at the block's
    time step a particle then moves 0.086 lo per step, so the bucket order of the GPU path is rebuilt every 2-3 steps and the
    reference's per-step re-bucketing does real work too).  Stream function psi = A sin(mx pi x/Lc) sin(mz pi z/H) g(y):
    divergence-free, tangential at the tank floor, the left wall and the column's free faces; g(y) = 0.6 + 0.4 cos(2 pi my y/W)
    makes it vary across the tank.  Deterministic, no RNG."""
    mz = max(1, int(round(H / (cell * lo)))); mx = max(1, int(round(Lc / (cell * lo)))); my = max(1, int(round(W / (2.0 * cell * lo))))
    kx = mx * math.pi / Lc; kz = mz * math.pi / H
    A = vmax / max(kx, kz)
    x, y, z = pos[:, 0], pos[:, 1], pos[:, 2]
    g = 0.6 + 0.4 * np.cos(2.0 * math.pi * my * y / W)
    v = np.zeros_like(pos)
    v[:, 0] = A * kz * np.sin(kx * x) * np.cos(kz * z) * g
    v[:, 2] = -A * kx * np.cos(kx * x) * np.sin(kz * z) * g
    return v


def dam3d_block(k: int, flow: bool = False, **kw) -> Case:
    """The benchmark block of SURVEY.md §8d: fluid column 2k x k x k (k=171 -> 10.0M, 292 -> 49.8M, 464 -> 199.8M).
    flow=True: the moving initial state (flow_field) instead of the column at rest."""
    c = dam3d(2 * k, k, k, **kw)
    if flow:
        lo = c.options["numerical"]["particle_dist"]
        c.vel = flow_field(c.pos, Lc=2 * k * lo, H=k * lo, W=k * lo, vmax=c.options["numerical"]["wall_repulsive_force"]["maxVel"], lo=lo)
        c.meta["flow"] = True
    return c


def dam3d_splash(nx: int, ny: int, nz: int, *, lo: float = 0.01, seed: int = 3, name: str | None = None, **kw) -> Case:
    """Particles LEAVE the domain: a dam3d column whose whole top layer rises at staggered speeds of 0.09-0.18 lo per step (inside
    the search slack (bucketSide - re)/2 = 0.21 lo a step may move, so the reference's own stale-bucket search stays complete)
    under a domain top set 0.55 lo above that layer, so that in the out-of-domain checks of steps 3-7 some particles cross
    z = domainMax - bucketSide (src/MpsParticle.cpp:286-296).  The top layer ends the particle array (x fastest, then y, then z),
    so the LAST live particle is among the leavers: the reference's serial swap-with-last loop copies a particle that is itself
    outside over a hole and does not re-test it in the same call (src/MpsParticle.cpp:298-396) — the order-dependent detail
    SURVEY H5 names."""
    tank = ((1.61 / 0.6) * nx * lo, ny * lo, (nz + 1.5) * lo)
    c = dam3d(nx, ny, nz, lo=lo, jitter=0.05, name=name or "splash", tank=tank, **kw)
    dt = c.options["numerical"]["time_step"]
    bucket = max(c.options["numerical"]["effective_radius"].values()) * lo * (1.0 + c.options["numerical"]["CFL_number"])
    # the reference pads the domain by 3 lo (src/MpsParticle.cpp:139-144): limit = domain max + 3 lo - bucketSide
    c.options["domain"]["max"]["z"] = float((nz - 0.5 + 0.55) * lo - 3.0 * lo + bucket)
    top = c.pos[:, 2] > (nz - 1) * lo
    rng = np.random.default_rng(seed)
    c.vel[top, 2] = rng.uniform(0.09, 0.18, size=int(top.sum())) * lo / dt
    return c


def dam3d_block_columns(k: int, i0: int, i1: int, *, lo: float = 0.01, flow: bool = False):
    """Lattice columns i0 <= i < i1 (along x) of the benchmark block dam3d_block(k), without building the whole block:
    (ptype, pos, vel, press, gid) with gid = the particle's index in dam3d_block(k)'s order (x fastest, then y, then z).
    A slab rank generates only its own part of a 50M / 200M-particle block this way (polymps_b200/slab.py)."""
    nx, ny, nz = 2 * k, k, k
    i0 = max(0, int(i0)); i1 = min(nx, int(i1)); ni = max(0, i1 - i0)
    m, j, i = np.meshgrid(np.arange(nz, dtype=np.int64), np.arange(ny, dtype=np.int64), np.arange(i0, i0 + ni, dtype=np.int64), indexing="ij")
    i = i.ravel(); j = j.ravel(); m = m.ravel()
    pos = np.stack([(i + 0.5) * lo, (j + 0.5) * lo, (m + 0.5) * lo], axis=1)
    n = len(pos)
    vel = np.zeros((n, 3))
    if flow:
        vel = flow_field(pos, Lc=nx * lo, H=nz * lo, W=ny * lo, vmax=2.0 * math.sqrt(9.81 * nz * lo), lo=lo)
    return {"ptype": np.zeros(n, dtype=np.int32), "pos": np.ascontiguousarray(pos), "vel": vel, "press": np.zeros(n),
            "gid": (m * ny + j) * nx + i}


def dam3d_block_options(k: int, *, lo: float = 0.01, incompressible: bool = False) -> Case:
    """dam3d_block(k) without particles: options, tank mesh and meta only."""
    c = dam3d(1, 1, 1, lo=lo)   # a one-particle case just for the key set
    nx, ny, nz = 2 * k, k, k
    H = nz * lo; tank = ((1.61 / 0.6) * nx * lo, ny * lo, 2.0 * nz * lo); c0 = 15.0 * math.sqrt(H / 0.3)
    kw = dict(dim=3, lo=lo, dt=0.375 * lo / c0, wall_type=1, c0=c0, max_vel=2.0 * math.sqrt(9.81 * H), dom_min=(0.0, 0.0, 0.0),
              dom_max=tank, mps_type=1, grad_type=3, grad_corr=False, shifting=2, pnd_type=2)
    if incompressible:   # InputDamINC, as dam3d(..., incompressible=True)
        kw.update(mps_type=2, grad_type=0, grad_corr=True, shifting=0, pnd_type=0, dt=2.0 * kw["dt"])
    return Case(name="block%d" % k, options=base_options(**kw), ptype=np.zeros(0, dtype=np.int32), pos=np.zeros((0, 3)), vel=np.zeros((0, 3)),
                press=np.zeros(0), tris=box_triangles(*tank).astype(np.float32).astype(np.float64), meta={"tank": tank, "H": H, "c0": c0})


def dam2d_particle(nx: int, nz: int, *, lo: float = 0.005, tank_nx: int | None = None, tank_nz: int | None = None,
                   incompressible: bool = False, name: str | None = None, **overrides) -> Case:
    """2D dam break with particle walls (dam1610_h300_lo0p0050_{WC,INC} option sets).
    The reference's 2D cases live in the x-y plane with gravity along -y (input/dam1610_h300_lo0p0050_WC.json)."""
    tank_nx = tank_nx or int(round(nx * 1.61 / 0.6))
    tank_nz = tank_nz or 2 * nz
    P, T = [], []
    for jy in range(-3, tank_nz):
        for ix in range(-3, tank_nx + 3):
            inside = 0 <= ix < tank_nx and jy >= 0
            if inside:
                if ix < nx and jy < nz:
                    P.append(((ix + 0.5) * lo, (jy + 0.5) * lo, 0.0)); T.append(0)
                continue
            layer = max(-ix, ix - tank_nx + 1, -jy)   # 1 = wall layer, 2..3 = dummy layers
            P.append(((ix + 0.5) * lo, (jy + 0.5) * lo, 0.0)); T.append(2 if layer == 1 else 3)
    pos = np.array(P); ptype = np.array(T, dtype=np.int32)
    H = nz * lo
    c0 = 15.0 * math.sqrt(H / 0.3)
    dt = 0.375 * lo / c0
    kw = dict(dim=2, lo=lo, dt=dt, wall_type=0, c0=c0, re_small=2.1, re_large=3.1, gravity=(0.0, -9.81, 0.0),
              grad_type=0, grad_corr=True, fs_type=0, fs_pnd=0.97, max_vel=2.0 * math.sqrt(9.81 * H),
              dom_min=(-4 * lo, -4 * lo, 0.0), dom_max=((tank_nx + 4) * lo, (tank_nz + 1) * lo, 0.0))
    if incompressible:
        kw.update(mps_type=2, shifting=0, pnd_type=0, dt=dt * 2.0)
    else:
        kw.update(mps_type=1, shifting=2, pnd_type=2)
    kw.update(overrides)
    n = len(pos)
    return Case(name=name or ("dam2d_%s_%dx%d" % ("inc" if incompressible else "wc", nx, nz)), options=base_options(**kw),
                ptype=ptype, pos=pos, vel=np.zeros((n, 3)), press=np.zeros(n), tris=None, meta={"H": H, "c0": c0})


def channel2d_periodic(nx: int, nz: int, *, lo: float = 0.005, hump: int = 0, drift: float = 1.0, window: float = 0.45,
                       incompressible: bool = False, name: str | None = None, **overrides) -> Case:
    """2D open channel, PERIODIC in x, particle walls (the boundary block of input/dam1610_BC.json: type 1, one direction): a floor
    of one wall + two dummy layers over nx columns, nz layers of fluid on top (+ `hump` more layers over the middle third, so the
    free surface is not flat), every fluid particle drifting with `drift` * expectMaxVelocity along -x.  The domain limits come from
    the json (limit 1; setParameters widens them by 3 lo each side first, src/MpsParticle.cpp:139-144) with the window shifted by
    `window` lo against the lattice: the first fluid column then starts 0.05 lo inside the left face of the physical domain and is
    wrapped to the right one by the first prediction (src/MpsParticle.cpp:238-269)."""
    P, T = [], []
    for jy in range(-3, nz + hump):
        for ix in range(nx):
            if jy < 0:
                P.append(((ix + 0.5) * lo, (jy + 0.5) * lo, 0.0)); T.append(2 if jy == -1 else 3)
            elif jy < nz or nx // 3 <= ix < 2 * nx // 3:
                P.append(((ix + 0.5) * lo, (jy + 0.5) * lo, 0.0)); T.append(0)
    pos = np.array(P); ptype = np.array(T, dtype=np.int32)
    H = (nz + hump) * lo
    c0 = 15.0 * math.sqrt(H / 0.3)
    dt = 0.375 * lo / c0
    vmax = 2.0 * math.sqrt(9.81 * H)
    x0, x1 = 0.5 * lo, (nx - 0.5) * lo
    kw = dict(dim=2, lo=lo, dt=dt, wall_type=0, c0=c0, re_small=2.1, re_large=3.1, gravity=(0.0, -9.81, 0.0),
              grad_type=0, grad_corr=True, fs_type=0, fs_pnd=0.97, max_vel=vmax,
              dom_min=(x0 + (3.0 + window) * lo, -0.5 * lo, 0.0), dom_max=(x1 - (3.0 - window) * lo, (2 * (nz + hump) - 2.5) * lo, 0.0))
    if incompressible:
        kw.update(mps_type=2, shifting=0, pnd_type=0, dt=dt * 2.0)
    else:
        kw.update(mps_type=1, shifting=2, pnd_type=2)
    kw.update(overrides)
    opts = base_options(**kw)
    opts["domain"]["boundary"] = {"type": 1, "limit": 1, "direction": {"x": True, "y": False, "z": False}}
    n = len(pos)
    vel = np.zeros((n, 3)); vel[ptype == 0, 0] = -drift * vmax
    return Case(name=name or "channel2d_periodic_%dx%d" % (nx, nz), options=opts, ptype=ptype, pos=pos, vel=vel, press=np.zeros(n),
                tris=None, meta={"H": H, "c0": c0})


def channel3d_periodic(nx: int, ny: int, nz: int, *, lo: float = 0.01, drift: float = 1.0, window: float = 0.45, jitter: float = 0.0,
                       seed: int = 11, incompressible: bool = False, name: str | None = None, **overrides) -> Case:
    """3D open channel along y, PERIODIC in y, polygon walls: floor z = 0 and side walls x = 0, x = 2 nx lo (triangles reaching far
    beyond both periodic faces), a fluid block of nx * ny * nz particles against the x = 0 wall that fills the whole periodic length
    ny lo and drifts with `drift` * expectMaxVelocity along -y while it collapses sideways.  Limits from the json as in
    channel2d_periodic (limit 1, window shifted by `window` lo against the lattice)."""
    Lx, Lz, far = 2.0 * nx * lo, 2.0 * nz * lo, 50.0 * ny * lo
    i, j, m = np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij")
    order = np.lexsort((i.ravel(), j.ravel(), m.ravel()))
    pos = np.stack([(i.ravel()[order] + 0.5) * lo, (j.ravel()[order] + 0.5) * lo, (m.ravel()[order] + 0.5) * lo], axis=1)
    if jitter > 0.0:
        rng = np.random.default_rng(seed); pos = pos + rng.uniform(-jitter * lo, jitter * lo, size=pos.shape)
    tris = np.array([
        [[Lx, far, 0], [0, -far, 0], [0, far, 0]], [[0, -far, 0], [Lx, far, 0], [Lx, -far, 0]],          # floor
        [[0, far, Lz], [0, -far, 0], [0, -far, Lz]], [[0, -far, 0], [0, far, Lz], [0, far, 0]],          # x = 0
        [[Lx, far, 0], [Lx, -far, Lz], [Lx, -far, 0]], [[Lx, -far, Lz], [Lx, far, 0], [Lx, far, Lz]],    # x = Lx
    ], dtype=np.float64)
    H = nz * lo
    c0 = 15.0 * math.sqrt(H / 0.3)
    dt = 0.375 * lo / c0
    vmax = 2.0 * math.sqrt(9.81 * H)
    y0, y1 = 0.5 * lo, (ny - 0.5) * lo
    kw = dict(dim=3, lo=lo, dt=dt, wall_type=1, c0=c0, max_vel=vmax,
              dom_min=(0.0, y0 + (3.0 + window) * lo, 0.0), dom_max=(Lx, y1 - (3.0 - window) * lo, Lz))
    if incompressible:
        kw.update(mps_type=2, grad_type=0, grad_corr=True, shifting=0, pnd_type=0, dt=dt * 2.0)
    else:
        kw.update(mps_type=1, grad_type=3, grad_corr=False, shifting=2, pnd_type=2)
    kw.update(overrides)
    opts = base_options(**kw)
    opts["domain"]["boundary"] = {"type": 1, "limit": 1, "direction": {"x": False, "y": True, "z": False}}
    n = len(pos)
    vel = np.zeros((n, 3)); vel[:, 1] = -drift * vmax
    return Case(name=name or "channel3d_periodic_%dx%dx%d" % (nx, ny, nz), options=opts, ptype=np.zeros(n, dtype=np.int32),
                pos=np.ascontiguousarray(pos), vel=vel, press=np.zeros(n), tris=tris.astype(np.float32).astype(np.float64),
                meta={"H": H, "c0": c0})


def slosh3d(nx: int, ny: int, nz: int, *, lo: float = 0.0093, name: str | None = None, jitter: float = 0.0, **overrides) -> Case:
    """Sloshing tank with MOVING polygon walls (option set of input/sloshing_3D_bulian_lo0p00930.json, BASELINE configs[3]):
    three meshes as in the reference input — rigid and deformable are single far-away triangles, the tank is the FORCED mesh
    that PolygonMesh::updateForcedPolygonMesh rotates about (0.45, 0) every step (src/PolygonMesh.cpp:174-366)."""
    tank = (0.9, ny * lo, 2.5 * nz * lo)
    c = dam3d(nx, ny, nz, lo=lo, tank=tank, jitter=jitter, name=name or "slosh3d", offset=(0.45 - 0.5 * nx * lo, 0.0, 0.0), **overrides)
    far = np.array([[[0.0, 0.0, -9.0], [0.01, 0.0, -9.0], [0.0, 0.01, -9.0]]])
    c.forced_tris = c.tris; c.tris = far; c.deformable_tris = far + np.array([0.0, 0.0, -1.0])
    c.options["flags"].update(fem_MESH=True, forced_MESH=True)
    return c


# rheological block of input/2D_dam_fraccarollo_lo02p50e-03_WC.json (BASELINE.json configs[0])
FRACCAROLLO_RHEO = {"power_law_index": 1.2, "consistency_index": 0.03, "phi": {"lower": 0.4, "wall": 0.4, "bed": 0.4, "second": 0.611},
                    "cohes_coeff": 0.0, "fraction_method": 2, "grain_size": 0.0035, "I0": 0.75, "mm": 300.0, "stress_calc_method": 1,
                    "viscosity": {"iter_num": 1, "error": 0.0, "average": 0.0}, "drag_coeff": 0.47, "volume_fraction": {"min": 0.25, "max": 0.69}}


def dam2d_erosion(nx: int, nz: int, bed: int, *, lo: float = 0.0025, tank_nx: int | None = None, tank_nz: int | None = None,
                  jitter: float = 0.0, seed: int = 20261104, name: str | None = None, **overrides) -> Case:
    """Erosional dam break over a mobile bed (Fraccarollo & Capart 2002; option set of input/2D_dam_fraccarollo_lo02p50e-03_WC.json:
    2D, particle walls, WC-MPS, NON-NEWTONIAN two-phase rheology).  A sediment bed of `bed` layers (grid id 1: the reader turns
    it into fluid particles of phase 2, src/MpsInputOutput.cpp:839-847) covers the tank floor; a water column of nx x nz
    particles stands on its left end."""
    tank_nx = tank_nx or int(round(nx * 2.5))
    tank_nz = tank_nz or (bed + 2 * nz)
    rng = np.random.default_rng(seed)
    P, T = [], []
    for jy in range(-3, tank_nz):
        for ix in range(-3, tank_nx + 3):
            inside = 0 <= ix < tank_nx and jy >= 0
            x = ((ix + 0.5) * lo, (jy + 0.5) * lo, 0.0)
            if inside:
                t = 1 if jy < bed else (0 if (ix < nx and jy < bed + nz) else None)
                if t is not None:
                    if jitter > 0.0:
                        d = rng.uniform(-jitter * lo, jitter * lo, size=2); x = (x[0] + d[0], x[1] + d[1], 0.0)
                    P.append(x); T.append(t)
                continue
            layer = max(-ix, ix - tank_nx + 1, -jy)
            P.append(x); T.append(2 if layer == 1 else 3)
    pos = np.array(P); ptype = np.array(T, dtype=np.int32)
    H = nz * lo
    kw = dict(dim=2, lo=lo, dt=0.02 * lo, wall_type=0, mps_type=1, c0=20.0, re_small=2.1, re_large=3.1, gravity=(0.0, -9.81, 0.0),
              grad_type=3, grad_corr=False, fs_type=0, fs_pnd=0.97, max_vel=4.0, shifting=2, pnd_type=2, fluid_type=1,
              rheo=FRACCAROLLO_RHEO, dom_min=(-4 * lo, -4 * lo, 0.0), dom_max=((tank_nx + 4) * lo, (tank_nz + 1) * lo, 0.0))
    kw.update(overrides)
    n = len(pos)
    return Case(name=name or "dam2d_erosion_%dx%d_%d" % (nx, nz, bed), options=base_options(**kw), ptype=ptype, pos=pos,
                vel=np.zeros((n, 3)), press=np.zeros(n), tris=None, meta={"H": H})


def dam3d_erosion(nx: int, ny: int, nz: int, bed: int, *, lo: float = 0.005, tank_nx: int | None = None, jitter: float = 0.0, seed: int = 20261105,
                  name: str | None = None, **overrides) -> Case:
    """3D erosional dam break in a POLYGON tank (option set of input/damErosion3D_WC.json: polygon walls, WC-MPS, non-Newtonian
    two-phase rheology — the configuration whose near-wall particles go through calcWall{Slip,NoSlip}ViscosityInteractionVal,
    src/MpsViscosity.cpp:703-1526).  A sediment bed of `bed` layers (grid id 1 = fluid of phase 2) covers the tank floor, a water column
    of nx x ny x nz particles stands on its left end; the tank is ny particles wide."""
    tank_nx = tank_nx or int(round(nx * 2.5))
    rng = np.random.default_rng(seed)
    P, T = [], []
    for m in range(bed + nz):
        for j in range(ny):
            for i in range(tank_nx):
                t = 1 if m < bed else (0 if i < nx else None)
                if t is None:
                    continue
                x = np.array([(i + 0.5) * lo, (j + 0.5) * lo, (m + 0.5) * lo])
                if jitter > 0.0:
                    x = x + rng.uniform(-jitter * lo, jitter * lo, size=3)
                P.append(x); T.append(t)
    pos = np.array(P); ptype = np.array(T, dtype=np.int32)
    tank = (tank_nx * lo, ny * lo, 2.0 * (bed + nz) * lo)
    kw = dict(dim=3, lo=lo, dt=0.02 * lo, wall_type=1, mps_type=1, c0=20.0, grad_type=3, grad_corr=False, fs_type=1, max_vel=4.0, shifting=2, pnd_type=2,
              fluid_type=1, rheo=FRACCAROLLO_RHEO, dom_min=(0.0, 0.0, 0.0), dom_max=tank)
    kw.update(overrides)
    n = len(pos)
    return Case(name=name or "dam3d_erosion_%dx%dx%d_%d" % (nx, ny, nz, bed), options=base_options(**kw), ptype=ptype, pos=np.ascontiguousarray(pos),
                vel=np.zeros((n, 3)), press=np.zeros(n), tris=box_triangles(*tank).astype(np.float32).astype(np.float64), meta={"H": nz * lo})


# ---------------------------------------------------------------------------------------------- files
def write_grid(path: str, ptype, pos, vel, press) -> None:
    """.grid: '0', N, then 'id x y z vx vy vz p pav' rows (reference README.md:65-74).  %.17g round-trips doubles."""
    with open(path, "w") as fh:
        fh.write("0\n%d\n" % len(ptype))
        for t, p, v, pr in zip(ptype, pos, vel, press):
            fh.write("%d %.17g %.17g %.17g %.17g %.17g %.17g %.17g 0\n" % (t, p[0], p[1], p[2], v[0], v[1], v[2], pr))


def read_grid(path: str):
    with open(path) as fh:
        toks = fh.read().split()
    n = int(toks[1])
    a = np.array(toks[2:2 + 9 * n], dtype=np.float64).reshape(n, 9)
    return a[:, 0].astype(np.int32), a[:, 1:4].copy(), a[:, 4:7].copy(), a[:, 7].copy()


def write_case(case: Case, root: str) -> str:
    """Lay the case out under `root` the way the reference expects it (cwd/input/<name>.json, hard-coded prefix
    src/MpsInputOutput.cpp:63,99).  Returns the JSON path."""
    os.makedirs(os.path.join(root, "input", "grid"), exist_ok=True)
    os.makedirs(os.path.join(root, "input", "stl"), exist_ok=True)
    os.makedirs(os.path.join(root, "output"), exist_ok=True)
    opts = json.loads(json.dumps(case.options))
    opts["pathNames"]["particle_grid_file"] = "input/grid/%s.grid" % case.name
    opts["pathNames"]["mesh_rigid_file"] = "input/stl/%s.stl" % case.name
    opts["pathNames"]["vtu_output_folder"] = "output/%s" % case.name
    write_grid(os.path.join(root, opts["pathNames"]["particle_grid_file"]), case.ptype, case.pos, case.vel, case.press)
    if case.tris is not None:
        write_binary_stl(os.path.join(root, opts["pathNames"]["mesh_rigid_file"]), case.tris)
    for key, tris in (("mesh_deformable_file", case.deformable_tris), ("mesh_forced_file", case.forced_tris)):
        if tris is not None:
            opts["pathNames"][key] = "input/stl/%s_%s.stl" % (case.name, key.split("_")[1])
            write_binary_stl(os.path.join(root, opts["pathNames"][key]), tris)
    jp = os.path.join(root, "input", case.name + ".json")
    with open(jp, "w") as fh:
        json.dump(opts, fh, indent=1)
    return jp
