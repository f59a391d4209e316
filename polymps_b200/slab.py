"""Slab decomposition of the particle hot path over the GPUs of one box.

The reference is a single-process OpenMP code: there is no distributed layer to mirror.  What this module adds on top
of the C ABI is the host side of the decomposition `north_star` asks for — one process per GPU (`torch.distributed`,
NCCL over NVLink; gloo in the CPU tests), the domain cut into x-slabs on BUCKET-COLUMN boundaries of the reference's
bucket grid (src/MpsBucket.cpp:20-67), so bucket ids stay globally consistent:

* every rank OWNS the particles whose x lies in its slab [xlo, xhi) and keeps HALO COPIES of the neighbouring ranks'
  particles within one search radius W = bucketSide of its cuts — exactly what the 27-bucket search can reach
  (src/MpsBucket.cpp:23).  Local particle table: [owned | halo from the left | halo from the right | unused (ghost)].
* `repartition()` runs whenever the engine wants to rebuild its bucket order (same step on every rank: the decision is
  taken from all-reduced maxima): owners decide by position where each particle lives now, MIGRANTS move to the
  neighbouring rank with their whole persistent state, and the halo sets are re-selected.  Both sides derive the
  per-step send / receive lists from the order of these messages — no index lists cross the wire afterwards.
* between rebuilds the particle sets are frozen; after every phase of the fused step (`pmps_step_phase`) the fields that
  phase produced are refreshed on the halo copies: pack kernel -> `isend/irecv` with the two neighbours -> unpack kernel
  (8 small messages per step and neighbour, 8-56 B per halo particle).
* velMax (src/MpsParticleVelPos.cpp:91), the displacement since the rebuild and the largest shift are MAX-all-reduced
  once per step (three doubles, in place in device memory).

There is no data-path collective besides those in the explicit / weakly-compressible schemes: the path shards (`"scaling":
"weak"` in bench.py).  The incompressible scheme adds the exchanges of its Krylov solve: the conjugate-gradient dot products
are SUM-all-reduced and the halo copies of the search direction refreshed before every matrix-vector product; libpmps calls
back for both (`pmps_slab_set_hooks`).

`SlabDomain` holds the partition / exchange logic over an abstract particle-table backend, so the CPU tests run the very
same code over gloo with a host backend (tests/test_slab_host.py); `SlabEngine` binds it to libpmps.
"""
from __future__ import annotations

import ctypes
import os
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch
import torch.distributed as dist

from .state import F64, I32, U8, VEC, PmpsHostArrays

HF_POS, HF_VEL, HF_PRESS, HF_PNDI, HF_CONC, HF_PNDKI, HF_BC, HF_PPE_X, HF_PPE_P = 1, 2, 4, 8, 16, 32, 64, 128, 256
MAX_PLANES = 14


def halo_planes(mask: int) -> int:
    return (4 if mask & HF_POS else 0) + (3 if mask & HF_VEL else 0) + sum(1 for b in (HF_PRESS, HF_PNDI, HF_CONC, HF_PNDKI, HF_BC, HF_PPE_X, HF_PPE_P) if mask & b)


# State that lives from one step into the next (everything else is recomputed inside a step before it is read): what a
# migrating particle takes along and what a new halo copy is created with.
TABLE_FIELDS = ["particleType", "PTYPE", "particleBC", "numNeigh", "nearMeshType", "numNeighWallContribution", "particleNearWall",
                "press", "pndi", "pndki", "pndski", "pndSmall", "RHO", "MEU", "distParticleWall2", "pndWallContribution", "concentration", "pressurePPE",
                "pos", "vel", "acc", "mirrorParticlePos", "particleAtWallPos", "polygonNormal"]


def field_width(name: str) -> int:
    return 3 if name in VEC else 1


def field_dtype(name: str) -> torch.dtype:
    return torch.int32 if name in I32 else torch.uint8 if name in U8 else torch.float64


# ------------------------------------------------------------------------------------------------ partition (host logic)
def choose_cuts(col_counts: Sequence[int], world: int, min_cols: int = 3) -> List[int]:
    """Bucket-column indices c_0 = 0 < c_1 < ... < c_world = nCols: rank r owns columns [c_r, c_{r+1}).  Balanced by the
    particle count per column (prefix sums), every slab at least `min_cols` columns wide (a slab must be wider than the
    two halos that reach into it)."""
    cnt = np.asarray(col_counts, dtype=np.int64)
    ncol = len(cnt)
    if world * min_cols > ncol:
        raise ValueError("%d bucket columns cannot be cut into %d slabs of at least %d columns" % (ncol, world, min_cols))
    pre = np.concatenate([[0], np.cumsum(cnt)])
    total = int(pre[-1])
    cuts = [0]
    for r in range(1, world):
        target = total * r / world
        c = int(np.searchsorted(pre, target, side="left"))
        if c > 0 and abs(pre[c - 1] - target) <= abs(pre[min(c, ncol)] - target):
            c -= 1
        c = max(c, cuts[-1] + min_cols)
        c = min(c, ncol - (world - r) * min_cols)
        cuts.append(c)
    cuts.append(ncol)
    return cuts


def column_of(x: torch.Tensor, dom_min_x: float, inv_bucket_side: float) -> torch.Tensor:
    """Bucket column of a position, the reference's expression (src/MpsBucket.cpp:333): (int)((x - domainMinX) * invBucketSide)."""
    return ((x - dom_min_x) * inv_bucket_side).to(torch.int64)


class TableBackend:
    """What SlabDomain needs from the thing that holds a rank's particle table (libpmps on a GPU; a host table in tests)."""
    n0: int
    device: torch.device

    def download_table(self, fields) -> Dict[str, torch.Tensor]:   # name -> [n0] or [n0, 3], local order
        raise NotImplementedError

    def upload_table(self, table: Dict[str, torch.Tensor], n_live: int, n_own: int) -> None:
        raise NotImplementedError

    def halo_pack(self, mask: int, index: torch.Tensor, out: torch.Tensor) -> None:   # out: [planes * n] float64
        raise NotImplementedError

    def halo_unpack(self, mask: int, first: int, n: int, buf: torch.Tensor) -> None:
        raise NotImplementedError

    def ghost_fill(self, name: str) -> float:
        return 0.0


class SlabDomain:
    """Partition state of one rank + the two exchanges (re-partition at rebuilds, field refresh between phases)."""

    def __init__(self, backend: TableBackend, rank: int, world: int, xlo: float, xhi: float, halo_width: float, group=None,
                 fields=TABLE_FIELDS, grid=None):
        """grid = (domainMinX, bucketSide, numBucketsX, cuts) enables RE-BALANCING: at every re-partition the cuts (bucket-column
        indices, identical on all ranks) move by at most ONE column towards the balanced partition of the current particle count
        per column — one column = one halo width, so every particle still only ever moves to a neighbouring rank."""
        self.b = backend; self.rank = rank; self.world = world; self.group = group
        self.xlo = float(xlo); self.xhi = float(xhi); self.W = float(halo_width)
        self.grid = grid; self.cuts = list(grid[3]) if grid else None; self.min_cols = 3
        if world > 1 and self.xhi - self.xlo <= 2.0 * self.W:
            raise ValueError("slab [%g, %g) is not wider than two halos of %g" % (xlo, xhi, halo_width))
        self.fields = list(fields)
        self.left = rank - 1 if rank > 0 else None
        self.right = rank + 1 if rank < world - 1 else None
        self.n_own = 0; self.n_live = 0
        self.gid: Optional[torch.Tensor] = None           # [n0] int64 global particle id, local order
        dev = backend.device
        self.send_idx = {"L": torch.zeros(0, dtype=torch.int32, device=dev), "R": torch.zeros(0, dtype=torch.int32, device=dev)}
        self.halo_first = {"L": 0, "R": 0}; self.halo_n = {"L": 0, "R": 0}
        self._buf = {}
        self.stage_host = world > 1 and dist.is_initialized() and dist.get_backend(group) == "gloo"
        self.stats = {"repartitions": 0, "migrants_out": 0, "halo": 0, "refreshes": 0, "refresh_bytes": 0}

    # -- wire helpers
    def _peer(self, side):
        return self.left if side == "L" else self.right

    def _global(self, r):
        return r if self.group is None else dist.get_global_rank(self.group, r)

    def _exchange(self, send: Dict[str, torch.Tensor], recv: Dict[str, torch.Tensor]) -> None:
        """One batched isend/irecv with both neighbours (either dict may lack a side at the ends of the row)."""
        if self.stage_host and self.b.device.type == "cuda":
            # gloo moves host memory only: used when two ranks share one GPU (single-GPU parity runs), never in the bench
            hs = {k: v.cpu() for k, v in send.items()}; hr = {k: torch.empty_like(v, device="cpu") for k, v in recv.items()}
            self._exchange_raw(hs, hr)
            for k, v in recv.items():
                if v.numel() > 0 and self._peer(k) is not None:
                    v.copy_(hr[k])
            return
        self._exchange_raw(send, recv)

    def _exchange_raw(self, send, recv) -> None:
        ops = []
        for side in ("L", "R"):
            p = self._peer(side)
            if p is None:
                continue
            if side in recv and recv[side].numel() > 0:
                ops.append(dist.P2POp(dist.irecv, recv[side], self._global(p), group=self.group))
            if side in send and send[side].numel() > 0:
                ops.append(dist.P2POp(dist.isend, send[side], self._global(p), group=self.group))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()

    # -- re-partition: migrants + new halo sets
    def repartition(self, gid: Optional[torch.Tensor] = None, n_own: Optional[int] = None) -> None:
        """Owners decide by x where each of their particles lives now.  The first call takes the rank's initial particles
        (all owned, any of them possibly outside the slab by less than one slab width) with their global ids."""
        b = self.b; dev = b.device; n0 = b.n0
        timing = os.environ.get("PMPS_SLAB_TIMING") is not None
        def tick(label, _t=[None]):
            if not timing:
                return
            import time
            if dev.type == "cuda":
                torch.cuda.synchronize(dev)
            now = time.perf_counter()
            if _t[0] is not None and self.rank == 0:
                print("[slab] repartition %-12s %.1f ms" % (label, 1e3 * (now - _t[0])), flush=True)
            _t[0] = now
        tick("start")
        if gid is not None:
            self.gid = gid.to(device=dev, dtype=torch.int64); self.n_own = int(n_own); self.n_live = int(n_own)
        elif hasattr(b, "download_rows") and os.environ.get("PMPS_SLAB_FULL_TABLE") is None:
            return self._repartition_rows(tick)
        T = b.download_table(self.fields)
        tick("download")
        nO = self.n_own
        x = T["pos"][:nO, 0]
        false = torch.zeros(nO, dtype=torch.bool, device=dev)
        if self.grid is not None and self.world > 1:
            self._rebalance(x if "particleType" not in T else x[T["particleType"][:nO] != int(b.ghost_fill("particleType"))])
            tick("rebalance")
        alive = T["particleType"][:nO] != int(b.ghost_fill("particleType")) if "particleType" in T else ~false   # removed by the out-of-domain test: dropped here
        goL = alive & (x < self.xlo) if self.left is not None else false
        goR = alive & (x >= self.xhi) if self.right is not None else false
        stay = alive & ~(goL | goR)
        nearL = stay & (x < self.xlo + self.W) if self.left is not None else false
        nearR = stay & (x >= self.xhi - self.W) if self.right is not None else false
        idx = {k: torch.nonzero(m, as_tuple=False).flatten() for k, m in (("stay", stay), ("goL", goL), ("goR", goR), ("nearL", nearL), ("nearR", nearR))}
        widths = [field_width(f) for f in self.fields]
        R = sum(widths) + 1   # + global id

        def records(ix):
            cols = [T[f][ix].to(torch.float64).reshape(len(ix), w) for f, w in zip(self.fields, widths)]
            cols.append(self.gid[ix].to(torch.float64).reshape(-1, 1))   # ids < 2^53: exact
            return torch.cat(cols, dim=1).contiguous()

        out = {"L": torch.cat([records(idx["goL"]), records(idx["nearL"])]), "R": torch.cat([records(idx["goR"]), records(idx["nearR"])])}
        cnt_out = {"L": torch.tensor([len(idx["goL"]), len(idx["nearL"])], dtype=torch.int64, device=dev),
                   "R": torch.tensor([len(idx["goR"]), len(idx["nearR"])], dtype=torch.int64, device=dev)}
        cnt_in = {"L": torch.zeros(2, dtype=torch.int64, device=dev), "R": torch.zeros(2, dtype=torch.int64, device=dev)}
        self._exchange(cnt_out, cnt_in)
        cin = {s: [int(v) for v in cnt_in[s].cpu()] for s in ("L", "R")}
        inc = {s: torch.empty((cin[s][0] + cin[s][1], R), dtype=torch.float64, device=dev) for s in ("L", "R")}
        tick("pack")
        self._exchange(out, inc)
        tick("exchange")
        mig = {s: inc[s][:cin[s][0]] for s in ("L", "R")}; hal = {s: inc[s][cin[s][0]:] for s in ("L", "R")}
        n_stay = len(idx["stay"]); mL, mR = cin["L"][0], cin["R"][0]
        new_own = n_stay + mL + mR
        nHL = len(idx["goL"]) + cin["L"][1]; nHR = len(idx["goR"]) + cin["R"][1]
        n_live = new_own + nHL + nHR
        if n_live > n0:
            raise RuntimeError("rank %d: %d particles (owned %d + halo %d) exceed the table capacity %d" % (self.rank, n_live, new_own, nHL + nHR, n0))
        # new table = [stay | migrants from L | migrants from R | halo L: my leavers to L, L's near | halo R: ... | ghosts]
        parts_local = [idx["stay"], None, None, idx["goL"], None, idx["goR"], None]
        parts_wire = [None, mig["L"], mig["R"], None, hal["L"], None, hal["R"]]
        newT = {}
        col = 0
        for f, w in zip(self.fields, widths):
            dt = field_dtype(f)
            segs = []
            for pl, pw in zip(parts_local, parts_wire):
                if pl is not None:
                    segs.append(T[f][pl].reshape(len(pl), w))
                else:
                    segs.append(pw[:, col:col + w].to(dt))
            fill = torch.full((n0 - n_live, w), b.ghost_fill(f), dtype=dt, device=dev)
            v = torch.cat(segs + [fill], dim=0)
            newT[f] = v.reshape(n0, 3).contiguous() if w == 3 else v.reshape(n0).contiguous()
            col += w
        gsegs = [self.gid[pl] if pl is not None else pw[:, col].to(torch.int64) for pl, pw in zip(parts_local, parts_wire)]
        self.gid = torch.cat(gsegs + [torch.full((n0 - n_live,), -1, dtype=torch.int64, device=dev)])
        # refresh lists.  The left neighbour's halo-R range is [its leavers to me = my migrants from L, in wire order | my
        # nearL, in wire order]; mine to the right likewise.  New local index of a staying particle = its rank among them.
        new_of_old = torch.cumsum(stay.to(torch.int64), 0) - 1
        i32 = lambda t: t.to(torch.int32).contiguous()
        ar = lambda a, n: torch.arange(a, a + n, dtype=torch.int64, device=dev)
        self.send_idx = {"L": i32(torch.cat([ar(n_stay, mL), new_of_old[idx["nearL"]]])),
                         "R": i32(torch.cat([ar(n_stay + mL, mR), new_of_old[idx["nearR"]]]))}
        self.halo_first = {"L": new_own, "R": new_own + nHL}; self.halo_n = {"L": nHL, "R": nHR}
        self.n_own = new_own; self.n_live = n_live
        tick("assemble")
        b.upload_table(newT, n_live, new_own)
        tick("upload")
        self.stats["repartitions"] += 1; self.stats["migrants_out"] += len(idx["goL"]) + len(idx["goR"]); self.stats["halo"] = nHL + nHR

    def _repartition_rows(self, tick) -> None:
        """The same re-partition with only the rows that change hands or places travelling between the engine and this code: the
        owned particles that stay keep their row; the rows of leavers (and of particles the out-of-domain test removed) are holes,
        filled by the migrants that arrive and — when more left than arrived — by the staying particles of the highest rows, so
        that the owned rows stay [0, n_own); the halo rows behind them are rewritten.  ~2 % of the table at 10 M particles per rank
        instead of two full round trips of all its fields."""
        b = self.b; dev = b.device; n0 = b.n0; nO = self.n_own
        i64 = lambda t: t.to(torch.int64)
        head = b.download_table(["pos", "particleType"])
        x = head["pos"][:nO, 0]
        false = torch.zeros(nO, dtype=torch.bool, device=dev)
        alive = head["particleType"][:nO] != int(b.ghost_fill("particleType"))
        tick("download")
        if self.grid is not None and self.world > 1:
            self._rebalance(x[alive])
            tick("rebalance")
        goL = alive & (x < self.xlo) if self.left is not None else false
        goR = alive & (x >= self.xhi) if self.right is not None else false
        stay = alive & ~(goL | goR)
        nearL = stay & (x < self.xlo + self.W) if self.left is not None else false
        nearR = stay & (x >= self.xhi - self.W) if self.right is not None else false
        idx = {k: torch.nonzero(m, as_tuple=False).flatten() for k, m in (("goL", goL), ("goR", goR), ("nearL", nearL), ("nearR", nearR))}
        n_stay = int(stay.sum())
        widths = [field_width(f) for f in self.fields]
        R = sum(widths) + 1   # + global id
        cnt_out = {"L": torch.tensor([len(idx["goL"]), len(idx["nearL"])], dtype=torch.int64, device=dev),
                   "R": torch.tensor([len(idx["goR"]), len(idx["nearR"])], dtype=torch.int64, device=dev)}
        cnt_in = {"L": torch.zeros(2, dtype=torch.int64, device=dev), "R": torch.zeros(2, dtype=torch.int64, device=dev)}
        self._exchange(cnt_out, cnt_in)
        cin = {s: [int(v) for v in cnt_in[s].cpu()] for s in ("L", "R")}
        mL, mR = cin["L"][0], cin["R"][0]
        new_own = n_stay + mL + mR
        nHL = len(idx["goL"]) + cin["L"][1]; nHR = len(idx["goR"]) + cin["R"][1]
        n_live = new_own + nHL + nHR
        if n_live > n0:
            raise RuntimeError("rank %d: %d particles (owned %d + halo %d) exceed the table capacity %d" % (self.rank, n_live, new_own, nHL + nHR, n0))
        # staying particles that sit above the new owned range move down into holes
        movers = torch.nonzero(stay[new_own:], as_tuple=False).flatten() + new_own if new_own < nO else torch.zeros(0, dtype=torch.int64, device=dev)
        # rows read from the engine: what goes on the wire (leavers, near-cut particles) and the movers
        rows_out = torch.cat([idx["goL"], idx["nearL"], idx["goR"], idx["nearR"], movers])
        Tr = b.download_rows(self.fields, rows_out.to(torch.int32).contiguous())
        rec = torch.cat([Tr[f].to(torch.float64).reshape(len(rows_out), w) for f, w in zip(self.fields, widths)] + [self.gid[rows_out].to(torch.float64).reshape(-1, 1)], dim=1).contiguous()
        o = [0, len(idx["goL"]), len(idx["goL"]) + len(idx["nearL"]), len(idx["goL"]) + len(idx["nearL"]) + len(idx["goR"]), len(rows_out) - len(movers), len(rows_out)]
        recs = {"goL": rec[o[0]:o[1]], "L": rec[o[0]:o[2]], "goR": rec[o[2]:o[3]], "R": rec[o[2]:o[4]], "movers": rec[o[4]:o[5]]}
        inc = {sd: torch.empty((cin[sd][0] + cin[sd][1], R), dtype=torch.float64, device=dev) for sd in ("L", "R")}
        tick("pack")
        self._exchange({"L": recs["L"].contiguous(), "R": recs["R"].contiguous()}, inc)
        tick("exchange")
        mig = {sd: inc[sd][:cin[sd][0]] for sd in ("L", "R")}; hal = {sd: inc[sd][cin[sd][0]:] for sd in ("L", "R")}
        # free rows of the new owned range, lowest first: the holes below new_own, then the rows above the old owned range
        holes = torch.nonzero(~stay[:min(new_own, nO)], as_tuple=False).flatten()
        free = torch.cat([holes, torch.arange(nO, new_own, dtype=torch.int64, device=dev)]) if new_own > nO else holes
        assert len(free) == mL + mR + len(movers), (len(free), mL, mR, len(movers))
        tgt_migL, tgt_migR, tgt_mov = free[:mL], free[mL:mL + mR], free[mL + mR:]
        # rows written: owned targets, the halo ranges, and the rows between the new and the old live count (they become ghosts)
        halo_rows = torch.arange(new_own, n_live, dtype=torch.int64, device=dev)
        dead_rows = torch.arange(n_live, max(self.n_live, n_live), dtype=torch.int64, device=dev)
        rows_in = torch.cat([tgt_migL, tgt_migR, tgt_mov, halo_rows, dead_rows])
        data = torch.cat([mig["L"], mig["R"], recs["movers"], recs["goL"], hal["L"], recs["goR"], hal["R"]], dim=0)
        newT = {}; col = 0
        for f, w in zip(self.fields, widths):
            dt = field_dtype(f)
            v = data[:, col:col + w].to(dt)
            if len(dead_rows):
                v = torch.cat([v, torch.full((len(dead_rows), w), b.ghost_fill(f), dtype=dt, device=dev)], dim=0)
            newT[f] = v.reshape(-1, 3).contiguous() if w == 3 else v.reshape(-1).contiguous()
            col += w
        if len(dead_rows) and "pos" in newT and hasattr(b, "_ghost_pos"):
            newT["pos"][len(data):] = torch.tensor(b._ghost_pos, dtype=torch.float64, device=dev)
        gid_new = data[:, col].to(torch.int64)
        self.gid[rows_in[:len(data)]] = gid_new
        if len(dead_rows):
            self.gid[dead_rows] = -1
        # refresh lists: the left neighbour's halo-R range is [its leavers to me = my migrants from L, in wire order | my nearL, in
        # wire order]; mine to the right likewise.  A staying particle keeps its row unless it moved down.
        new_of_old = torch.arange(nO, dtype=torch.int64, device=dev)
        new_of_old[movers] = tgt_mov
        i32 = lambda t: t.to(torch.int32).contiguous()
        self.send_idx = {"L": i32(torch.cat([tgt_migL, new_of_old[idx["nearL"]]])), "R": i32(torch.cat([tgt_migR, new_of_old[idx["nearR"]]]))}
        self.halo_first = {"L": new_own, "R": new_own + nHL}; self.halo_n = {"L": nHL, "R": nHR}
        self.n_own = new_own; self.n_live = n_live
        tick("assemble")
        b.upload_rows(newT, i32(rows_in), n_live, new_own)
        tick("upload")
        self.stats["repartitions"] += 1; self.stats["migrants_out"] += len(idx["goL"]) + len(idx["goR"]); self.stats["halo"] = nHL + nHR

    def _rebalance(self, x_alive: torch.Tensor) -> None:
        x0, side, ncol, _ = self.grid
        col = ((x_alive - x0) / side).to(torch.int64).clamp_(0, ncol - 1)
        hist = torch.bincount(col, minlength=ncol).to(torch.float64)
        self.all_sum(hist)
        target = choose_cuts(hist.cpu().numpy().astype(np.int64), self.world, self.min_cols)
        new = list(self.cuts)
        for r in range(1, self.world):
            new[r] += (target[r] > new[r]) - (target[r] < new[r])
        for r in range(1, self.world):
            new[r] = max(new[r], new[r - 1] + self.min_cols)
        for r in range(self.world - 1, 0, -1):
            new[r] = min(new[r], new[r + 1] - self.min_cols)
        if new != self.cuts:
            self.stats["cut_moves"] = self.stats.get("cut_moves", 0) + sum(1 for a, c in zip(new, self.cuts) if a != c)
        self.cuts = new
        self.xlo = x0 + new[self.rank] * side; self.xhi = x0 + new[self.rank + 1] * side

    # -- per-phase refresh of the halo copies
    def refresh(self, mask: int) -> None:
        if not mask or self.world == 1:
            return
        b = self.b; P = halo_planes(mask)
        send, recv = {}, {}
        for side in ("L", "R"):
            if self._peer(side) is None:
                continue
            ns, nr = int(self.send_idx[side].numel()), self.halo_n[side]
            sb = self._buffer("s" + side, MAX_PLANES * ns)[:P * ns]; rb = self._buffer("r" + side, MAX_PLANES * nr)[:P * nr]
            if ns:
                b.halo_pack(mask, self.send_idx[side], sb)
            send[side] = sb; recv[side] = rb
            self.stats["refresh_bytes"] += 8 * P * ns
        self._exchange(send, recv)
        for side in ("L", "R"):
            if side in recv and self.halo_n[side]:
                b.halo_unpack(mask, self.halo_first[side], self.halo_n[side], recv[side])
        self.stats["refreshes"] += 1

    def _buffer(self, key, n):
        t = self._buf.get(key)
        if t is None or t.numel() < n:
            t = torch.empty(max(n, 1), dtype=torch.float64, device=self.b.device); self._buf[key] = t
        return t

    def all_sum(self, t: torch.Tensor) -> torch.Tensor:
        if self.world > 1:
            if self.stage_host and t.device.type == "cuda":
                c = t.cpu(); dist.all_reduce(c, op=dist.ReduceOp.SUM, group=self.group); t.copy_(c)
            else:
                dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return t

    def all_max(self, t: torch.Tensor) -> torch.Tensor:
        if self.world > 1:
            if self.stage_host and t.device.type == "cuda":
                c = t.cpu(); dist.all_reduce(c, op=dist.ReduceOp.MAX, group=self.group); t.copy_(c)
            else:
                dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        return t


# ------------------------------------------------------------------------------------------------ libpmps binding
SUM_FN = ctypes.CFUNCTYPE(None, ctypes.c_void_p, ctypes.c_int)
REFRESH_FN = ctypes.CFUNCTYPE(None, ctypes.c_void_p, ctypes.c_uint)


class EngineBackend(TableBackend):
    """The rank's table lives in a libpmps handle; transfers are device-to-device on the engine's stream."""

    def __init__(self, eng, n0: int, device: torch.device):
        self.eng = eng; self.n0 = int(n0); self.device = device
        L = eng.lib
        L.pmps_download_device.argtypes = [ctypes.c_void_p, ctypes.POINTER(PmpsHostArrays)]
        L.pmps_upload_values_device.argtypes = [ctypes.c_void_p, ctypes.POINTER(PmpsHostArrays)]
        L.pmps_download_rows_device.argtypes = [ctypes.c_void_p, ctypes.POINTER(PmpsHostArrays), ctypes.c_int, ctypes.c_void_p]
        L.pmps_upload_rows_device.argtypes = [ctypes.c_void_p, ctypes.POINTER(PmpsHostArrays), ctypes.c_int, ctypes.c_void_p, ctypes.c_int]
        L.pmps_slab_begin.argtypes = [ctypes.c_void_p, ctypes.c_int]
        L.pmps_halo_pack.argtypes = [ctypes.c_void_p, ctypes.c_uint, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
        L.pmps_halo_unpack.argtypes = [ctypes.c_void_p, ctypes.c_uint, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]
        L.pmps_step_phase.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.POINTER(ctypes.c_uint)]
        L.pmps_slab_collect.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.pmps_slab_apply.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.pmps_slab_wants_rebuild.argtypes = [ctypes.c_void_p]
        L.pmps_slab_mid_check.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        L.pmps_slab_mid_rebuild.argtypes = [ctypes.c_void_p]
        L.pmps_slab_set_hooks.argtypes = [ctypes.c_void_p, ctypes.c_void_p, SUM_FN, REFRESH_FN, ctypes.c_void_p]
        ps = eng.ps
        self._ghost_pos = [ps.domainMaxX - ps.partDist, ps.domainMaxY - ps.partDist, 0.0 if int(ps.dim) == 2 else ps.domainMaxZ - ps.partDist]
        self._ghost = {"particleType": float(ps.ghost), "particleBC": float(ps.other), "distParticleWall2": 10e8 * ps.partDist}

    def ghost_fill(self, name):
        return self._ghost.get(name, 0.0)

    def _cstruct(self, table, n_live):
        c = PmpsHostArrays(); c.numParticles = int(n_live); c.numParticlesZero = self.n0
        for f, t in table.items():
            assert t.is_contiguous() and t.device == self.device and t.dtype == field_dtype(f), f
            ct = ctypes.c_int32 if f in I32 else ctypes.c_uint8 if f in U8 else ctypes.c_double
            setattr(c, f, ctypes.cast(ctypes.c_void_p(t.data_ptr()), ctypes.POINTER(ct)))
        return c

    def download_table(self, fields):
        T = {f: torch.empty((self.n0, 3) if f in VEC else (self.n0,), dtype=field_dtype(f), device=self.device) for f in fields}
        c = self._cstruct(T, 0)
        self.eng._chk(self.eng.lib.pmps_download_device(self.eng.h, ctypes.byref(c)))
        return T

    def upload_table(self, table, n_live, n_own, same_rows=False):
        """same_rows: the table the engine handed out, with new values (a host-driven loop between two re-partitions) — the engine keeps
        its bucket order and lists and lets the displacement trackers decide (pmps_upload_values_device)."""
        if n_live < self.n0 and "pos" in table:   # unused tail: the reference's ghost (src/MpsParticle.cpp:300-340)
            table["pos"][n_live:] = torch.tensor(self._ghost_pos, dtype=torch.float64, device=self.device)
        c = self._cstruct(table, n_live)
        self.eng.remember_capacity(self.n0)
        up = self.eng.lib.pmps_upload_values_device if same_rows else self.eng.lib.pmps_upload_device
        self.eng._chk(up(self.eng.h, ctypes.byref(c)))
        self.eng._chk(self.eng.lib.pmps_slab_begin(self.eng.h, int(n_own)))
        self.eng._chk(self.eng.lib.pmps_synchronize(self.eng.h))   # the table tensors may be freed after this returns

    # row sets: only the rows a re-partition touches travel (SlabDomain._repartition_rows)
    def download_rows(self, fields, rows: torch.Tensor) -> Dict[str, torch.Tensor]:
        n = int(rows.numel())
        T = {f: torch.empty((n, 3) if f in VEC else (n,), dtype=field_dtype(f), device=self.device) for f in fields}
        if n:
            c = self._cstruct(T, 0); c.numParticlesZero = n
            self.eng._chk(self.eng.lib.pmps_download_rows_device(self.eng.h, ctypes.byref(c), n, rows.data_ptr()))
        return T

    def upload_rows(self, table: Dict[str, torch.Tensor], rows: torch.Tensor, n_live: int, n_own: int) -> None:
        n = int(rows.numel())
        c = self._cstruct(table, n_live); c.numParticlesZero = n
        self.eng._chk(self.eng.lib.pmps_upload_rows_device(self.eng.h, ctypes.byref(c), n, rows.data_ptr(), int(n_live)))
        self.eng._chk(self.eng.lib.pmps_slab_begin(self.eng.h, int(n_own)))
        self.eng._chk(self.eng.lib.pmps_synchronize(self.eng.h))   # the row tensors may be freed after this returns

    def halo_pack(self, mask, index, out):
        self.eng._chk(self.eng.lib.pmps_halo_pack(self.eng.h, mask, index.data_ptr(), int(index.numel()), out.data_ptr()))

    def halo_unpack(self, mask, first, n, buf):
        self.eng._chk(self.eng.lib.pmps_halo_unpack(self.eng.h, mask, int(first), int(n), buf.data_ptr()))


class SlabEngine:
    """One rank of a slab-decomposed simulation: `Engine` + `SlabDomain`.

    `local` holds this rank's INITIAL particles (numpy: ptype, pos, vel, press, gid) — any set whose particles are not
    farther than one slab from where they belong; the first re-partition moves strays and builds the halos."""

    def __init__(self, ps, meshes, local: dict, cuts_x: Sequence[float], rank: int, world: int, device: int = 0, group=None,
                 capacity: Optional[int] = None, shifting_jacobi: bool = True, cuts: Optional[Sequence[int]] = None):
        """cuts = the bucket-column indices behind cuts_x: given, the slabs re-balance themselves as the particles move."""
        from .engine import Engine
        self.ps = ps; self.rank = rank; self.world = world
        self.dev = torch.device("cuda", device)
        self.eng = Engine(ps, meshes, device=device, shifting_jacobi=shifting_jacobi)
        self.stream = torch.cuda.ExternalStream(self.eng.stream, device=self.dev)
        n = len(local["ptype"])
        n0 = int(capacity) if capacity else int(n * 1.25) + 200000
        self.backend = EngineBackend(self.eng, n0, self.dev)
        grid = (ps.domainMinX, ps.bucketSide, ps.numBucketsX, list(cuts)) if cuts is not None else None
        self.dom = SlabDomain(self.backend, rank, world, cuts_x[rank], cuts_x[rank + 1], ps.bucketSide * (1.0 + 1e-9), group=group, grid=grid)
        self.scal = torch.zeros(4, dtype=torch.float64, device=self.dev)
        self.nph = int(self.eng.lib.pmps_step_phases())
        self.midstep = os.environ.get("PMPS_SLAB_STEPSTART") is None   # rebuild decision after the prediction (default) or at the step start (round 1)
        # the Poisson solve of the incompressible scheme calls back for its all-reduces and the halo refresh of the search direction
        self.work = torch.zeros(8, dtype=torch.float64, device=self.dev)
        def _sum(user, n):
            self.dom.all_sum(self.work[:n])

        def _refresh(user, fields):
            self.dom.refresh(int(fields))
        self._sum_cb = SUM_FN(_sum); self._refresh_cb = REFRESH_FN(_refresh)
        self.eng._chk(self.eng.lib.pmps_slab_set_hooks(self.eng.h, self.work.data_ptr(), self._sum_cb, self._refresh_cb, None))
        with torch.cuda.stream(self.stream):
            T = {}
            for f in TABLE_FIELDS:
                T[f] = torch.zeros((n0, 3) if f in VEC else (n0,), dtype=field_dtype(f), device=self.dev)
                g = self.backend.ghost_fill(f)
                if g:
                    T[f].fill_(g)
            up = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a)).to(device=self.dev, dtype=dt)
            T["particleType"][:n] = up(local["ptype"], torch.int32)
            T["pos"][:n] = up(local["pos"], torch.float64); T["vel"][:n] = up(local["vel"], torch.float64)
            T["press"][:n] = up(local["press"], torch.float64)
            if ps.fluidType != 0:
                raise NotImplementedError("non-Newtonian fluids are out of scope this round (SURVEY §8f row 1)")
            T["RHO"][:n] = ps.DNS_FL1; T["PTYPE"][:n] = 1; T["MEU"][:n] = ps.KNM_VS1 * ps.DNS_FL1   # src/MpsInputOutput.cpp:800-857
            self.backend.upload_table(T, n, n)
            self.dom.repartition(gid=torch.from_numpy(np.ascontiguousarray(local["gid"])), n_own=n)

    # -- stepping
    def _stage(self, name, mask=0):
        getattr(self.eng, name)()
        self.dom.refresh(mask)

    def init_step0(self):
        """The 0-th step (src/main.cpp:121-180) with the halo copies refreshed after every pass that neighbours read."""
        poly = self.ps.wallType == 1
        with torch.cuda.stream(self.stream):
            self._stage("update_buckets"); self._stage("check_out_domain")
            if poly:
                self._stage("wall_closest_point"); self._stage("wall_npcd")
            m = HF_PNDI | HF_PNDKI | HF_PRESS
            self._stage("initial_pnd", m)
            if poly:
                self._stage("mean_wall_pnd", m)
            self._stage("mean_pnd", m); self._stage("mean_neigh_fluid_pnd", m)
            if self.ps.freeSurfType == 2:   # ARC free-surface detection: the particle normals first (src/main.cpp:170-176); they read positions only
                self._stage("calc_normals")
                if poly:
                    self._stage("calc_wall_normals")
            self._stage("update_bc", HF_BC); self._stage("calc_press", m | HF_PPE_X)

    def step(self, n: int = 1):
        L = self.eng.lib; h = self.eng.h
        with torch.cuda.stream(self.stream):
            for _ in range(n):
                if L.pmps_slab_wants_rebuild(h) == 1:
                    self.dom.repartition()
                m = ctypes.c_uint(0)
                timing = os.environ.get("PMPS_SLAB_TIMING") is not None and self.rank == 0
                for ph in range(self.nph):
                    if timing:
                        import time
                        torch.cuda.synchronize(self.dev); t0 = time.perf_counter()
                    self.eng._chk(L.pmps_step_phase(h, ph, ctypes.byref(m)))
                    if timing:
                        torch.cuda.synchronize(self.dev); t1 = time.perf_counter()
                    self.dom.refresh(int(m.value))
                    if timing:
                        torch.cuda.synchronize(self.dev); t2 = time.perf_counter()
                        print("[slab] phase %d: kernels %.1f ms, refresh %.1f ms" % (ph, 1e3 * (t1 - t0), 1e3 * (t2 - t1)), flush=True)
                    if ph == 0 and self.midstep:
                        # the rebuild decision, where a single GPU takes it: after the prediction, from the maxima over all ranks
                        self.eng._chk(L.pmps_slab_collect(h, self.scal.data_ptr()))
                        self.dom.all_max(self.scal)
                        if self.eng._chk(L.pmps_slab_mid_check(h, self.scal.data_ptr())) == 1:
                            self.dom.repartition()
                            self.eng._chk(L.pmps_slab_mid_rebuild(h))
                self.eng._chk(L.pmps_slab_collect(h, self.scal.data_ptr()))
                self.dom.all_max(self.scal)
                self.eng._chk(L.pmps_slab_apply(h, self.scal.data_ptr()))

    # -- results
    @property
    def n_own(self):
        return self.dom.n_own

    def download_owned(self, fields) -> Dict[str, np.ndarray]:
        """Owned particles' fields (numpy) + "gid"."""
        with torch.cuda.stream(self.stream):
            T = self.backend.download_table(list(fields))
            self.eng.synchronize()
            out = {f: T[f][:self.dom.n_own].cpu().numpy() for f in fields}
            out["gid"] = self.dom.gid[:self.dom.n_own].cpu().numpy()
        return out

    def scalars(self):
        return self.eng.scalars()

    def close(self):
        """Drops every tensor that lives on the engine's stream BEFORE the handle (and with it the stream) is destroyed."""
        import gc
        torch.cuda.synchronize(self.dev)
        self.dom._buf.clear(); self.dom.gid = None; self.dom.send_idx = {}; self.scal = None; self.work = None
        gc.collect(); torch.cuda.empty_cache()
        self.eng.close()
