"""Host side of the slab decomposition (polymps_b200/slab.py) over gloo, world sizes 2 and 3, on the CPU.

A host table backend stands in for libpmps (the device half — pack / unpack kernels, owned-only bookkeeping — is covered by
the GPU tests); the partition, migration, halo selection and send/receive list logic under test is the product code.
Checked after every re-partition and refresh:
  * every particle is owned by exactly one rank (none lost, none duplicated) even when particles sit on a cut;
  * a rank's halo copies are exactly the neighbours' owned particles within W of its cuts, plus its own leavers;
  * after a refresh every halo copy carries its owner's current values (the send / receive orders agree);
  * choose_cuts balances particle counts and respects the minimum slab width.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

from polymps_b200 import slab  # noqa: E402
from polymps_b200.slab import HF_POS, HF_VEL, HF_PRESS, HF_PNDI, HF_CONC, HF_PNDKI, SlabDomain, TableBackend, choose_cuts  # noqa: E402

FIELDS = ["particleType", "particleBC", "press", "pndi", "pndki", "concentration", "pos", "vel"]
PLANES = [(HF_POS, "pos", 4), (HF_VEL, "vel", 3), (HF_PRESS, "press", 1), (HF_PNDI, "pndi", 1), (HF_CONC, "concentration", 1), (HF_PNDKI, "pndki", 1)]


class HostTable(TableBackend):
    """Particle table in host memory, local order == storage order (slotOfHost = identity)."""

    def __init__(self, n0):
        self.n0 = n0; self.device = torch.device("cpu")
        self.T = {f: torch.zeros((n0, 3) if slab.field_width(f) == 3 else (n0,), dtype=slab.field_dtype(f)) for f in FIELDS}
        self.n_live = 0; self.n_own = 0

    def ghost_fill(self, name):
        return -1.0 if name == "particleType" else 0.0

    def download_table(self, fields):
        return {f: self.T[f].clone() for f in fields}

    def upload_table(self, table, n_live, n_own):
        for f, t in table.items():
            self.T[f] = t.clone()
        self.n_live = n_live; self.n_own = n_own

    def halo_pack(self, mask, index, out):
        n = index.numel(); ix = index.long(); o = 0
        for bit, f, w in PLANES:
            if not mask & bit:
                continue
            if f == "pos":
                v = torch.cat([self.T["pos"][ix].T.reshape(-1), self.T["particleType"][ix].double()])
            elif w == 3:
                v = self.T[f][ix].T.reshape(-1)
            else:
                v = self.T[f][ix]
            out[o:o + w * n] = v; o += w * n

    def halo_unpack(self, mask, first, n, buf):
        assert first >= self.n_own, "unpack onto an owned particle"
        o = 0
        for bit, f, w in PLANES:
            if not mask & bit:
                continue
            v = buf[o:o + w * n].reshape(w, n); o += w * n
            if f == "pos":
                self.T["pos"][first:first + n] = v[:3].T; self.T["particleType"][first:first + n] = v[3].to(torch.int32)
            elif w == 3:
                self.T[f][first:first + n] = v.T
            else:
                self.T[f][first:first + n] = v[0]


def _gather(obj, world):
    out = [None] * world
    dist.all_gather_object(out, obj)
    return out


def _worker(rank, world, port, n_total, seed, rounds, rebalance=False):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(seed)
        L = 30.0 * world; W = 2.5
        cuts = np.linspace(0.0, L, world + 1)
        pos = rng.uniform(0.0, 1.0, size=(n_total, 3)) * np.array([L, 5.0, 5.0])
        pos[:world - 1, 0] = cuts[1:-1]            # particles exactly on the cuts: x >= xhi belongs to the right
        pos[world - 1:2 * (world - 1), 0] = np.nextafter(cuts[1:-1], -np.inf)
        mine = np.nonzero((pos[:, 0] >= cuts[rank]) & (pos[:, 0] < cuts[rank + 1]))[0]
        n = len(mine); n0 = n_total
        be = HostTable(n0)
        be.T["pos"][:n] = torch.from_numpy(pos[mine]); be.T["particleType"][:] = -1; be.T["particleType"][:n] = 0
        be.T["press"][:n] = torch.from_numpy(mine.astype(np.float64)) * 10.0
        be.n_live = be.n_own = n
        grid = (0.0, W, int(round(L / W)), [int(round(c / W)) for c in cuts]) if rebalance else None
        dom = SlabDomain(be, rank, world, cuts[rank], cuts[rank + 1], W, fields=FIELDS, grid=grid)
        if rebalance:
            vel_bias = 12.0   # everybody drifts to the right: the cuts must follow
        dom.repartition(gid=torch.from_numpy(mine.astype(np.int64)), n_own=n)
        vel = rng.normal(0.0, 1.0, size=(n_total, 3)) * np.array([0.6, 0.05, 0.05])   # global velocity field by gid
        if rebalance:
            vel[:, 0] = np.abs(vel[:, 0]) * 0.2 + vel_bias * (pos[:, 0] < 0.5 * L)   # the left half streams right
        for rd in range(rounds):
            # ---- invariants of the partition
            own_g = dom.gid[:dom.n_own].numpy(); own_x = be.T["pos"][:dom.n_own, 0].numpy()
            allown = _gather((own_g, own_x), world)
            if rank == 0:
                g = np.concatenate([a for a, _ in allown])
                assert len(g) == n_total and len(np.unique(g)) == n_total, "round %d: %d owned ids, %d unique, %d expected" % (rd, len(g), len(np.unique(g)), n_total)
            for side, nb in (("L", rank - 1), ("R", rank + 1)):
                if nb < 0 or nb >= world:
                    assert dom.halo_n[side] == 0
                    continue
                f0, nh = dom.halo_first[side], dom.halo_n[side]
                hg = set(dom.gid[f0:f0 + nh].numpy().tolist())
                ng, nx = allown[nb]
                cut = dom.xlo if side == "L" else dom.xhi
                near = set(ng[np.abs(nx - cut) < W * 0.999].tolist())
                assert near <= hg, "round %d rank %d side %s: %d of the neighbour's near particles have no halo copy" % (rd, rank, side, len(near - hg))
                assert hg <= set(ng.tolist()), "halo copy of a particle the neighbour does not own"
            assert (be.T["particleType"][dom.n_live:] == -1).all() and (be.T["particleType"][:dom.n_live] == 0).all()
            # ---- a "phase": owners change their fields (functions of gid and round), halo copies go stale, refresh restores them
            g_own = dom.gid[:dom.n_own].double()
            be.T["press"][:dom.n_own] = g_own * 10.0 + rd; be.T["pndi"][:dom.n_own] = g_own * 0.5 - rd
            be.T["concentration"][:dom.n_own] = g_own + 0.25 * rd; be.T["pndki"][:dom.n_own] = -g_own
            be.T["vel"][:dom.n_own] = torch.from_numpy(vel[dom.gid[:dom.n_own].numpy()]) * ((1 + rd) if not rebalance else 1.0)
            be.T["pos"][:dom.n_own] += 0.02 * be.T["vel"][:dom.n_own]
            for f in ("press", "pndi", "concentration", "pndki"):
                be.T[f][dom.n_own:dom.n_live] = -777.0
            be.T["pos"][dom.n_own:dom.n_live] += 99.0
            for mask in (HF_POS | HF_VEL, HF_PRESS | HF_PNDI, HF_CONC, HF_PNDKI):
                dom.refresh(mask)
            allstate = _gather({int(k): (p.numpy().copy(), float(pr), float(c)) for k, p, pr, c in
                                zip(dom.gid[:dom.n_own], be.T["pos"][:dom.n_own], be.T["press"][:dom.n_own], be.T["concentration"][:dom.n_own])}, world)
            truth = {}
            for d in allstate:
                truth.update(d)
            for k in range(dom.n_own, dom.n_live):
                p, pr, c = truth[int(dom.gid[k])]
                assert np.array_equal(be.T["pos"][k].numpy(), p) and float(be.T["press"][k]) == pr and float(be.T["concentration"][k]) == c, \
                    "round %d rank %d: halo copy %d (gid %d) is stale after the refresh" % (rd, rank, k, int(dom.gid[k]))
                assert float(be.T["pndki"][k]) == -float(dom.gid[k]) and int(be.T["particleType"][k]) == 0
            # ---- next rebuild: particles have moved (some across the cuts)
            dom.repartition()
        assert dom.stats["repartitions"] == rounds + 1
        if rebalance:
            moved = torch.tensor([float(dom.stats.get("cut_moves", 0))]); dist.all_reduce(moved)
            assert moved.item() > 0, "the cuts never moved"
            own_x = be.T["pos"][:dom.n_own, 0]
            lo_ok = rank == 0 or float(own_x.min()) >= dom.xlo          # the end slabs keep whatever lies beyond the domain
            hi_ok = rank == world - 1 or float(own_x.max()) < dom.xhi
            assert dom.n_own == 0 or (lo_ok and hi_ok), "an owned particle lies outside the slab"
        tot = torch.tensor([float(dom.stats["migrants_out"])]); dist.all_reduce(tot)
        assert tot.item() > 0, "the test moved nobody across a cut"
        m = dom.all_max(torch.tensor([float(rank)])); assert m.item() == world - 1
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_slab_partition_over_gloo(world):
    port = 29500 + (os.getpid() % 2000) + world
    mp.spawn(_worker, args=(world, port, 6000, 7 + world, 4), nprocs=world, join=True)


def test_slab_rebalancing_over_gloo():
    """The left half of the particles streams to the right: the cuts move (one bucket column per re-partition) and every
    invariant of the partition keeps holding."""
    port = 29500 + (os.getpid() % 2000) + 11
    mp.spawn(_worker, args=(3, port, 6000, 5, 8, True), nprocs=3, join=True)


def test_choose_cuts_balances_and_respects_width():
    cnt = [0, 0, 100, 100, 100, 100, 100, 100, 100, 100, 0, 0, 0, 0, 0, 0]
    assert choose_cuts(cnt, 2) == [0, 6, 16]
    c4 = choose_cuts(cnt, 4)
    assert c4[0] == 0 and c4[-1] == 16 and all(b - a >= 3 for a, b in zip(c4, c4[1:]))
    with pytest.raises(ValueError):
        choose_cuts([1] * 5, 2)
    # all the fluid in the first columns (dam break at rest): the right slabs still get their minimum width
    c = choose_cuts([50] * 4 + [0] * 20, 4)
    assert all(b - a >= 3 for a, b in zip(c, c[1:]))
