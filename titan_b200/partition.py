"""Partition one large mass-spring system over several engines (BASELINE config C4).

Every engine holds its OWN masses plus GHOST copies of the foreign endpoints of its cut springs, and every spring
with at least one own endpoint (cut springs are stored by both sides, in the global spring order, so per-mass
summation order -- and with TB_MATH_STRICT the bits -- equal the single-engine run).  Per step exactly one thing
crosses the cut: the new state of the boundary masses (the step is Jacobi-style, reference src/sim.cu:1801-1809).

Two partitioners, one interface (`local`, `ghost_owner`, `local_of_global`):
  GraphPartition  any graph, any owner map (numpy over the global arrays)
  LatticeSlab     a reference-ordered lattice cut into i-slabs without ever materialising the global arrays
                  (300^3 = 27 M masses / 349 M springs would cost ~40 GB per rank)
`wire()` turns what the ranks publish about their ghosts into the tb_halo_connect / tb_halo_set_exports arguments.
Host-side plumbing only: numpy + any allgather (torch.distributed in bench.py, plain lists in the tests).
"""
import numpy as np

from . import lattice


class GraphPartition:
    """Generic partition by an owner map.  Local mass order: ascending global id (own and ghosts interleaved)."""

    def __init__(self, nm, left, right, owner, rank):
        left = np.asarray(left, dtype=np.int64)
        right = np.asarray(right, dtype=np.int64)
        owner = np.asarray(owner, dtype=np.int32)
        self.rank, self.nm_global, self._owner = int(rank), int(nm), owner
        mine_l, mine_r = owner[left] == rank, owner[right] == rank
        keep = mine_l | mine_r
        self.spring_gid = np.nonzero(keep)[0]                       # ascending: global spring order preserved
        touched = np.zeros(nm, dtype=bool)
        touched[left[keep]] = True
        touched[right[keep]] = True
        touched[owner == rank] = True                               # own masses without springs stay
        self.mass_gid = np.nonzero(touched)[0]
        g2l = np.full(nm, -1, dtype=np.int64)
        g2l[self.mass_gid] = np.arange(self.mass_gid.size)
        self._g2l = g2l
        self.left = g2l[left[keep]].astype(np.int32)
        self.right = g2l[right[keep]].astype(np.int32)
        self.is_ghost = owner[self.mass_gid] != rank
        self.ghost_local = np.nonzero(self.is_ghost)[0].astype(np.int32)
        self.ghost_gid = self.mass_gid[self.ghost_local]
        self.ghost_owner = owner[self.ghost_gid]
        self.own_local = np.nonzero(~self.is_ghost)[0].astype(np.int32)
        self.ns_owned = int(mine_l.sum())                           # global totals count a spring with its left endpoint's owner

    def local_of_global(self, gid):
        return self._g2l[np.asarray(gid, dtype=np.int64)]

    def owner_of(self, gid):
        return self._owner[np.asarray(gid, dtype=np.int64)]


def slab_bounds(nx, world):
    """i-plane ranges [lo, hi) per rank: contiguous, sizes differ by at most one plane."""
    base, extra = divmod(nx, world)
    lo = [r * base + min(r, extra) for r in range(world)]
    return [(lo[r], lo[r] + base + (1 if r < extra else 0)) for r in range(world)]


class LatticeSlab:
    """Rank `rank`'s share of an nx x ny x nz reference lattice cut into `world` i-slabs: its own planes plus one
    ghost plane per neighbouring slab.  Built directly as a small lattice of (own + ghost) planes."""

    def __init__(self, center, dims, nx, ny, nz, rank, world):
        if nx < world:
            raise ValueError("fewer i-planes than ranks")
        self.rank, self.world = int(rank), int(world)
        self.n = (nx, ny, nz)
        self.bounds = slab_bounds(nx, world)
        lo, hi = self.bounds[rank]
        self.g0 = lo - (1 if rank > 0 else 0)
        self.g1 = hi + (1 if rank < world - 1 else 0)
        npl = ny * nz
        self.plane = npl
        nloc = self.g1 - self.g0
        self.pos = lattice.lattice_positions(center, dims, nx, ny, nz, i_range=(self.g0, self.g1))
        left, right = lattice.lattice_springs(nloc, ny, nz)
        # drop springs with both endpoints in a ghost plane (in-plane springs of the ghost planes)
        own_lo, own_hi = (lo - self.g0) * npl, (hi - self.g0) * npl
        own_l = (left >= own_lo) & (left < own_hi)
        own_r = (right >= own_lo) & (right < own_hi)
        keep = own_l | own_r
        self.left, self.right = left[keep], right[keep]
        self.rest = lattice.rest_lengths(self.pos, self.left, self.right)
        self.nm_local = nloc * npl
        self.mass_gid = np.arange(self.g0 * npl, self.g1 * npl, dtype=np.int64)
        loc = np.arange(self.nm_local)
        self.is_ghost = (loc < own_lo) | (loc >= own_hi)
        self.ghost_local = np.nonzero(self.is_ghost)[0].astype(np.int32)
        self.ghost_gid = self.mass_gid[self.ghost_local]
        self.ghost_owner = np.where(self.ghost_local < own_lo, rank - 1, rank + 1).astype(np.int32)
        self.own_local = np.nonzero(~self.is_ghost)[0].astype(np.int32)
        # springs this rank accounts for in global totals: those whose LEFT endpoint it owns (left has the lower i)
        self.ns_owned = int(own_l[keep].sum())
        self.nm_own = (hi - lo) * npl

    def local_of_global(self, gid):
        gid = np.asarray(gid, dtype=np.int64)
        out = gid - self.g0 * self.plane
        out[(gid < self.g0 * self.plane) | (gid >= self.g1 * self.plane)] = -1
        return out

    def owner_of(self, gid):
        plane = np.asarray(gid, dtype=np.int64) // self.plane
        starts = np.array([b[0] for b in self.bounds])
        return (np.searchsorted(starts, plane, side="right") - 1).astype(np.int32)

    def tile_hint(self, tile=128):
        """Brick ids for tb_set_tile_hint over the local (own + ghost) planes; the engine itself keeps ghosts in
        tiles of their own and moves boundary tiles to the front."""
        lo, hi = self.bounds[self.rank]
        nx, ny, nz = self.n
        hint = np.empty(self.nm_local, dtype=np.int32)
        own = lattice.brick_hint(hi - lo, ny, nz, tile=tile, base=0)
        base = int(own.max()) + 1
        a = (lo - self.g0) * self.plane
        hint[a:a + own.size] = own
        for sl in ((0, a), (a + own.size, self.nm_local)):
            if sl[1] > sl[0]:
                g = lattice.brick_hint(1, ny, nz, tile=tile, base=base)
                hint[sl[0]:sl[1]] = g
                base = int(g.max()) + 1
        return hint


def publish(part, handle, ghost_offsets):
    """What one rank tells the others (an allgather of these dicts is all the plumbing the exchange needs)."""
    return {"rank": part.rank, "handle": handle, "ghost_gid": np.asarray(part.ghost_gid, dtype=np.int64),
            "ghost_owner": np.asarray(part.ghost_owner, dtype=np.int32),
            "ghost_offset": np.asarray(ghost_offsets, dtype=np.int64)}


def peers_of(rank, infos):
    """Ranks `rank` exchanges boundary state with: owners of its ghosts and ranks holding ghosts it owns."""
    me = infos[rank]
    p = set(int(x) for x in np.unique(me["ghost_owner"]))
    for q, inf in enumerate(infos):
        if q != rank and (inf["ghost_owner"] == rank).any():
            p.add(q)
    p.discard(rank)
    return sorted(p)


def wire(part, infos):
    """Arguments for tb_halo_connect and tb_halo_set_exports of rank part.rank from the published infos."""
    r = part.rank
    peers = peers_of(r, infos)
    handles = [infos[p]["handle"] for p in peers]
    flag_slot = [peers_of(p, infos).index(r) for p in peers]
    e_idx, e_peer, e_off = [], [], []
    for j, p in enumerate(peers):
        sel = infos[p]["ghost_owner"] == r
        loc = part.local_of_global(infos[p]["ghost_gid"][sel])
        assert (loc >= 0).all(), "a peer holds a ghost of a mass this rank does not have"
        e_idx.append(loc.astype(np.int32))
        e_peer.append(np.full(loc.size, j, dtype=np.int32))
        e_off.append(infos[p]["ghost_offset"][sel])
    cat = lambda xs, dt: np.concatenate(xs).astype(dt) if xs else np.zeros(0, dtype=dt)  # noqa: E731
    return {"peers": peers, "handles": handles, "flag_slot": flag_slot, "export_idx": cat(e_idx, np.int32),
            "export_peer": cat(e_peer, np.int32), "export_offset": cat(e_off, np.int64)}


def connect(engines, parts, allgather=None):
    """Handshake.  In-process form: `engines` and `parts` are lists over all ranks.  Distributed form: pass this
    rank's engine and partition (not lists) plus `allgather(obj) -> list of every rank's obj`."""
    if allgather is None:
        handles = [en.halo_prepare() for en in engines]
        infos = [publish(pt, h, en.halo_state_offsets(pt.ghost_local)) for en, pt, h in zip(engines, parts, handles)]
        wires = [wire(pt, infos) for pt in parts]
        for en, w in zip(engines, wires):
            en.halo_connect(w["handles"], w["flag_slot"])
            en.halo_set_exports(w["export_idx"], w["export_peer"], w["export_offset"])
        return wires
    handle = engines.halo_prepare()
    infos = allgather(publish(parts, handle, engines.halo_state_offsets(parts.ghost_local)))
    w = wire(parts, infos)
    engines.halo_connect(w["handles"], w["flag_slot"])
    engines.halo_set_exports(w["export_idx"], w["export_peer"], w["export_offset"])
    return w
