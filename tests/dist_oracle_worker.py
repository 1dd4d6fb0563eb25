"""Worker of tests/test_partitioned.py::test_two_process_oracle_partition_is_exact (launched with torchrun, gloo).
Each rank steps its half of a lattice with the CPU oracle; ghosts are frozen (fixed) masses whose state is replaced
after every step by what the owner sends.  Own masses must equal the single-domain oracle bit for bit."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

import oracle
import scenes
from titan_b200 import partition


def main():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    nx, ny, nz, steps = 10, 4, 4, 40
    c, d = (0.0, 0.0, 0.24), (0.9, 0.3, 0.3)
    full = scenes.lattice_scene(c, d, (nx, ny, nz))
    n = np.arange(full.nm)
    full.vel[:, 0] = 0.2 + 0.001 * (n % 7)
    full.vel[:, 2] = -0.3 + 0.002 * (n % 5)
    full.damping[:] = np.where(np.arange(full.ns) % 3 == 0, 0.05, 0.0)
    full.planes.append(oracle.make_plane((0, 0, 1), 0, 0.5, 0.6))

    part = partition.LatticeSlab(c, d, nx, ny, nz, rank, world)
    owner = part.owner_of(np.arange(full.nm))
    gp = partition.GraphPartition(full.nm, full.left, full.right, owner, rank)
    loc = oracle.Scene(part.pos, part.left, part.right, part.rest)
    loc.vel[:] = full.vel[part.mass_gid]
    loc.damping[:] = full.damping[gp.spring_gid]
    loc.planes.append(oracle.make_plane((0, 0, 1), 0, 0.5, 0.6))
    loc.fixed[part.ghost_local] = 1

    infos = [None] * world
    dist.all_gather_object(infos, partition.publish(part, b"", np.zeros(part.ghost_local.size, dtype=np.int64)))
    w = partition.wire(part, infos)
    for _ in range(steps):
        loc.step(1)
        reqs, recv = [], {}
        for j, p in enumerate(w["peers"]):
            sel = w["export_idx"][w["export_peer"] == j]
            out = torch.from_numpy(np.concatenate([loc.pos[sel], loc.vel[sel]], axis=1).copy())
            reqs.append(dist.isend(out, p))
            mine = part.ghost_local[part.ghost_owner == p]
            recv[p] = (mine, torch.empty((mine.size, 6), dtype=torch.float64))
            reqs.append(dist.irecv(recv[p][1], p))
        for q in reqs:
            q.wait()
        for p, (mine, buf) in recv.items():
            loc.pos[mine] = buf[:, :3].numpy()
            loc.vel[mine] = buf[:, 3:].numpy()
    full.step(steps)
    own = part.own_local
    ok = loc.pos[own].tobytes() == full.pos[part.mass_gid[own]].tobytes() and \
        loc.vel[own].tobytes() == full.vel[part.mass_gid[own]].tobytes()
    print("rank %d: %s" % (rank, "partition exact" if ok else "MISMATCH"), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
