"""One process per GPU: the torch.distributed plumbing around a partitioned engine (bench.py, multi-GPU tests).

torch is used for rendezvous, barriers and the allgather of the halo handshake only; the per-step boundary
exchange itself happens inside the step kernel over NVLink peer memory (tb_halo_*), never through NCCL or the host.
"""
import numpy as np

from . import capi, partition


def allgather_objects(obj):
    import torch.distributed as dist
    out = [None] * dist.get_world_size()
    dist.all_gather_object(out, obj)
    return out


def upload_slab(eng, part, planes=(), g=(0.0, 0.0, -9.81), dt=1e-4, k=None, m=None, hint=True):
    """Upload rank part.rank's piece of a lattice (partition.LatticeSlab) and mark its ghost planes."""
    eng.set_topology(part.nm_local, part.left, part.right)
    if hint:
        eng.set_tile_hint(part.tile_hint())
    eng.set_mass_state(pos=part.pos, m=None if m is None else np.full(part.nm_local, float(m)))
    eng.set_spring_params(rest=part.rest, k=None if k is None else np.full(part.left.size, float(k)))
    eng.set_planes([capi.Plane((capi.C.c_double * 3)(*n), off, fk, fs) for (n, off, fk, fs) in planes])
    eng.set_globals(g, dt)
    eng.halo_set_ghosts(part.ghost_local)


def connect(eng, part):
    """Collective: every rank of the default process group calls this once after its uploads."""
    import torch.distributed as dist
    w = partition.connect(eng, part, allgather=allgather_objects)
    dist.barrier()
    return w
