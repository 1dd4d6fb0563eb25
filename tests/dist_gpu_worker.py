"""torchrun worker: one process per GPU steps its i-slab of a lattice (halo exchange over CUDA-IPC peer memory);
rank 0 also steps the whole lattice on one engine and checks the gathered result -- bit-exact under TB_MATH_STRICT,
<= 1e-12 under TB_MATH_FAST.  Used by tests/test_partitioned.py (needs >= 2 GPUs) and scripts on the GPU box."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

import titan_b200 as tb
from titan_b200 import distributed, lattice, partition

PLANE = ((0.0, 0.0, 1.0), 0.0, 0.5, 0.6)


def main():
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    n = (int(os.environ.get("TB_TEST_NX", "48")), 24, 24)
    steps = int(os.environ.get("TB_TEST_STEPS", "230"))
    c, d = (0.0, 0.0, 0.1 * (n[2] - 1) / 2 - 0.001), (0.1 * (n[0] - 1), 0.1 * (n[1] - 1), 0.1 * (n[2] - 1))
    ok = True
    os.environ["TB_TPM"] = "1"   # fast arithmetic: same per-mass summation split as the partitioned engines (which use 1)
    for flags, tol in ((tb.MATH_STRICT, 0.0), (tb.MATH_FAST, 1e-12)):
        part = partition.LatticeSlab(c, d, n[0], n[1], n[2], rank, world)
        eng = tb.Engine(local, flags)
        distributed.upload_slab(eng, part, planes=[PLANE])
        distributed.connect(eng, part)
        t = 0.0
        for chunk in (1, 29, steps - 30):
            eng.run_async(chunk, t)
            eng.sync()
            for _ in range(chunk):
                t += 1e-4
        dist.barrier()
        st = eng.get_mass_state(idx=part.own_local, fields=("pos", "vel"))
        eng.close()
        gathered = [None] * world
        dist.all_gather_object(gathered, (part.mass_gid[part.own_local], st["pos"], st["vel"]))
        if rank == 0:
            L = lattice.build_lattice(c, d, *n)
            ref = tb.Engine(local, flags)
            ref.set_topology(L["pos"].shape[0], L["left"], L["right"])
            ref.set_tile_hint(lattice.brick_hint(*n))
            ref.set_mass_state(pos=L["pos"])
            ref.set_spring_params(rest=L["rest"])
            ref.set_planes([tb.Plane((tb.capi.C.c_double * 3)(*PLANE[0]), *PLANE[1:])])
            ref.set_globals((0.0, 0.0, -9.81), 1e-4)
            ref.run(steps, 0.0)
            want = ref.get_mass_state(fields=("pos", "vel"))
            ref.close()
            err = 0.0
            for gid, pos, vel in gathered:
                err = max(err, float(np.max(np.abs(pos - want["pos"][gid]))), float(np.max(np.abs(vel - want["vel"][gid]))))
            good = err <= tol
            ok = ok and good
            print("world %d flags %d: max abs diff vs single engine %.3e (tol %.0e) %s" % (world, flags, err, tol, "OK" if good else "MISMATCH"), flush=True)
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, 0)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) else 1)


if __name__ == "__main__":
    main()
