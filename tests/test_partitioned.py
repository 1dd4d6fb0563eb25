"""Partitioned simulation: one system split over several engines with boundary-mass halo exchange fused into the
step kernel (include/titan_b200.h tb_halo_*; titan_b200/partition.py).

CPU: the partitioners (generic owner map vs. lattice i-slabs) agree with each other and with the global lattice
bit for bit, and a 2-process gloo run of the ORACLE over the two halves of a lattice -- ghosts exchanged with
send/recv after every step -- equals the single-domain oracle bit for bit (the decomposition itself is exact).
GPU (single box, all engines on cuda:0 with small persistent grids so they are co-resident): the partitioned
engines equal the single engine -- bit-exact under TB_MATH_STRICT, <= 1e-12 under TB_MATH_FAST.
"""
import os
import subprocess
import sys

import numpy as np
import pytest

import oracle
import scenes
import titan_b200 as tb
from titan_b200 import lattice, partition

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def owner_slabs(nx, ny, nz, world):
    owner = np.zeros(nx * ny * nz, dtype=np.int32)
    for r, (lo, hi) in enumerate(partition.slab_bounds(nx, world)):
        owner[lo * ny * nz:hi * ny * nz] = r
    return owner


@pytest.mark.parametrize("world", [2, 3, 4])
def test_slab_partition_equals_generic_partition(world):
    nx, ny, nz = 9, 4, 5
    c, d = (0.0, 0.0, 1.0), (2.0, 1.0, 1.5)
    L = lattice.build_lattice(c, d, nx, ny, nz)
    owner = owner_slabs(nx, ny, nz, world)
    total, seen = 0, np.zeros(nx * ny * nz, dtype=int)
    for r in range(world):
        g = partition.GraphPartition(nx * ny * nz, L["left"], L["right"], owner, r)
        s = partition.LatticeSlab(c, d, nx, ny, nz, r, world)
        assert (g.mass_gid == s.mass_gid).all() and (g.left == s.left).all() and (g.right == s.right).all()
        assert (g.is_ghost == s.is_ghost).all() and (g.ghost_owner == s.ghost_owner).all()
        assert s.pos.tobytes() == L["pos"][s.mass_gid].tobytes()              # same doubles as the global lattice
        assert s.rest.tobytes() == L["rest"][g.spring_gid].tobytes()
        assert (s.owner_of(s.mass_gid) == owner[s.mass_gid]).all()
        assert g.ns_owned == s.ns_owned
        seen[s.mass_gid[~s.is_ghost]] += 1
        total += s.ns_owned
        assert s.tile_hint().shape == (s.nm_local,)
    assert (seen == 1).all() and total == L["left"].size                     # every mass / spring owned exactly once


def test_wire_is_symmetric():
    """What rank a exports to rank b is exactly b's ghost list, entry for entry."""
    nx, ny, nz, world = 8, 3, 3, 3
    parts = [partition.LatticeSlab((0, 0, 0), (1, 1, 1), nx, ny, nz, r, world) for r in range(world)]
    infos = [partition.publish(p, b"h%d" % p.rank, 1000 * p.rank + np.arange(p.ghost_local.size)) for p in parts]
    for p in parts:
        w = partition.wire(p, infos)
        assert w["peers"] == [q for q in (p.rank - 1, p.rank + 1) if 0 <= q < world]
        for j, q in enumerate(w["peers"]):
            sel = w["export_peer"] == j
            exported_gid = p.mass_gid[w["export_idx"][sel]]
            want = infos[q]["ghost_owner"] == p.rank
            assert (exported_gid == infos[q]["ghost_gid"][want]).all()
            assert (w["export_offset"][sel] == infos[q]["ghost_offset"][want]).all()
            assert partition.peers_of(q, infos)[w["flag_slot"][j]] == p.rank


def test_two_process_oracle_partition_is_exact(tmp_path):
    """world_size 2 over gloo: the decomposition (ghosts + one position/velocity exchange per step) is exact."""
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.path.join(ROOT, "tests"), OMP_NUM_THREADS="1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
                        "127.0.0.1", "--master-port", "29611", os.path.join(ROOT, "tests", "dist_oracle_worker.py")],
                       env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:]
    assert r.stdout.count("partition exact") == 2, r.stdout[-3000:]


# ------------------------------------------------------------------------------------------------ GPU ----------
def make_scene(kind):
    nx, ny, nz = 12, 6, 6
    sc = scenes.lattice_scene((0, 0, 0.29), (1.1, 0.5, 0.5), (nx, ny, nz))
    n, s = np.arange(sc.nm), np.arange(sc.ns)
    sc.vel[:, 0] = 0.2 + 0.001 * (n % 7)
    sc.vel[:, 2] = -0.3 + 0.002 * (n % 5)
    sc.planes.append(oracle.make_plane((0, 0, 1), 0, 0.5, 0.6))
    if kind == "damped_active":
        sc.damping[:] = np.where(s % 3 == 0, 0.05, 0.0)
        sc.type[:] = np.where(s % 4 == 1, 2, np.where(s % 4 == 2, 3, 0))
        sc.omega[:] = 4 * np.pi
    if kind == "actuated":      # RobotLink-style springs (device-side _rest mutation), many of them across the slab cuts
        act = s % 11 == 3
        sc.type[act] = 4 + (s[act] % 2)
        sc.l_max[act] = sc.rest[act] * 1.004
        sc.l_min[act] = sc.rest[act] * 0.996
        sc.rate[act] = 0.05 + 0.01 * (s[act] % 5)
    return sc, (nx, ny, nz)


def upload_part(eng, sc, part, sel_m, sel_s):
    """Upload the local piece of a global oracle scene (sel_m / sel_s: global ids of local masses / springs)."""
    eng.set_topology(sel_m.size, part.left, part.right)
    eng.set_mass_state(pos=sc.pos[sel_m], vel=sc.vel[sel_m], m=sc.m[sel_m])
    eng.set_spring_params(k=sc.k[sel_s], rest=sc.rest[sel_s], damping=sc.damping[sel_s], type=sc.type[sel_s], omega=sc.omega[sel_s],
                          l_max=sc.l_max[sel_s], l_min=sc.l_min[sel_s], rate=sc.rate[sel_s])
    eng.set_planes([tb.Plane((tb.capi.C.c_double * 3)(*p.normal), p.offset, p.friction_k, p.friction_s) for p in sc.planes])
    eng.set_globals(sc.g, sc.dt)
    eng.halo_set_ghosts(part.ghost_local)


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["plain", "damped_active", "actuated"])
@pytest.mark.parametrize("world,flags,tol", [(2, tb.MATH_STRICT, 0.0), (3, tb.MATH_STRICT, 0.0), (3, tb.MATH_FAST, 1e-12)])
def test_partitioned_engines_match_single_engine(kind, world, flags, tol, monkeypatch):
    monkeypatch.setenv("TB_GRID", "16")     # all engines of the partition co-resident on one GPU
    sc, (nx, ny, nz) = make_scene(kind)
    steps = 120
    ref = tb.Engine(0, flags)
    scenes.upload(ref, sc)
    ref.run(steps, 0.0)
    want = ref.get_mass_state()
    want_rest = ref.get_spring_rest()
    ref.close()
    if kind == "plain" and flags == tb.MATH_STRICT:
        o = sc.copy(); o.step(steps)
        assert o.pos.tobytes() == want["pos"].tobytes()      # and the single engine equals the oracle, bit for bit

    owner = owner_slabs(nx, ny, nz, world)
    parts = [partition.GraphPartition(sc.nm, sc.left, sc.right, owner, r) for r in range(world)]
    engines = [tb.Engine(0, flags) for _ in range(world)]
    for en, pt in zip(engines, parts):
        upload_part(en, sc, pt, pt.mass_gid, pt.spring_gid)
    partition.connect(engines, parts)
    t = 0.0
    for chunk in (1, 7, 62, 50):                              # single launches, a tail, and a whole CUDA graph
        for en in engines:
            en.run_async(chunk, t)
        for en in engines:
            en.sync()
        for _ in range(chunk):
            t += sc.dt
    got_pos, got_vel = np.zeros_like(want["pos"]), np.zeros_like(want["vel"])
    for en, pt in zip(engines, parts):
        st = en.get_mass_state(fields=("pos", "vel"))
        got_pos[pt.mass_gid[pt.own_local]] = st["pos"][pt.own_local]
        got_vel[pt.mass_gid[pt.own_local]] = st["vel"][pt.own_local]
        # every engine advanced the rest lengths of its springs (cut ones on both sides) exactly like the single engine
        assert en.get_spring_rest().tobytes() == want_rest[pt.spring_gid].tobytes()
        # ghosts hold the owner's latest state
        gh = en.get_mass_state(idx=pt.ghost_local, fields=("pos",))["pos"]
        assert np.max(np.abs(gh - want["pos"][pt.ghost_gid])) <= max(tol, 0.0) * 10 + (0 if tol == 0 else 1e-12)
    for en in engines:
        en.close()
    err = max(float(np.max(np.abs(got_pos - want["pos"]))), float(np.max(np.abs(got_vel - want["vel"]))))
    print(kind, world, flags, "max abs diff", err)
    assert err <= tol


@pytest.mark.gpu
def test_parameter_upload_after_connect_rebuilds_records_in_place(monkeypatch):
    """Partitioned engines run the dictionary-record layout (fast arithmetic, one thread per mass).  A stiffness upload
    that splits parameter classes rebuilds the records while the mass slots, state buffers and exchange tables the
    peers hold pointers into stay where they are."""
    monkeypatch.setenv("TB_GRID", "16")
    monkeypatch.setenv("TB_TPM", "1")
    sc, (nx, ny, nz) = make_scene("plain")
    k2 = sc.k.copy()
    k2[::7] *= 1.75
    ref = tb.Engine(0, tb.MATH_FAST)
    scenes.upload(ref, sc)
    ref.run(40, 0.0)
    t40 = 0.0
    for _ in range(40):
        t40 += sc.dt
    ref.set_spring_params(k=k2)
    ref.run(60, t40)
    want = ref.get_mass_state()
    ref.close()

    world = 2
    owner = owner_slabs(nx, ny, nz, world)
    parts = [partition.GraphPartition(sc.nm, sc.left, sc.right, owner, r) for r in range(world)]
    engines = [tb.Engine(0, tb.MATH_FAST) for _ in range(world)]
    for en, pt in zip(engines, parts):
        upload_part(en, sc, pt, pt.mass_gid, pt.spring_gid)
    partition.connect(engines, parts)
    for en in engines:
        en.run_async(40, 0.0)
    for en in engines:
        en.sync()
    assert all(en.info()["dict_records"] > 0 for en in engines)
    for en, pt in zip(engines, parts):
        en.set_spring_params(k=k2[pt.spring_gid])
    # These engines share ONE GPU: the rebuild's cudaMalloc/cudaFree wait for the whole device, so every engine
    # brings its layout up to date (zero-step run) before any of them launches steps that wait for a peer.
    for en in engines:
        en.run_async(0, t40)
    for en in engines:
        en.run_async(60, t40)
    for en in engines:
        en.sync()
    assert all(en.info()["dict_records"] > 0 for en in engines)
    err = 0.0
    for en, pt in zip(engines, parts):
        st = en.get_mass_state(fields=("pos", "vel"))
        for f in ("pos", "vel"):
            err = max(err, float(np.max(np.abs(st[f][pt.own_local] - want[f][pt.mass_gid[pt.own_local]]))))
        en.close()
    assert err <= 1e-12, err


@pytest.mark.gpu
def test_upload_while_paused_then_push(monkeypatch):
    """New state uploaded to the owners reaches the ghosts through tb_halo_push_now."""
    monkeypatch.setenv("TB_GRID", "16")
    sc, (nx, ny, nz) = make_scene("plain")
    world = 2
    parts = [partition.LatticeSlab((0, 0, 0.29), (1.1, 0.5, 0.5), nx, ny, nz, r, world) for r in range(world)]
    owner = owner_slabs(nx, ny, nz, world)
    gparts = [partition.GraphPartition(sc.nm, sc.left, sc.right, owner, r) for r in range(world)]
    engines = [tb.Engine(0, tb.MATH_STRICT) for _ in range(world)]
    for en, pt, gp in zip(engines, parts, gparts):
        upload_part(en, sc, pt, pt.mass_gid, gp.spring_gid)
        en.set_tile_hint(pt.tile_hint())
    partition.connect(engines, parts)
    for en in engines:
        en.run_async(10, 0.0)
    for en in engines:
        en.sync()
    # restart from the initial scene with every mass moved: owners get the new state, ghosts are NOT uploaded
    sc2 = sc.copy()
    sc2.pos[:, 0] += 0.125
    for en, pt in zip(engines, parts):
        en.set_mass_state(idx=pt.own_local, pos=sc2.pos[pt.mass_gid[pt.own_local]], vel=sc2.vel[pt.mass_gid[pt.own_local]])
    for en in engines:
        en.halo_push_now()
    for en in engines:
        en.run_async(25, 0.0)
    for en in engines:
        en.sync()
    o = sc2.copy(); o.step(25)
    for en, pt in zip(engines, parts):
        st = en.get_mass_state(fields=("pos", "vel"))
        own = pt.own_local
        assert st["pos"][own].tobytes() == o.pos[pt.mass_gid[own]].tobytes()
        assert st["vel"][own].tobytes() == o.vel[pt.mass_gid[own]].tobytes()
        en.close()


@pytest.mark.gpu
def test_lone_engine_of_a_partition_reports_instead_of_hanging(monkeypatch):
    monkeypatch.setenv("TB_GRID", "16")
    sc, (nx, ny, nz) = make_scene("plain")
    owner = owner_slabs(nx, ny, nz, 2)
    parts = [partition.GraphPartition(sc.nm, sc.left, sc.right, owner, r) for r in range(2)]
    engines = [tb.Engine(0, tb.MATH_FAST) for _ in range(2)]
    for en, pt in zip(engines, parts):
        upload_part(en, sc, pt, pt.mass_gid, pt.spring_gid)
    partition.connect(engines, parts)
    engines[0].run_async(3, 0.0)                 # its peer never runs: step 2 cannot get the peer's step-1 state
    with pytest.raises(tb.TitanError, match="did not signal"):
        engines[0].sync()
    for en in engines:
        en.close()


@pytest.mark.gpu
def test_one_process_per_gpu_over_cuda_ipc():
    """The real deployment shape: torchrun, one process per GPU, peer memory opened through CUDA IPC."""
    if tb.device_count() < 2:
        pytest.skip("needs at least 2 GPUs")
    world = min(4, tb.device_count())
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.path.join(ROOT, "tests"))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=%d" % world, "--master-addr",
                        "127.0.0.1", "--master-port", "29612", os.path.join(ROOT, "tests", "dist_gpu_worker.py")],
                       env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    print(r.stdout[-2000:])
    assert r.returncode == 0, r.stdout[-3000:]
    assert r.stdout.count(" OK") == 2
