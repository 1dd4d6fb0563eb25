"""Topology edits while paused without re-uploading the system: tb_add_masses / tb_add_springs rebuild only
the tiles the new springs touch, tb_set_mass_state(valid = 0) freezes a mass and turns its springs into tombstones.
Reference: Simulation::createMass / createSpring after start (src/sim.cu:231-253,300-322), deleteMass (:343-380).

The edited engine must equal, bit for bit, an engine that is uploaded FRESH with the final topology and the same state -- the
per-mass summation order is by spring index whatever the tiling -- in strict and in fast arithmetic (one thread per mass).
"""
import time

import numpy as np
import pytest

import scenes
import titan_b200 as tb

pytestmark = pytest.mark.gpu


def same_bits(a, b):
    return bool((np.ascontiguousarray(a + 0.0).view(np.uint64) == np.ascontiguousarray(b + 0.0).view(np.uint64)).all())


def fresh_engine(flags, sc, pos, vel, left, right, k, rest, damping, nm):
    eng = tb.Engine(0, flags)
    eng.set_topology(nm, left, right)
    m = np.full(nm, 0.1)
    m[:sc.nm] = sc.m
    eng.set_mass_state(pos=pos, vel=vel, m=m)
    eng.set_spring_params(k=k, rest=rest, damping=damping)
    eng.set_planes([tb.Plane((tb.capi.C.c_double * 3)(*p.normal), p.offset, p.friction_k, p.friction_s) for p in sc.planes])
    eng.set_globals(sc.g, sc.dt)
    return eng


@pytest.mark.parametrize("flags", [tb.MATH_STRICT, tb.MATH_FAST], ids=["strict", "fast"])
@pytest.mark.parametrize("scene", ["contact_friction", "lattice24"])
def test_added_springs_and_masses_equal_a_fresh_upload(flags, scene, monkeypatch):
    monkeypatch.setenv("TB_TPM", "1")
    monkeypatch.setenv("TB_RESIDENT", "0")
    sc = scenes.contact_friction() if scene == "contact_friction" else scenes.c2_lattice(24)
    rng = np.random.RandomState(7)
    A = tb.Engine(0, flags)
    scenes.upload(A, sc)
    A.run(30, 0.0)
    t = 0.0
    for _ in range(30):
        t += sc.dt
    info0 = A.info()
    # (1) springs between existing masses (some far apart: new halo entries, higher degree)
    a = rng.randint(0, sc.nm, 60).astype(np.int32)
    b = ((a + rng.randint(1, 40, 60)) % sc.nm).astype(np.int32)
    first_s = A.add_springs(a, b)
    assert first_s == sc.ns
    k_new, rest_new, d_new = np.full(60, 800.0), 0.05 + 0.01 * rng.rand(60), np.where(np.arange(60) % 2 == 0, 0.02, 0.0)
    A.set_spring_params(first=first_s, count=60, k=k_new, rest=rest_new, damping=d_new)
    # (2) new masses hanging on the body and on each other (more than one tile of them)
    n_new = 150
    first_m = A.add_masses(n_new)
    assert first_m == sc.nm
    st = A.get_mass_state(first=0, count=sc.nm, fields=("pos", "vel"))
    anchor = rng.randint(0, sc.nm, n_new)
    pos_new = st["pos"][anchor] + np.array([0.0, 0.0, 0.05]) + 0.01 * rng.rand(n_new, 3)
    A.set_mass_state(first=first_m, count=n_new, pos=pos_new, vel=np.zeros((n_new, 3)), m=np.full(n_new, 0.1))
    l2 = np.concatenate([anchor, first_m + np.arange(n_new - 1)]).astype(np.int32)
    r2 = np.concatenate([first_m + np.arange(n_new), first_m + 1 + np.arange(n_new - 1)]).astype(np.int32)
    first_s2 = A.add_springs(l2, r2)
    k2, rest2 = np.full(l2.size, 300.0), np.full(l2.size, 0.06)
    A.set_spring_params(first=first_s2, count=l2.size, k=k2, rest=rest2, damping=np.zeros(l2.size))
    A.run(50, t)
    info1 = A.info()
    assert info1["kernel"] == 1 and info1["nm"] == sc.nm + n_new and info1["ns"] == sc.ns + 60 + l2.size
    assert info1["n_tiles"] == info0["n_tiles"] + 2          # the new masses live in two tiles of their own
    got = A.get_mass_state(fields=("pos", "vel"))
    A.close()
    # the same system uploaded fresh at step 30
    left = np.concatenate([sc.left, a, l2]).astype(np.int32)
    right = np.concatenate([sc.right, b, r2]).astype(np.int32)
    B = fresh_engine(flags, sc, np.vstack([st["pos"], pos_new]), np.vstack([st["vel"], np.zeros((n_new, 3))]), left, right,
                     np.concatenate([sc.k, k_new, k2]), np.concatenate([sc.rest, rest_new, rest2]),
                     np.concatenate([sc.damping, d_new, np.zeros(l2.size)]), sc.nm + n_new)
    B.run(50, t)
    want = B.get_mass_state(fields=("pos", "vel"))
    B.close()
    assert same_bits(got["pos"], want["pos"]) and same_bits(got["vel"], want["vel"])


@pytest.mark.parametrize("flags", [tb.MATH_STRICT, tb.MATH_FAST], ids=["strict", "fast"])
def test_deleted_masses_freeze_and_their_springs_vanish(flags, monkeypatch):
    monkeypatch.setenv("TB_TPM", "1")
    sc = scenes.contact_friction()
    o = sc.copy()
    eng = tb.Engine(0, flags | tb.NO_RESIDENT)
    scenes.upload(eng, sc)
    eng.run(25, 0.0)
    o.step(25)
    builds = eng.info()["launches"]
    gone = np.array([3, 77, 140, 141], dtype=np.int32)
    eng.set_mass_state(idx=gone, valid=np.zeros(gone.size, dtype=np.uint8))
    o.valid[gone] = 0
    eng.run(60, o.T_sim)
    o.step(60)
    st = eng.get_mass_state()
    assert eng.info()["ns_removed"] > 0 and eng.info()["launches"] > builds
    eng.close()
    tol = 0.0 if flags & tb.MATH_STRICT else 1e-9
    for f in ("pos", "vel"):
        err = float(np.max(np.abs(st[f] - getattr(o, f))))
        assert err <= tol, (f, err)
    assert same_bits(st["pos"][gone], o.pos[gone])           # frozen where they were


@pytest.mark.parametrize("n_edits", [5, 24], ids=["5 edits", "24 edits"])
@pytest.mark.parametrize("flags", [tb.MATH_STRICT, tb.MATH_FAST], ids=["strict", "fast"])
def test_a_series_of_edits_goes_into_the_room_left_in_the_tables(flags, n_edits, monkeypatch):
    """Five edits in a row, the layout brought up to date after each (tb_run_async of zero steps): the first one re-allocates
    the per-spring / per-mass arrays with room, the later ones fill that room -- arrays and tile tables alike -- and nothing is
    re-uploaded.  The result equals a fresh upload of the final system bit for bit.  24 edits: the lists left behind by rebuilt
    tiles pile up until the engine builds the layout afresh (or a table outgrows its room and moves) -- same result."""
    monkeypatch.setenv("TB_TPM", "1")
    monkeypatch.setenv("TB_RESIDENT", "0")
    sc = scenes.c2_lattice(24)
    rng = np.random.RandomState(11)
    A = tb.Engine(0, flags)
    scenes.upload(A, sc)
    A.run(30, 0.0)
    t = 0.0
    for _ in range(30):
        t += sc.dt
    st = A.get_mass_state(first=0, count=sc.nm, fields=("pos", "vel"))
    left, right, k, rest, damping = [sc.left], [sc.right], [sc.k], [sc.rest], [sc.damping]
    pos, vel = [st["pos"]], [st["vel"]]
    nm, ns = sc.nm, sc.ns
    uploads = A.info()["layout_uploads"]
    for edit in range(n_edits):
        n_s = 25
        a = rng.randint(0, 400, n_s).astype(np.int32)              # local edits: a dozen tiles each
        b = ((a + rng.randint(1, 60, n_s)) % sc.nm).astype(np.int32)
        assert A.add_springs(a, b) == ns
        kk, rr = np.full(n_s, 500.0 + 10 * edit), 0.05 + 0.01 * rng.rand(n_s)
        A.set_spring_params(first=ns, count=n_s, k=kk, rest=rr, damping=np.zeros(n_s))
        ns += n_s
        n_m = 40
        assert A.add_masses(n_m) == nm
        anchor = rng.randint(0, 400, n_m)
        pn = st["pos"][anchor] + np.array([0.0, 0.0, 0.04]) + 0.01 * rng.rand(n_m, 3)
        A.set_mass_state(first=nm, count=n_m, pos=pn, vel=np.zeros((n_m, 3)), m=np.full(n_m, 0.1))
        l2, r2 = anchor.astype(np.int32), (nm + np.arange(n_m)).astype(np.int32)
        assert A.add_springs(l2, r2) == ns
        A.set_spring_params(first=ns, count=n_m, k=np.full(n_m, 300.0), rest=np.full(n_m, 0.06), damping=np.zeros(n_m))
        ns += n_m; nm += n_m
        left += [a, l2]; right += [b, r2]; k += [kk, np.full(n_m, 300.0)]; rest += [rr, np.full(n_m, 0.06)]
        damping += [np.zeros(n_s), np.zeros(n_m)]; pos.append(pn); vel.append(np.zeros((n_m, 3)))
        A.run_async(0, t); A.sync()
    if n_edits == 5:
        assert A.info()["layout_uploads"] == uploads and A.info()["layout_patches"] == 5     # the tile tables were never re-uploaded
    else:
        assert A.info()["layout_uploads"] + A.info()["layout_patches"] >= uploads + 2
    A.run(50, t)
    got = A.get_mass_state(fields=("pos", "vel"))
    assert A.info()["kernel"] == 1 and A.info()["nm"] == nm and A.info()["ns"] == ns
    A.close()
    B = fresh_engine(flags, sc, np.vstack(pos), np.vstack(vel), np.concatenate(left).astype(np.int32), np.concatenate(right).astype(np.int32),
                     np.concatenate(k), np.concatenate(rest), np.concatenate(damping), nm)
    B.run(50, t)
    want = B.get_mass_state(fields=("pos", "vel"))
    B.close()
    assert same_bits(got["pos"], want["pos"]) and same_bits(got["vel"], want["vel"])


def test_edit_cost_on_a_large_lattice():
    """60^3 lattice (216,000 masses, 2.7 M springs): 1,000 springs + 100 masses added in one region, then the layout brought up
    to date: only the tiles holding an endpoint are rebuilt (the full build runs on all host cores and is the yardstick)."""
    import bench
    W = bench.workload_c2(60)
    eng = tb.Engine(0, tb.MATH_FAST)
    eng.set_topology(W["pos"].shape[0], W["left"], W["right"])
    eng.set_tile_hint(W["tile_hint"])
    eng.set_mass_state(pos=W["pos"])
    eng.set_spring_params(rest=W["rest"])
    eng.set_planes([tb.Plane((tb.capi.C.c_double * 3)(0, 0, 1), 0.0, 0.5, 0.6)])
    eng.set_globals((0, 0, -9.81), 1e-4)
    t0 = time.perf_counter()
    eng.run(10, 0.0)
    full = time.perf_counter() - t0
    nm = W["pos"].shape[0]
    rng = np.random.RandomState(3)
    a = rng.randint(0, 20000, 1000).astype(np.int32)       # a local edit: the first few i-planes of the lattice
    t1 = time.perf_counter()
    s0 = eng.add_springs(a, a + 1 + rng.randint(0, 100, 1000).astype(np.int32))
    eng.set_spring_params(first=s0, count=1000, k=np.full(1000, 50.0), rest=np.full(1000, 0.1))
    m0 = eng.add_masses(100)
    eng.set_mass_state(first=m0, count=100, pos=W["pos"][:100] + (0, 0, 6.5))
    eng.add_springs(np.arange(m0, m0 + 99, dtype=np.int32), np.arange(m0 + 1, m0 + 100, dtype=np.int32))
    eng.run_async(0, 1e-3)           # brings the layout up to date
    eng.sync()
    edit = time.perf_counter() - t1
    eng.run(10, 1e-3)
    info = eng.info()
    st = eng.get_mass_state(first=0, count=10, fields=("pos",))
    eng.close()
    print("first build + 10 steps %.3f s, edit (1,099 springs, 100 masses) %.4f s" % (full, edit))
    assert info["kernel"] == 1 and np.isfinite(st["pos"]).all()
    assert edit < 2.0          # wall clock, informational: printed above; scripts/edit_cost_c2.py measures C2 (0.12 s vs 1 s for a build)


def test_removed_springs_leave_no_dead_words_after_compaction(monkeypatch):
    """A small removal is compacted on the spot (the tiles holding the springs are rebuilt without them); a large one stays in
    place with k = 0 until tb_compact.  Either way the results equal the oracle without those springs."""
    import oracle
    monkeypatch.setenv("TB_TPM", "1")
    sc = scenes.c2_lattice(24)
    eng = tb.Engine(0, tb.MATH_FAST)
    scenes.upload(eng, sc)
    eng.run(20, 0.0)
    bytes0 = eng.info()["tiled_bytes_per_step"]
    o = sc.copy()
    o.step(20)
    small = np.arange(100, 160, dtype=np.int32)                       # <= 4096: patched tile by tile
    large = np.arange(1000, sc.ns, 17, dtype=np.int32)                # > 4096: tombstones in place
    assert large.size > 4096
    eng.remove_springs(small)
    eng.run(20, o.T_sim)
    eng.remove_springs(large)
    eng.run(20, o.T_sim + 20 * sc.dt)
    eng.compact()
    eng.run(20, o.T_sim + 40 * sc.dt)
    got = eng.get_mass_state(fields=("pos", "vel"))
    info = eng.info()
    eng.close()
    assert info["ns_removed"] == small.size + large.size and info["kernel"] == 1
    # the oracle: same schedule with the springs physically absent
    def without(o_prev, gone_so_far):
        keep = np.setdiff1d(np.arange(sc.ns), gone_so_far)
        o2 = oracle.Scene(o_prev.pos.copy(), sc.left[keep], sc.right[keep], sc.rest[keep], k=1.0, m=0.1, dt=sc.dt, g=sc.g)
        o2.k[:] = sc.k[keep]; o2.vel[:] = o_prev.vel; o2.planes.extend(sc.planes); o2.T_sim = o_prev.T_sim
        return o2
    o1 = without(o, small); o1.step(20)
    o2 = without(o1, np.concatenate([small, large])); o2.step(40)
    for f in ("pos", "vel"):
        err = float(np.max(np.abs(got[f] - getattr(o2, f))))
        assert err <= 1e-9, (f, err)
