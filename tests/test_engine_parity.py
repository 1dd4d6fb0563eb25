"""GPU parity tests proper: the CUDA engine, driven through the C ABI, against the CPU oracle on the same inputs.

Tolerances (BASELINE.json north_star: <= 1e-9 relative per step in fp64; indexing bit-exact):
  * TB_MATH_STRICT: bit-identical to the oracle for every scenario without sin() actuation; with actuation the
    device sin() may differ from glibc's by 1 ulp, so <= 1e-13 relative after the test's step count.
  * TB_MATH_FAST (what bench.py times): pos/vel <= 1e-12 relative after 1 step, <= 1e-9 after 100 steps.
Relative error is max|a-b| / max(1, max|b|) per field.
"""
import os

import numpy as np
import pytest

import scenes
import titan_b200 as tb

pytestmark = pytest.mark.gpu

FIELDS = ("pos", "vel", "acc")


def rel(a, b):
    return float(np.max(np.abs(a - b)) / max(1.0, float(np.max(np.abs(b)))))


def same_bits(a, b):
    return bool((np.ascontiguousarray(a).view(np.uint64) == np.ascontiguousarray(b).view(np.uint64)).all())


def run_engine(sc, nsteps, flags, chunks=1):
    eng = tb.Engine(0, flags)
    scenes.upload(eng, sc)
    t = sc.T_sim
    per = nsteps // chunks
    done = 0
    for c in range(chunks):
        n = per if c < chunks - 1 else nsteps - done
        eng.run(n, t)
        for _ in range(n):
            t += sc.dt
        done += n
    st = eng.get_mass_state()
    st["rest"] = eng.get_spring_rest() if sc.ns else np.zeros(0)
    st["info"] = eng.info()
    eng.close()
    return st


def oracle_after(sc, nsteps):
    o = sc.copy()
    o.step(nsteps)
    return o


# Small scenes (<= 8 tiles) run on the resident cluster kernel (all steps of a run in one launch, state in shared memory)
# unless NO_RESIDENT asks for the streaming kernel: both are checked.
KERNELS = [0, tb.NO_RESIDENT]


@pytest.mark.parametrize("name", ["c1_readme", "contact_friction", "local_constraints", "cube_beam"])
@pytest.mark.parametrize("nsteps", [1, 10, 137])
@pytest.mark.parametrize("kern", KERNELS)
def test_strict_is_bit_identical(name, nsteps, kern):
    sc = scenes.SCENARIOS[name]()
    o = oracle_after(sc, nsteps)
    g = run_engine(sc, nsteps, tb.MATH_STRICT | kern)
    for f in FIELDS:
        assert rel(g[f], getattr(o, f)) == 0.0, (name, f, rel(g[f], getattr(o, f)))
        assert same_bits(g[f] + 0.0, getattr(o, f) + 0.0), (name, f)  # "+ 0.0" folds -0 into +0
    assert same_bits(g["T"], o.T)
    assert same_bits(g["rest"], o.rest)
    assert same_bits(g["extern_force"] + 0.0, o.ext + 0.0)


@pytest.mark.parametrize("nsteps", [1, 10, 137])
def test_strict_actuated(nsteps):
    sc = scenes.actuated()
    o = oracle_after(sc, nsteps)
    g = run_engine(sc, nsteps, tb.MATH_STRICT)
    for f in FIELDS:
        assert rel(g[f], getattr(o, f)) <= 1e-13, (f, rel(g[f], getattr(o, f)))
    assert same_bits(g["rest"], o.rest)  # the _rest mutation of ACTUATED_* springs involves no sin()


@pytest.mark.parametrize("name", sorted(scenes.SCENARIOS))
@pytest.mark.parametrize("kern", KERNELS)
def test_fast_within_tolerance(name, kern):
    sc = scenes.SCENARIOS[name]()
    for nsteps, tol in ((1, 1e-12), (100, 1e-9)):
        o = oracle_after(sc, nsteps)
        g = run_engine(sc, nsteps, tb.MATH_FAST | kern)
        for f in ("pos", "vel"):
            assert rel(g[f], getattr(o, f)) <= tol, (name, nsteps, f, rel(g[f], getattr(o, f)))
        # acc = force/m amplifies the cancellation in k*(rest - L): its rounding noise lives on the force
        # scale k*L/m (~1e5 here), so it is compared on that scale
        assert rel(g["acc"], o.acc) <= 1e3 * max(tol, 1e-11), (name, nsteps, rel(g["acc"], o.acc))
        assert same_bits(g["T"], o.T)


@pytest.mark.parametrize("flags", [tb.MATH_STRICT, tb.MATH_STRICT | tb.NO_GRAPH, tb.MATH_STRICT | tb.KERNEL_CSR,
                                   tb.MATH_STRICT | tb.NO_RESIDENT, tb.MATH_STRICT | tb.NO_RESIDENT | tb.NO_GRAPH])
def test_graph_batches_and_chunking_do_not_change_results(flags):
    sc = scenes.contact_friction()
    ref = run_engine(sc, 237, tb.MATH_STRICT | tb.NO_GRAPH | tb.KERNEL_CSR)
    for chunks in (1, 3):
        g = run_engine(sc, 237, flags, chunks=chunks)
        for f in ("pos", "vel"):
            assert same_bits(g[f], ref[f]), (flags, chunks, f)


def test_pause_step_count_matches_reference_loop():
    # Simulation::execute runs while !(bp <= T): pause(1.0) at dt=1e-4 is 10,001 steps
    sc = scenes.c1_readme()
    o = sc.copy()
    n = o.run_until(1.0)
    assert n == 10001
    g = run_engine(sc, n, tb.MATH_STRICT)
    for f in FIELDS:
        assert same_bits(g[f] + 0.0, getattr(o, f) + 0.0), f
    assert same_bits(g["T"], o.T)


def test_external_force_acts_for_one_step_only():
    sc = scenes.lattice_scene((0, 0, 5), (1, 1, 1), 3)
    sc.ext[:] = (1.0, 2.0, 3.0)
    sc.fixed[4] = 1
    o = oracle_after(sc, 5)
    g = run_engine(sc, 5, tb.MATH_STRICT)
    assert same_bits(g["pos"], o.pos) and same_bits(g["vel"], o.vel)
    assert same_bits(g["extern_force"] + 0.0, o.ext + 0.0)
    assert (g["extern_force"][4] == (1.0, 2.0, 3.0)).all()          # fixed masses keep theirs (sim.cu:1294)
    assert (np.delete(g["extern_force"], 4, axis=0) == 0).all()


def test_partial_upload_download_by_index():
    sc = scenes.c1_readme()
    eng = tb.Engine(0, tb.MATH_STRICT)
    scenes.upload(eng, sc)
    eng.run(7, 0.0)
    o = oracle_after(sc, 7)
    idx = np.array([3, 100, 17, 64], dtype=np.int32)
    part = eng.get_mass_state(idx=idx)
    assert same_bits(part["pos"], o.pos[idx]) and same_bits(part["vel"], o.vel[idx])
    # container-style set: move four masses, zero their velocity, advance both sides
    newpos = o.pos[idx] + 0.125
    o.pos[idx] = newpos
    o.vel[idx] = 0.0
    eng.set_mass_state(idx=idx, pos=newpos, vel=np.zeros((4, 3)))
    eng.run(9, o.T_sim)
    o.step(9)
    st = eng.get_mass_state()
    assert same_bits(st["pos"], o.pos) and same_bits(st["vel"], o.vel)
    # spring parameter edit by index
    sidx = np.array([0, 500, 1035], dtype=np.int32)
    o.k[sidx] = (123.0, 4567.0, 8.5)
    eng.set_spring_params(idx=sidx, k=o.k[sidx])
    eng.run(4, o.T_sim)
    o.step(4)
    st = eng.get_mass_state()
    assert same_bits(st["pos"], o.pos) and same_bits(st["vel"], o.vel)
    eng.close()


def test_invalid_masses_and_null_springs_are_skipped():
    sc = scenes.lattice_scene((0, 0, 3), (1, 1, 1), 4)
    sc.valid[5] = 0
    sc.left[10] = -1
    sc.right[77] = -1
    sc.vel[:, 0] = 0.1
    o = oracle_after(sc, 20)
    g = run_engine(sc, 20, tb.MATH_STRICT)
    assert same_bits(g["pos"], o.pos) and same_bits(g["vel"], o.vel)


def test_time_step_change_and_per_mass_T():
    sc = scenes.c1_readme()
    eng = tb.Engine(0, tb.MATH_STRICT)
    scenes.upload(eng, sc)
    o = sc.copy()
    eng.run(11, o.T_sim); o.step(11)
    eng.set_globals(None, 2.5e-4); o.dt = 2.5e-4
    eng.run(13, o.T_sim); o.step(13)
    st = eng.get_mass_state()
    assert same_bits(st["T"], o.T) and same_bits(st["pos"], o.pos)
    eng.close()


def test_more_global_constraints_than_fit_inline():
    sc = scenes.lattice_scene((0, 0, 0.3), (0.5, 0.5, 0.5), 3)
    import oracle
    for j in range(7):
        sc.planes.append(oracle.make_plane((0.1 * j, 0.05 * j, 1), 0.05 * j, 0.1 * j, 0.2))
        sc.balls.append(oracle.make_ball((0.3 * j - 1, 0, 0), 0.4))
    o = oracle_after(sc, 15)
    g = run_engine(sc, 15, tb.MATH_STRICT)
    assert same_bits(g["pos"], o.pos) and same_bits(g["vel"], o.vel)


def test_full_size_c2_two_steps_against_oracle():
    # BASELINE config C2 (100^3, 1M masses, 12.7M springs) at full size: the oracle needs ~1 s per step
    sc = scenes.c2_lattice(100)
    assert sc.nm == 1000000 and sc.ns == 12731796
    o = oracle_after(sc, 2)
    g = run_engine(sc, 2, tb.MATH_STRICT)
    assert same_bits(g["pos"], o.pos) and same_bits(g["vel"] + 0.0, o.vel + 0.0)
    f = run_engine(sc, 2, tb.MATH_FAST)
    assert rel(f["pos"], o.pos) <= 1e-12 and rel(f["vel"], o.vel) <= 1e-12
    assert f["info"]["algorithmic_bytes_per_step"] == 409563104.0


def test_determinism_run_to_run():
    sc = scenes.c2_lattice(40)
    a = run_engine(sc, 120, tb.MATH_FAST)
    b = run_engine(sc, 120, tb.MATH_FAST)
    assert same_bits(a["pos"], b["pos"]) and same_bits(a["vel"], b["vel"])


@pytest.mark.parametrize("kernel", [0, tb.KERNEL_CSR])
@pytest.mark.parametrize("nsteps", [1, 10, 137])
def test_magnet_pass_strict_is_bit_identical(nsteps, kernel):
    """Occupancy-grid + magnet pass (reference src/sim.cu:865-906,1223-1281) on the RobotLink scene
    whose reference-CUDA dump pins the oracle (tests/golden/ref_cuda/ref_magnets.bin); in front of the tiled kernel (the
    pass reads the tile-blocked state through the mass -> slot map) and of the generic one."""
    sc = scenes.magnets()
    o = oracle_after(sc, nsteps)
    g = run_engine(sc, nsteps, tb.MATH_STRICT | kernel, chunks=2 if nsteps > 10 else 1)
    assert g["info"]["kernel"] == (0 if kernel else 1)
    for f in FIELDS:
        assert same_bits(g[f] + 0.0, getattr(o, f) + 0.0), (f, rel(g[f], getattr(o, f)))
    assert same_bits(g["T"], o.T)
    assert same_bits(g["extern_force"] + 0.0, o.ext + 0.0)


def magnet_cloud(n=4000, seed=12345):
    """n free magnetic masses dropped at random into a 0.6 x 0.6 x 0.3 m box (about 18 per occupancy-grid cell, many pairs
    inside the 0.14 m cut-off, some shells intersecting) + a sparse random spring net; one corner cell is crowded
    beyond the 128-per-cell capacity."""
    rng = np.random.default_rng(seed)
    pos = rng.uniform((-0.3, -0.3, 0.2), (0.3, 0.3, 0.5), size=(n, 3))
    crowd = 150
    pos[:crowd, 0] = rng.uniform(0.401, 0.439, crowd)     # one cell: x in [0.40, 0.44), y in [0.40, 0.44)
    pos[:crowd, 1] = rng.uniform(0.401, 0.439, crowd)
    pos[:crowd, 2] = rng.uniform(0.0, 3.0, crowd)
    for j in range(4):                                    # outside the grid: coordinates clamp into cell (255, 0)
        pos[crowd + j] = (6.0 + 0.01 * j, -7.0, 0.3 + 0.01 * j)
    left = rng.integers(0, n, size=n // 2).astype(np.int32)
    right = ((left + 1 + rng.integers(0, n - 1, size=n // 2)) % n).astype(np.int32)
    import oracle
    rest = np.linalg.norm(pos[left] - pos[right], axis=1)
    sc = oracle.Scene(pos, left, right, rest, k=50.0, m=0.05)
    sc.rad[:] = rng.choice([0.0, 0.01, 0.015], size=n)
    sc.stiffness[:] = rng.choice([1000.0, 5000.0], size=n)
    sc.max_mag[:] = rng.choice([0.0, 0.002, 0.004], size=n)
    sc.mag_scale[:] = rng.choice([0.0, 0.5, 1.0], size=n)
    sc.fixed[rng.integers(0, n, size=n // 10)] = 1
    sc.ext[::7] = (0.01, -0.02, 0.03)                     # uploaded one-shot force is folded into the pass's output
    sc.planes.append(oracle.make_plane((0, 0, 1), 0, 0.5, 0.6))
    sc.magnets = True
    return sc


@pytest.mark.parametrize("flags", [tb.MATH_STRICT, tb.MATH_STRICT | tb.NO_GRAPH, tb.MATH_STRICT | tb.KERNEL_CSR])
def test_magnet_cloud_strict_is_bit_identical(flags):
    """Free-free pairs, cell overflow (> 128 masses in a cell keeps the 128 lowest indices), clamped cell coordinates:
    the engine's deterministic pass (start-of-step positions, ascending index inside a cell) against the oracle's."""
    sc = magnet_cloud()
    o = oracle_after(sc, 60)
    g = run_engine(sc, 60, flags)
    assert g["info"]["kernel"] == (0 if flags & tb.KERNEL_CSR else 1)   # isolated masses and a random spring net fit the tiles
    for f in FIELDS:
        assert same_bits(g[f] + 0.0, getattr(o, f) + 0.0), (f, rel(g[f], getattr(o, f)))
    assert same_bits(g["extern_force"] + 0.0, o.ext + 0.0)


@pytest.mark.parametrize("shift", [(2.0e4, 0.0, 0.0), (0.0, -3.0e5, 1.0e3)])
def test_magnet_pass_far_from_the_origin(shift):
    """Masses tens of kilometres out: every cell coordinate on that axis clamps to the grid's border, and the pair
    kernel's single-precision neighbour pre-test is switched off (its rounding would exceed the gate's slack) -- every
    listed neighbour goes through the exact fp64 test.  Still bit-identical to the oracle."""
    sc = magnet_cloud(n=700, seed=99)
    sc.pos += np.array(shift)
    sc.planes.clear()
    o = oracle_after(sc, 20)
    for kernel in (0, tb.KERNEL_CSR):
        g = run_engine(sc, 20, tb.MATH_STRICT | kernel)
        for f in FIELDS:
            assert same_bits(g[f] + 0.0, getattr(o, f) + 0.0), (kernel, f, rel(g[f], getattr(o, f)))
    moved = oracle_after(sc, 1)
    still = sc.copy(); still.magnets = False; still.step(1)
    assert not same_bits(moved.vel, still.vel)            # pairs inside the cut-off exist out there too


def test_magnet_cloud_fast_within_tolerance():
    sc = magnet_cloud()
    for nsteps, tol in ((1, 1e-12), (30, 1e-9)):
        o = oracle_after(sc, nsteps)
        g = run_engine(sc, nsteps, tb.MATH_FAST)
        for f in ("pos", "vel"):
            assert rel(g[f], getattr(o, f)) <= tol, (nsteps, f, rel(g[f], getattr(o, f)))


def test_plain_masses_feel_a_magnetic_neighbour():
    """A mass with rad = 0 and max_mag_force = 0 still gets a shell-intersection force from a neighbour with rad > 0
    (its own stiffness scales it, src/sim.cu:1230-1232): the lazily stored uniform stiffness must reach the pass."""
    sc = scenes.lattice_scene((0, 0, 1), (0.02, 0.02, 0.02), 3)   # 27 masses 1 cm apart
    sc.stiffness[:] = 1000.0
    eng = tb.Engine(0, tb.MATH_STRICT)
    scenes.upload(eng, sc)                                 # uniform, non-magnetic: nothing stored, pass elided
    eng.run(3, 0.0)
    o = oracle_after(sc, 3)
    one = np.array([13], dtype=np.int32)
    eng.set_mass_magnet(idx=one, rad=np.array([0.015]), stiffness=np.array([5000.0]), max_mag_force=np.array([0.0]),
                        mag_scale=np.array([0.0]))
    o.rad[13] = 0.015; o.stiffness[13] = 5000.0; o.magnets = True
    eng.run(5, o.T_sim); o.step(5)
    st = eng.get_mass_state()
    eng.close()
    assert same_bits(st["pos"], o.pos) and same_bits(st["vel"] + 0.0, o.vel + 0.0)
    q = o.copy(); q.magnets = False; q.step(1)
    o.step(1)
    assert not same_bits(q.vel, o.vel)                     # the pass does act on this scene


def test_headless_export_buffers():
    """Row N3: fp32 packed xyz + line list in device memory (updateVertices / updateIndices, src/sim.cu:1944-1980)
    without a graphics API; also after steps on the tiled kernel (state is brought back to API order first)."""
    sc = scenes.contact_friction()
    sc.valid[7] = 0
    sc.left[3] = -1
    eng = tb.Engine(0, tb.MATH_FAST)
    scenes.upload(eng, sc)
    eng.run(25, 0.0)
    gone = np.array([10, 11], dtype=np.int32)
    eng.remove_springs(gone)
    st = eng.get_mass_state()
    xyz = eng.read_device(eng.export_positions_f32_dev(), (sc.nm, 3), np.float32)
    assert (xyz == st["pos"].astype(np.float32)).all()
    idx = eng.read_device(eng.export_indices_dev(), (sc.ns, 2), np.uint32)
    want = np.stack([sc.left, sc.right], axis=1).astype(np.int64)
    skip = (sc.left < 0) | (sc.right < 0) | (sc.left == 7) | (sc.right == 7)
    skip[gone] = True
    want[skip] = 0
    assert (idx == want.astype(np.uint32)).all()
    eng.close()


VERLET = tb.INTEGRATOR_VERLET


@pytest.mark.parametrize("name", ["c1_readme", "contact_friction", "local_constraints", "cube_beam", "magnets"])
@pytest.mark.parametrize("nsteps", [1, 10, 137])
@pytest.mark.parametrize("kernel", [0, tb.KERNEL_CSR, tb.NO_RESIDENT])
def test_verlet_strict_is_bit_identical(name, nsteps, kernel):
    """The reference's VERLET build option (src/sim.cu:1350-1354): acc is persistent state, no clamp.
    Both step kernels (tiled: the previous acceleration is read from / stored to the canonical array by the mass's
    thread; generic)."""
    sc = scenes.SCENARIOS[name]()
    sc.verlet = True
    o = oracle_after(sc, nsteps)
    g = run_engine(sc, nsteps, tb.MATH_STRICT | VERLET | kernel, chunks=2 if nsteps > 10 else 1)
    assert g["info"]["kernel"] == (0 if kernel == tb.KERNEL_CSR else 1)
    for f in FIELDS:
        assert same_bits(g[f] + 0.0, getattr(o, f) + 0.0), (name, f, rel(g[f], getattr(o, f)))
    assert same_bits(g["T"], o.T)


@pytest.mark.parametrize("name", ["c1_readme", "contact_friction", "actuated", "local_constraints"])
def test_verlet_fast_within_tolerance(name):
    sc = scenes.SCENARIOS[name]()
    sc.verlet = True
    for nsteps, tol in ((1, 1e-12), (100, 1e-9)):
        o = oracle_after(sc, nsteps)
        g = run_engine(sc, nsteps, tb.MATH_FAST | VERLET)
        for f in ("pos", "vel"):
            assert rel(g[f], getattr(o, f)) <= tol, (name, nsteps, f, rel(g[f], getattr(o, f)))
        assert rel(g["acc"], o.acc) <= 1e3 * max(tol, 1e-11), (name, nsteps, rel(g["acc"], o.acc))


@pytest.mark.parametrize("nsteps", [1, 2, 3, 50, 51])
def test_resident_kernel_equals_streaming_kernel_bitwise(nsteps):
    """Odd and even step counts (the resident kernel leaves the state in either ping-pong buffer), chunked runs, an
    uploaded one-shot external force (consumed by one streaming step first), actuated + damped springs."""
    sc = scenes.actuated()
    sc.type[sc.type >= 4] = 0                    # ACTUATED_* springs run on the generic kernel
    sc.ext[::5] = (0.3, -0.2, 0.1)
    a = run_engine(sc, nsteps, tb.MATH_STRICT, chunks=2 if nsteps > 3 else 1)
    b = run_engine(sc, nsteps, tb.MATH_STRICT | tb.NO_RESIDENT, chunks=2 if nsteps > 3 else 1)
    assert a["info"]["kernel"] == 1 and b["info"]["kernel"] == 1
    for f in FIELDS + ("T", "extern_force"):
        assert same_bits(a[f] + 0.0, b[f] + 0.0), f
    fa = run_engine(sc, nsteps, tb.MATH_FAST)
    fb = run_engine(sc, nsteps, tb.MATH_FAST | tb.NO_RESIDENT)
    for f in ("pos", "vel"):
        assert rel(fa[f], fb[f]) <= 1e-12, f


def test_more_actuation_frequencies_than_the_sine_table_holds():
    """Every ACTIVE_* spring its own omega (> 64 distinct values): sin(omega t) is evaluated per spring, on both tiled
    kernels; the oracle's glibc sin may differ from the device's by an ulp."""
    sc = scenes.actuated()
    sc.type[sc.type >= 4] = 0
    sc.omega[:] = 10.0 + 0.37 * np.arange(sc.ns)
    o = oracle_after(sc, 40)
    a = run_engine(sc, 40, tb.MATH_STRICT)
    b = run_engine(sc, 40, tb.MATH_STRICT | tb.NO_RESIDENT)
    for f in ("pos", "vel"):
        assert same_bits(a[f], b[f]), f
        assert rel(a[f], getattr(o, f)) <= 1e-13, (f, rel(a[f], getattr(o, f)))
    fa = run_engine(sc, 40, tb.MATH_FAST)
    for f in ("pos", "vel"):
        assert rel(fa[f], getattr(o, f)) <= 1e-10, (f, rel(fa[f], getattr(o, f)))


def test_resident_kernel_runs_independent_systems_as_separate_clusters():
    """Four independent 10 x 10 x 5 robots in one engine (bricks of 5^3: four tiles each): the resident kernel launches
    one cluster per robot; bit-identical to the streaming kernel and to the oracle."""
    import oracle
    from titan_b200 import lattice
    parts, hints, nm1 = [], [], 500
    for r in range(4):
        L = lattice.build_lattice((1.5 * r, 0.0, 0.25 + 0.01 * r), (0.9, 0.9, 0.4), 10, 10, 5)
        parts.append(L)
        h = lattice.brick_hint(10, 10, 5)
        hints.append(h + r * (int(h.max()) + 1))
    pos = np.vstack([p["pos"] for p in parts])
    left = np.concatenate([p["left"] + r * nm1 for r, p in enumerate(parts)])
    right = np.concatenate([p["right"] + r * nm1 for r, p in enumerate(parts)])
    rest = np.concatenate([p["rest"] for p in parts])
    sc = oracle.Scene(pos, left, right, rest)
    sc.vel[:, 2] = -0.3
    sc.planes.append(oracle.make_plane((0, 0, 1), 0, 0.5, 0.6))
    o = oracle_after(sc, 61)
    out = []
    for flags in (tb.MATH_STRICT, tb.MATH_STRICT | tb.NO_RESIDENT):
        eng = tb.Engine(0, flags)
        scenes.upload(eng, sc)
        eng.set_tile_hint(np.concatenate(hints).astype(np.int32))
        eng.run(61, 0.0)
        out.append(eng.get_mass_state())
        assert eng.info()["n_tiles"] == 16 and eng.info()["kernel"] == 1
        eng.close()
    for f in ("pos", "vel"):
        assert same_bits(out[0][f], out[1][f]) and same_bits(out[0][f] + 0.0, getattr(o, f) + 0.0), f


def test_verlet_tiled_dictionary_kernel_within_tolerance(monkeypatch):
    """Verlet on the configuration bench.py runs (fast arithmetic, one thread per mass, dictionary records)."""
    monkeypatch.setenv("TB_TPM", "1")               # one thread per mass also for this small scene
    sc = scenes.c2_lattice(24)
    sc.verlet = True
    o = oracle_after(sc, 100)
    g = run_engine(sc, 100, tb.MATH_FAST | VERLET)
    assert g["info"]["kernel"] == 1 and g["info"]["dict_records"] > 0
    for f in ("pos", "vel"):
        assert rel(g[f], getattr(o, f)) <= 1e-9, (f, rel(g[f], getattr(o, f)))
    assert rel(g["acc"], o.acc) <= 1e-8 * 1e3


def test_verlet_state_upload_resets_acceleration():
    """set(Mass*) replaces the whole CUDA_MASS by a freshly constructed one whose acc is zero (src/sim.cu:603-616,
    src/mass.cu:74-100); under Verlet that acc feeds the next velocity update."""
    sc = scenes.c1_readme()
    sc.verlet = True
    eng = tb.Engine(0, tb.MATH_STRICT | VERLET)
    scenes.upload(eng, sc)
    eng.run(20, 0.0)
    o = oracle_after(sc, 20)
    idx = np.array([0, 31, 124], dtype=np.int32)
    st = eng.get_mass_state(idx=idx)
    eng.set_mass_state(idx=idx, pos=st["pos"], vel=st["vel"])
    o.acc[idx] = 0.0
    eng.run(20, o.T_sim); o.step(20)
    st = eng.get_mass_state()
    eng.close()
    assert same_bits(st["pos"], o.pos) and same_bits(st["vel"], o.vel) and same_bits(st["acc"] + 0.0, o.acc + 0.0)


@pytest.mark.parametrize("flags", [tb.MATH_STRICT, tb.MATH_STRICT | tb.KERNEL_CSR, tb.MATH_FAST])
def test_removed_springs_are_bit_identical_to_absent_springs(flags):
    """tb_remove_springs (tombstones; BASELINE config C5 / Simulation::deleteSpring after start, src/sim.cu:382-412):
    the system steps exactly as if the springs had never existed, also after a later full k upload."""
    sc = scenes.SCENARIOS["contact_friction"]()
    gone = np.arange(3, sc.ns, 17).astype(np.int32)
    keep = np.setdiff1d(np.arange(sc.ns), gone)
    # oracle: 40 steps with all springs, then the springs physically absent, stiffer survivors
    o = sc.copy()
    o.step(40)
    import oracle
    o2 = oracle.Scene(o.pos.copy(), sc.left[keep], sc.right[keep], sc.rest[keep], k=1.0, m=0.1, dt=sc.dt, g=sc.g)
    o2.vel[:] = o.vel
    o2.k[:] = sc.k[keep] * 1.25
    o2.damping[:] = sc.damping[keep]
    o2.planes.extend(sc.planes)
    o2.T_sim = o.T_sim
    o2.step(60)
    eng = tb.Engine(0, flags)
    scenes.upload(eng, sc)
    eng.run(40, 0.0)
    eng.remove_springs(gone)
    eng.set_spring_params(k=sc.k * 1.25)          # touches the tombstones too: they must stay dead
    eng.run(60, o.T_sim)
    st = eng.get_mass_state()
    assert eng.info()["ns_removed"] == gone.size
    eng.close()
    tol = 0.0 if flags & tb.MATH_STRICT else 1e-9
    for f in ("pos", "vel"):
        assert rel(st[f], getattr(o2, f)) <= tol, (f, rel(st[f], getattr(o2, f)))


@pytest.mark.parametrize("name", ["c2_small", "contact_friction", "active"])
def test_dictionary_records_are_bit_identical_to_plain_records(name, monkeypatch):
    """Dictionary mode of the tiled kernel (springs with bit-identical parameters share a record)
    changes where a spring's (k, rest, damping, omega) is read from, not the arithmetic: results are bit-identical
    to the one-record-per-spring layout -- also after parameter uploads that split a class (records rebuilt in
    place) and after an upload that makes every spring different (falls back to plain records)."""
    if name == "c2_small":
        sc = scenes.c2_lattice(24)
    elif name == "active":
        sc = scenes.SCENARIOS["contact_friction"]()
        sc.type[:] = np.arange(sc.ns) % 4          # PASSIVE, ACTIVE_*, ...: table mode of the sine
        sc.omega[:] = 20.0 * (1 + np.arange(sc.ns) % 3)
    else:
        sc = scenes.SCENARIOS[name]()
    monkeypatch.setenv("TB_TPM", "1")               # one thread per mass also for these small scenes
    engines = []
    for dict_on in ("0", "1"):
        monkeypatch.setenv("TB_DICT", dict_on)      # read when the layout is built, i.e. at the first run
        eng = tb.Engine(0, tb.MATH_FAST | tb.NO_RESIDENT)   # the streaming kernel also for the small scenes
        scenes.upload(eng, sc)
        eng.run(60, 0.0)
        engines.append(eng)
    monkeypatch.delenv("TB_DICT")
    plain, dic = engines
    assert plain.info()["dict_records"] == 0 and dic.info()["dict_records"] > 0
    assert plain.info()["kernel"] == 1 and dic.info()["threads_per_mass"] == 1

    def compare(tag):
        a, b = plain.get_mass_state(), dic.get_mass_state()
        for f in ("pos", "vel"):
            assert same_bits(a[f], b[f]), (tag, f, rel(a[f], b[f]))

    compare("initial classes")
    t = 0.0
    for _ in range(60):
        t += sc.dt
    # (1) a few springs get their own stiffness: classes split, the dictionary is rebuilt with the slots kept
    k2 = sc.k.copy()
    k2[5:40:3] *= 1.5
    for e in engines:
        e.set_spring_params(k=k2)
        e.run(50, t)
    compare("split classes")
    assert dic.info()["dict_records"] > 0
    for _ in range(50):
        t += sc.dt
    # (2) same values uploaded again: nothing to rebuild
    for e in engines:
        e.set_spring_params(k=k2)
        e.run(10, t)
    compare("same values again")
    for _ in range(10):
        t += sc.dt
    # (3) every spring different: more classes than a tile's byte map addresses -> plain records from now on
    k3 = sc.k * (1.0 + 1e-3 * np.arange(sc.ns) / sc.ns)
    for e in engines:
        e.set_spring_params(k=k3)
        e.run(50, t)
    compare("all different")
    # ... but the rest lengths (and damping, omega) still fall into a few classes: split dictionary -- those stay shared,
    # the stiffness becomes one value per record slot (BASELINE config C5 rewrites every k every 100 steps)
    assert plain.info()["dict_records"] == 0 and plain.info()["dict_split_k"] == 0
    if sc.ns > 4096:
        assert dic.info()["dict_records"] > 0 and dic.info()["dict_split_k"] == 1
    for _ in range(50):
        t += sc.dt
    # (4) in that mode a new k for every spring needs no rebuild, and removed springs (tombstones: k = 0) stay removed
    gone = np.arange(7, sc.ns, 23).astype(np.int32)
    k4 = k3 * (0.9 + 0.2 * ((np.arange(sc.ns) * 2654435761 % 1000) / 1000.0))
    for e in engines:
        e.remove_springs(gone)
        e.set_spring_params(k=k4)
        e.run(50, t)
    compare("k rewrite + tombstones")
    if sc.ns > 4096:
        assert dic.info()["dict_split_k"] == 1
    for _ in range(50):
        t += sc.dt
    # (5) a rest-length edit that splits a class of the split dictionary: records rebuilt, results unchanged
    r5 = sc.rest.copy()
    r5[11:300:7] *= 1.001
    for e in engines:
        e.set_spring_params(rest=r5)
        e.run(30, t)
    compare("rest edit in split mode")
    for e in engines:
        e.close()


def lattice_with_actuated_links(n=16):
    """An n^3 lattice resting on a plane plus ACTUATED_EXPAND / ACTUATED_CONTRACT springs (what RobotLink::expand /
    contract switch on, src/object.cu:386-421) between some of its masses, one of them across two bricks."""
    sc = scenes.c2_lattice(n)
    import oracle
    pairs = [(0, 1), (5, 5 + n * n), (n ** 3 - 1, n ** 3 - 2), (7 * n * n + 3 * n + 2, 12 * n * n + 3 * n + 2)]
    left = np.concatenate([sc.left, [a for a, _ in pairs]]).astype(np.int32)
    right = np.concatenate([sc.right, [b for _, b in pairs]]).astype(np.int32)
    rest_new = [float(np.linalg.norm(sc.pos[a] - sc.pos[b])) for a, b in pairs]
    rest = np.concatenate([sc.rest, rest_new])
    s2 = oracle.Scene(sc.pos, left, right, rest, k=10000.0, m=0.1, dt=sc.dt, g=sc.g)
    s2.planes.extend(sc.planes)
    for q in range(len(pairs)):
        s = sc.ns + q
        s2.type[s] = 4 + (q % 2)
        s2.k[s] = 500.0
        s2.l_max[s] = rest_new[q] * 1.01
        s2.l_min[s] = rest_new[q] * 0.99
        s2.rate[s] = 0.02 + 0.01 * q
    return s2


def test_actuated_springs_stay_on_the_tiled_kernel(monkeypatch):
    """ACTUATED_* springs (device-side _rest mutation, src/sim.cu:1174-1181) no longer send the system to the generic
    kernel: the mutation runs over the list of such springs and writes their private record in the tile blobs."""
    # (1) the small all-types scene, strict: bit-identical rest lengths, state within the sin() ulp
    sc = scenes.actuated()
    for nsteps in (1, 60):
        o = oracle_after(sc, nsteps)
        g = run_engine(sc, nsteps, tb.MATH_STRICT | tb.NO_RESIDENT)
        assert g["info"]["kernel"] == 1
        assert same_bits(g["rest"], o.rest)
        for f in ("pos", "vel"):
            assert rel(g[f], getattr(o, f)) <= 1e-13, (nsteps, f)
    # (2) a lattice with a few actuated links, fast arithmetic, one thread per mass: dictionary records for the lattice
    # springs, a private record per actuated spring
    monkeypatch.setenv("TB_TPM", "1")
    sc = lattice_with_actuated_links(16)
    for flags in (tb.MATH_STRICT, tb.MATH_FAST):
        for nsteps, tol in ((1, 1e-12), (100, 1e-9)):
            o = oracle_after(sc, nsteps)
            g = run_engine(sc, nsteps, flags | tb.NO_RESIDENT, chunks=2 if nsteps > 1 else 1)
            assert g["info"]["kernel"] == 1
            if flags == tb.MATH_FAST:
                assert g["info"]["dict_records"] > 0
            assert same_bits(g["rest"], o.rest), (flags, nsteps)
            for f in ("pos", "vel"):
                e = rel(g[f], getattr(o, f))
                assert e <= (0.0 if flags == tb.MATH_STRICT else tol), (flags, nsteps, f, e)
    # (3) the generic kernel walks the same list
    o = oracle_after(sc, 30)
    g = run_engine(sc, 30, tb.MATH_STRICT | tb.KERNEL_CSR)
    assert g["info"]["kernel"] == 0 and same_bits(g["rest"], o.rest) and same_bits(g["pos"], o.pos)


@pytest.mark.parametrize("name", ["c1_readme", "contact_friction", "local_constraints", "cube_beam", "actuated"])
@pytest.mark.parametrize("kern", [0, tb.KERNEL_CSR], ids=["tiled", "generic"])
def test_rk2_integrator(name, kern):
    """TB_INTEGRATOR_RK2 (the reference's -DRK2 build with its call sites repaired, src/sim.cu:1336-1349,1778-1799): two force
    evaluations per step, T += 0.5 dt per half.  Strict arithmetic is bit-identical to the oracle's RK2 branch (1e-13 where
    sin() is involved), fast arithmetic within 1e-12 / 1e-9; per-mass T and the rest lengths of ACTUATED_* springs (advanced in
    both halves, as computeSpringForces does) are bit-identical."""
    sc = scenes.SCENARIOS[name]()
    sc.rk2 = True
    for nsteps in (1, 60):
        o = oracle_after(sc, nsteps)
        g = run_engine(sc, nsteps, tb.MATH_STRICT | tb.INTEGRATOR_RK2 | kern)
        assert g["info"]["kernel"] == (0 if kern else 1)
        assert same_bits(g["T"], o.T) and same_bits(g["rest"], o.rest)
        for f in FIELDS:
            if name == "actuated":
                assert rel(g[f], getattr(o, f)) <= (1e-13 if f != "acc" else 1e-9), (name, nsteps, f)
            else:
                assert same_bits(g[f] + 0.0, getattr(o, f) + 0.0), (name, nsteps, f)
        f_ = run_engine(sc, nsteps, tb.MATH_FAST | tb.INTEGRATOR_RK2 | kern, chunks=2 if nsteps > 1 else 1)
        for f in ("pos", "vel"):
            assert rel(f_[f], getattr(o, f)) <= (1e-12 if nsteps == 1 else 1e-9), (name, nsteps, f)
        assert same_bits(f_["T"], o.T)


def test_rk2_refuses_what_it_is_not_built_for():
    sc = scenes.magnets()
    eng = tb.Engine(0, tb.MATH_FAST | tb.INTEGRATOR_RK2)
    scenes.upload(eng, sc)
    with pytest.raises(tb.TitanError) as ei:
        eng.run(1, 0.0)
    assert ei.value.status == 4 and "RK2" in str(ei.value)
    eng.close()
