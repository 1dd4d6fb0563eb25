"""tb_group: several engines in one process behind one handle (include/titan_b200.h: tb_create_partitioned,
tb_batch_create, tb_group_*), against ONE engine stepping the same system.  Bit-identical under TB_MATH_STRICT (each engine
keeps the global spring order per mass), <= 1e-12 under TB_MATH_FAST with one thread per mass on both sides.

On a single-GPU box all engines of a group sit on device 0 (small persistent grids, TB_GRID, so that they are co-resident);
with two or more GPUs the same checks also run with one engine per device over peer access.
"""
import numpy as np
import pytest

import scenes
import titan_b200 as tb
from test_partitioned import make_scene

pytestmark = pytest.mark.gpu


def run_single(sc, steps, flags):
    eng = tb.Engine(0, flags)
    scenes.upload(eng, sc)
    eng.run(steps, 0.0)
    st = eng.get_mass_state()
    st["rest"] = eng.get_spring_rest()
    eng.close()
    return st


def devices_cases():
    cases = [[0, 0], [0, 0, 0]]
    if tb.device_count() >= 2:
        cases.append([0, 1])
    if tb.device_count() >= 4:
        cases.append([0, 1, 2, 3])
    return cases


@pytest.mark.parametrize("kind", ["plain", "damped_active", "actuated"])
@pytest.mark.parametrize("devices", devices_cases(), ids=lambda d: "dev" + "".join(map(str, d)))
@pytest.mark.parametrize("flags,tol", [(tb.MATH_STRICT, 0.0), (tb.MATH_FAST, 1e-12)])
def test_partitioned_group_matches_single_engine(kind, devices, flags, tol, monkeypatch):
    monkeypatch.setenv("TB_GRID", "16")
    monkeypatch.setenv("TB_TPM", "1")
    sc, _ = make_scene(kind)
    steps = 120
    want = run_single(sc, steps, flags)
    grp = tb.Group(devices, flags)
    scenes.upload(grp, sc)                       # owner == NULL: equal contiguous index ranges = i-slabs of the lattice
    t = 0.0
    for chunk in (1, 7, 62, 50):
        grp.run_async(chunk, t)
        grp.sync()
        for _ in range(chunk):
            t += sc.dt
    got = grp.get_mass_state()
    rest = grp.get_spring_rest()
    infos = grp.info()
    grp.close()
    assert len(infos) == len(devices) and all(i["kernel"] == 1 for i in infos)
    assert sum(i["nm"] for i in infos) > sc.nm      # ghost copies exist: the springs are cut
    for f in ("pos", "vel"):
        err = float(np.max(np.abs(got[f] - want[f])))
        assert err <= tol, (kind, devices, f, err)
    assert rest.tobytes() == want["rest"].tobytes()
    assert got["T"].tobytes() == want["T"].tobytes()


def robots(n_robots):
    import oracle
    one = scenes.lattice_scene((0, 0, 0.3), (0.4, 0.4, 0.4), 5)
    pos = np.concatenate([one.pos + (1.0 * r, 0.0, 0.01 * r) for r in range(n_robots)])
    left = np.concatenate([one.left + r * one.nm for r in range(n_robots)]).astype(np.int32)
    right = np.concatenate([one.right + r * one.nm for r in range(n_robots)]).astype(np.int32)
    sc = oracle.Scene(pos, left, right, np.tile(one.rest, n_robots))
    s = np.arange(sc.ns)
    sc.type[:] = np.where(s % 3 == 1, 2, np.where(s % 3 == 2, 3, 0))
    sc.omega[:] = 4 * np.pi
    sc.planes.append(oracle.make_plane((0, 0, 1), 0, 0.5, 0.6))
    return sc, one.nm


@pytest.mark.parametrize("devices", devices_cases(), ids=lambda d: "dev" + "".join(map(str, d)))
def test_batch_group_shards_whole_simulations(devices, monkeypatch):
    monkeypatch.setenv("TB_GRID", "16")
    monkeypatch.setenv("TB_TPM", "1")
    monkeypatch.setenv("TB_RESIDENT", "0")
    sc, nm1 = robots(7)
    want = run_single(sc, 80, tb.MATH_STRICT)
    grp = tb.Group(devices, tb.MATH_STRICT, batch_sizes=[nm1] * 7)
    scenes.upload(grp, sc)
    grp.run(80, 0.0)
    got = grp.get_mass_state()
    infos = grp.info()
    grp.close()
    assert sum(i["nm"] for i in infos) == sc.nm       # whole robots per engine: no ghosts, nothing exchanged
    assert all(i["nm"] % nm1 == 0 and i["nm"] > 0 for i in infos)
    assert got["pos"].tobytes() == want["pos"].tobytes() and got["vel"].tobytes() == want["vel"].tobytes()


def test_group_routes_partial_uploads_and_edits(monkeypatch):
    monkeypatch.setenv("TB_GRID", "16")
    sc, (nx, ny, nz) = make_scene("plain")
    eng = tb.Engine(0, tb.MATH_STRICT)
    grp = tb.Group([0, 0, 0], tb.MATH_STRICT)
    for x in (eng, grp):
        scenes.upload(x, sc)
        x.run(20, 0.0)
    t = 0.0
    for _ in range(20):
        t += sc.dt
    # masses on both sides of a cut (an i-plane boundary of the 3-way split), by index; a custom owner-independent order
    plane = ny * nz
    idx = np.array([4 * plane - 1, 4 * plane, 4 * plane + 1, 3, 8 * plane + 5, sc.nm - 1], dtype=np.int32)
    a, b = eng.get_mass_state(idx=idx), grp.get_mass_state(idx=idx)
    assert a["pos"].tobytes() == b["pos"].tobytes() and a["vel"].tobytes() == b["vel"].tobytes()
    newpos = a["pos"] + 0.01
    sidx = np.array([0, sc.ns // 2, sc.ns - 1, 17], dtype=np.int32)
    gone = np.arange(5, sc.ns, 97).astype(np.int32)
    for x in (eng, grp):
        x.set_mass_state(idx=idx, pos=newpos, vel=np.zeros((idx.size, 3)))
        x.set_spring_params(idx=sidx, k=np.array([123.0, 4567.0, 8.5, 2e4]))
        x.remove_springs(gone)
        x.run(40, t)
    a, b = eng.get_mass_state(), grp.get_mass_state()
    assert a["pos"].tobytes() == b["pos"].tobytes() and a["vel"].tobytes() == b["vel"].tobytes()
    eng.close()
    grp.close()
