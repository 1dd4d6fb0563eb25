"""The oracle against state dumps written by the REFERENCE's own CUDA kernels on a B200 (tests/golden/ref_cuda,
see tests/golden/README.md).  This pins the end-to-end step semantics of oracle/titan_oracle.c -- step count of
pause(t), spring/mass kernels, constraints, actuation, one-shot external force, fixed masses, per-mass T.

Tolerance: the reference sums spring forces with fp64 atomicAdd (order varies run to run) and its nvcc build
contracts a*b+c into FMA, while the oracle is unfused and sums in ascending spring order.  Differences are
therefore rounding-level: <= 1e-9 relative (north_star's per-step bound, here after up to 1000 steps);
observed values are printed by `pytest -s`.
"""
import os
import struct

import numpy as np
import pytest

import scenes

GOLD = os.path.join(os.path.dirname(__file__), "golden", "ref_cuda")
DTYPES = {0: np.float64, 1: np.int32, 2: np.uint8}


def read_dump(path):
    out, data = {}, open(path, "rb").read()
    o = 0
    while o < len(data):
        (n,) = struct.unpack_from("<I", data, o); o += 4
        name = data[o:o + n].decode(); o += n
        dt, cnt = struct.unpack_from("<IQ", data, o); o += 12
        a = np.frombuffer(data, dtype=DTYPES[dt], count=cnt, offset=o); o += a.nbytes
        out[name] = a
    return out


def rel(a, b):
    return float(np.max(np.abs(a - b)) / max(1.0, float(np.max(np.abs(b)))))


def check_against_dump(name, path, verlet=False, rk2=False):
    d = read_dump(path)
    sc = scenes.SCENARIOS[name]()
    sc.verlet = verlet
    sc.rk2 = rk2
    # 1. both descriptions of the scenario agree on the initial scene, bit for bit
    same = lambda a, b: (np.ascontiguousarray(a).ravel().view(np.uint64) == b.view(np.uint64)).all()  # noqa: E731
    assert same(sc.pos, d["init.pos"]) and same(sc.vel, d["init.vel"]) and same(sc.m, d["init.m"])
    assert (sc.left == d["init.left"]).all() and (sc.right == d["init.right"]).all()
    assert (sc.type == d["init.type"]).all()
    for f in ("k", "rest", "damping", "omega", "l_max", "l_min", "rate"):
        assert same(getattr(sc, f), d["init." + f]), f
    if sc.magnets:   # magnet fields as the reference's RobotLink / Mass constructors set them
        for f in ("rad", "stiffness", "max_mag", "mag_scale"):
            assert same(getattr(sc, f), d["init." + f]), f
    else:
        assert "init.rad" not in d
    # 2. every snapshot: same number of steps (time), same state within tolerance
    worst = {}
    for i, t_pause in enumerate(scenes.PAUSES[name]):
        p = "snap%d." % i
        assert d[p + "time"][0] == t_pause
        n = sc.run_until(t_pause)
        assert sc.T_sim == d[p + "time"][1], (name, i, n)          # identical T += dt accumulation
        # RK2 has no speed clamp and evaluates the contact laws twice per step: on the frictional scene the sliding regime of
        # single masses flips between steps 91 and 399 (3e-13 in vel at step 91, 1e-2 at step 399 -- the discontinuity of
        # src/object.cu:82-98); the last snapshot of that scene is only bounded loosely
        loose = rk2 and name == "contact_friction" and i == len(scenes.PAUSES[name]) - 1
        for f in ("pos", "vel", "acc"):
            e = rel(getattr(sc, f).ravel(), d[p + f])
            if loose:
                assert f != "pos" or e <= 1e-4, (name, i, f, e)
                continue
            worst[f] = max(worst.get(f, 0.0), e)
        assert rel(sc.T, d[p + "T"]) == 0.0                         # per-mass T (fixed masses frozen)
        assert rel(sc.rest, d[p + "rest"]) <= 1e-15                 # ACTUATED_* rest mutation
    print(name, "rk2" if rk2 else "verlet" if verlet else "euler", worst)
    assert worst["pos"] <= 1e-9 and worst["vel"] <= 1e-9
    # acc = force/m amplifies cancellation in k*(rest-L): compare on the force scale k*L/m ~ 1e5
    assert worst["acc"] <= 1e-9 * 1e3


@pytest.mark.parametrize("name", sorted(scenes.SCENARIOS))
def test_oracle_matches_reference_cuda(name):
    check_against_dump(name, os.path.join(GOLD, "ref_%s.bin" % name))


@pytest.mark.parametrize("name", scenes.VERLET_SCENARIOS)
def test_oracle_verlet_matches_reference_verlet_build(name):
    """The oracle's Verlet branch against the reference compiled with its VERLET option (CMakeLists.txt:13,90-93;
    src/sim.cu:1350-1354), same scenes, same driver (oracle/_ref/titan_ref_cuda_verlet)."""
    check_against_dump(name, os.path.join(GOLD, "ref_verlet_%s.bin" % name), verlet=True)


@pytest.mark.parametrize("name", scenes.RK2_SCENARIOS)
def test_oracle_rk2_matches_reference_rk2_build(name):
    """The oracle's RK2 branch against the reference compiled with its RK2 option (CMakeLists.txt:14,95-100; src/sim.cu:1336-1349,
    1778-1799) after oracle/patch_ref.py repaired the two call sites that do not compile as shipped (they pass 6 arguments to
    the 12-argument kernel); same scenes, same driver (oracle/_ref/titan_ref_cuda_rk2)."""
    path = os.path.join(GOLD, "ref_rk2_%s.bin" % name)
    if not os.path.exists(path):
        pytest.skip("no dump of the reference's RK2 build for this scene yet")
    check_against_dump(name, path, rk2=True)
