"""The C++ drop-in (include/Titan/*.h + titan_b200/libTitan.so) exercised through the reference's own public API.

* tests/cpp/facade_tests.cpp: behaviour tests in the shape of the reference's test/physics/*_unittest.cpp.
* tests/cpp/bin/titan_facade_dump: tests/cpp/titan_api_driver.cu -- the SAME driver source that, linked against the
  reference's own CUDA engine, produced tests/golden/ref_cuda/*.bin -- compiled unchanged against this facade.
  Its dumps are compared record by record with those golden files (the reference's own outputs on a B200).

Tolerances: builders / initial scene bit-exact; step count of pause(t) (simulation time) bit-exact; per-mass T
bit-exact; pos, vel <= 1e-9 relative after up to 1,000 steps (the reference's atomicAdd summation order is
run-dependent and the facade uses TB_MATH_FAST); acc compared on the force scale; _rest <= 1e-15.
"""
import os
import subprocess

import numpy as np
import pytest

import scenes
from test_oracle_vs_reference_cuda import GOLD, read_dump, rel

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "tests", "cpp", "bin")
LIB = os.path.join(ROOT, "titan_b200", "libTitan.so")


def run(cmd, **kw):
    return subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=180, **kw)


def test_library_exports_the_reference_api():
    """Every public member function of titan::Simulation (reference include/Titan/sim.h:41-122) is exported."""
    out = run(["nm", "-D", "--defined-only", "-C", LIB]).stdout
    names = ["createMass", "createSpring", "deleteMass", "deleteSpring", "get", "set", "getAll", "setAll", "createPlane",
             "createBall", "clearConstraints", "createContainer", "deleteContainer", "createCube", "createLattice",
             "createRobotLink", "createBeam", "importFromSTL", "setAllSpringConstantValues", "setAllMassValues",
             "setTimeStep", "getTimeStep", "setGlobalAcceleration", "defaultRestLengths", "start", "stop", "pause",
             "resume", "reset", "setBreakpoint", "wait", "waitUntil", "waitForEvent", "time", "running",
             "printPositions", "getSpringByIndex", "getMassByIndex", "getContainerByIndex"]
    for n in names:
        assert "titan::Simulation::%s(" % n in out, n
    for n in ["titan::Lattice::Lattice(", "titan::Cube::Cube(", "titan::Beam::Beam(", "titan::RobotLink::RobotLink(",
              "titan::Container::rotate(", "titan::Container::translate(", "titan::Mass::addConstraint(",
              "titan::Mass::setDrag(", "titan::Mass::fix(", "titan::Spring::defaultLength(", "titan::Spring::setLeft("]:
        assert n in out, n
    # the facade reaches the GPU only through the C ABI
    und = run(["nm", "-D", "--undefined-only", LIB]).stdout
    assert "tb_run_async" in und and "tb_set_topology" in und and "cuda" not in und.lower()


@pytest.mark.parametrize("shape", [(5, 5, 5), (10, 10, 10), (38, 60, 60), (37, 30, 41), (100, 100, 100), (3, 200, 2)])
def test_facade_tile_hint_equals_the_array_side_builder(shape):
    """start() gives the engine the same brick decomposition (layered shapes, super-block numbering) as
    titan_b200.lattice.brick_hint, which bench.py and the partitioned path use."""
    import ctypes
    from titan_b200 import lattice
    L = ctypes.CDLL(LIB)
    out = np.zeros(shape[0] * shape[1] * shape[2], dtype=np.int32)
    n = L.titan_b200_brick_hint(*shape, out.ctypes.data_as(ctypes.c_void_p))
    ref = lattice.brick_hint(*shape)
    assert n == ref.max() + 1
    assert (out == ref).all()


def test_host_behaviour():
    r = run([os.path.join(BIN, "titan_facade_tests"), "host"])
    assert r.returncode == 0, r.stdout
    assert "0 failed" in r.stdout


@pytest.mark.parametrize("name", sorted(scenes.SCENARIOS))
def test_builders_bit_exact_against_reference_dump(name, tmp_path):
    """Initial scene built through the facade's createLattice/Cube/Beam/... == the reference's, bit for bit."""
    out = str(tmp_path / "init.bin")
    r = run([os.path.join(BIN, "titan_facade_dump"), "init", name, out])
    assert r.returncode == 0, r.stdout
    a, g = read_dump(out), read_dump(os.path.join(GOLD, "ref_%s.bin" % name))
    assert len(a) == (19 if name == "magnets" else 15)
    for k, v in a.items():
        assert v.tobytes() == g[k].tobytes(), k


@pytest.mark.gpu
def test_gpu_behaviour():
    r = run([os.path.join(BIN, "titan_facade_tests"), "gpu"])
    print(r.stdout[-3000:])
    assert r.returncode == 0, r.stdout[-3000:]


@pytest.mark.gpu
def test_simulation_over_several_gpus():
    """titan::Simulation(std::vector<int> devices): same API, same results; on a single-GPU box both engines share device 0."""
    import titan_b200 as tb
    second = 1 if tb.device_count() >= 2 else 0
    env = dict(os.environ, TB_GRID="16", TB_TPM="1")
    r = subprocess.run([os.path.join(BIN, "titan_facade_tests"), "multi", str(second)], env=env, stdout=subprocess.PIPE,
                       stderr=subprocess.STDOUT, text=True, timeout=600)
    print(r.stdout[-3000:])
    assert r.returncode == 0, r.stdout[-3000:]


CASES = ([(n, "") for n in sorted(scenes.SCENARIOS)] + [(n, "verlet") for n in scenes.VERLET_SCENARIOS] +
         [(n, "rk2") for n in scenes.RK2_SCENARIOS])


@pytest.mark.gpu
@pytest.mark.parametrize("name,verlet", CASES)
def test_facade_matches_reference_cuda_dump(name, verlet, tmp_path):
    """verlet / rk2: the driver compiled with -DVERLET / -DRK2 against this facade vs the reference built with that option."""
    out = str(tmp_path / "dump.bin")
    gold = os.path.join(GOLD, ("ref_%s_%s.bin" % (verlet, name)) if verlet else "ref_%s.bin" % name)
    if not os.path.exists(gold):
        pytest.skip("no dump of the reference's %s build for this scene yet" % verlet)
    r = run([os.path.join(BIN, "titan_facade_dump" + ("_" + verlet if verlet else "")), "dump", name, out])
    assert r.returncode == 0, r.stdout[-2000:]
    a, g = read_dump(out), read_dump(gold)
    assert set(a) == set(g)
    worst = {}
    for i, t_pause in enumerate(scenes.PAUSES[name]):
        p = "snap%d." % i
        assert a[p + "time"].tobytes() == g[p + "time"].tobytes()        # same number of steps: identical T += dt
        assert a[p + "T"].tobytes() == g[p + "T"].tobytes()              # per-mass clocks (fixed masses frozen)
        assert rel(a[p + "rest"], g[p + "rest"]) <= 1e-15
        if verlet == "rk2" and name == "contact_friction" and i == len(scenes.PAUSES[name]) - 1:
            assert rel(a[p + "pos"], g[p + "pos"]) <= 1e-4     # sliding-contact regime flips after ~100 steps (test_oracle_vs_reference_cuda.py)
            continue
        for f in ("pos", "vel", "acc"):
            worst[f] = max(worst.get(f, 0.0), rel(a[p + f], g[p + f]))
    print(name, worst)
    assert worst["pos"] <= 1e-9 and worst["vel"] <= 1e-9
    assert worst["acc"] <= 1e-9 * 1e3      # acc = force/m lives on the force scale k*L/m ~ 1e5
