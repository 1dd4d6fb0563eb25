"""One magnetic mass in a C2-style lattice switches the occupancy-grid + magnet pass on for every mass: device time per
step with the pass in front of the tiled kernel, in front of the generic kernel, and without the pass."""
import sys, json
import numpy as np
sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import titan_b200 as tb
import scenes

n = int(sys.argv[1]) if len(sys.argv) > 1 else 60
sc = scenes.c2_lattice(n)
for label, flags, magnet in (("tiled+pass", tb.MATH_FAST, True), ("generic+pass", tb.MATH_FAST | tb.KERNEL_CSR, True),
                             ("tiled", tb.MATH_FAST, False), ("generic", tb.MATH_FAST | tb.KERNEL_CSR, False)):
    eng = tb.Engine(0, flags)
    scenes.upload(eng, sc)
    if magnet:
        eng.set_mass_magnet(idx=np.array([0], dtype=np.int32), rad=np.array([0.01]), stiffness=np.array([1000.0]),
                            max_mag_force=np.array([0.001]), mag_scale=np.array([1.0]))
    eng.run(100, 0.0)
    eng.run(400, 100 * sc.dt)
    ms = eng.last_run_ms()
    info = eng.info()
    print(json.dumps({"case": label, "lattice": n, "masses": sc.nm, "springs": sc.ns, "us_per_step": 1e3 * ms / 400,
                      "spring_steps_per_s": sc.ns * 400 / (ms * 1e-3), "kernel": info["kernel"]}))
    eng.close()
