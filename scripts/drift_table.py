#!/usr/bin/env python3
"""Long-run divergence of TB_MATH_FAST (what bench.py runs) from the CPU oracle: max-norm of the difference after
100 / 1,000 / 10,000 steps.  C1 = README 5^3 lattice dropped from z = 10 onto the plane (the speed clamp of
src/sim.cu:1359-1360 holds it at 1 m/s: it reaches the plane after ~75,000 steps, so a second C1 starts 5 mm above the
plane and goes through the bounce); 20^3 lattice resting on a frictional plane (energy_unittest.cpp geometry scaled to
0.1 m spacing).  Test infrastructure (imports oracle); run on the GPU box:  python scripts/drift_table.py out.json"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle  # noqa: E402
import scenes  # noqa: E402
import titan_b200 as tb  # noqa: E402


def c1_bounce():
    sc = scenes.lattice_scene((0, 0, 2.505), (5, 5, 5), 5)
    sc.vel[:, 2] = -1.0
    sc.planes.append(oracle.make_plane((0, 0, 1), 0))
    return sc


CASES = {"c1_readme_free_fall": scenes.c1_readme, "c1_through_the_bounce": c1_bounce,
         "lattice20_resting_contact_friction": lambda: scenes.c2_lattice(20)}
rows = []
for name, make in CASES.items():
    for flags, label in ((tb.MATH_FAST, "fast"), (tb.MATH_STRICT, "strict")):
        sc = make()
        o = sc.copy()
        eng = tb.Engine(0, flags)
        scenes.upload(eng, sc)
        done = 0
        for n in (100, 1000, 10000):
            eng.run(n - done, o.T_sim)
            o.step(n - done)
            done = n
            st = eng.get_mass_state(fields=("pos", "vel"))
            row = {"case": name, "math": label, "steps": n, "kernel": eng.info()["kernel"],
                   "pos_max_abs": float(np.max(np.abs(st["pos"] - o.pos))), "vel_max_abs": float(np.max(np.abs(st["vel"] - o.vel))),
                   "pos_scale": float(np.max(np.abs(o.pos))), "min_z": float(o.pos[:, 2].min()), "max_speed": float(np.max(np.linalg.norm(o.vel, axis=1)))}
            rows.append(row)
            print(json.dumps(row), flush=True)
        eng.close()
if len(sys.argv) > 1:
    json.dump(rows, open(sys.argv[1], "w"), indent=1)
