"""A few steps of the occupancy-grid + magnet pass on a lattice, without graphs, for an ncu launch list."""
import sys
import numpy as np
sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import titan_b200 as tb
import scenes
n = int(sys.argv[1]) if len(sys.argv) > 1 else 60
sc = scenes.c2_lattice(n)
eng = tb.Engine(0, tb.MATH_FAST | tb.NO_GRAPH)
scenes.upload(eng, sc)
eng.set_mass_magnet(idx=np.array([0], dtype=np.int32), rad=np.array([0.01]), stiffness=np.array([1000.0]),
                    max_mag_force=np.array([0.001]), mag_scale=np.array([1.0]))
eng.run(6, 0.0)
eng.close()
