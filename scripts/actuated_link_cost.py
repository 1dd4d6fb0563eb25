import sys, numpy as np; sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
import bench, titan_b200 as tb
W = bench.workload_c2(60)
nm, ns = W["pos"].shape[0], W["left"].shape[0]
for with_link in (False, True):
    left, right, rest = W["left"], W["right"], W["rest"]
    if with_link:
        left = np.concatenate([left, [0]]).astype(np.int32); right = np.concatenate([right, [1]]).astype(np.int32); rest = np.concatenate([rest, [0.1]])
    eng = tb.Engine(0, tb.MATH_FAST)
    eng.set_topology(nm, left, right); eng.set_tile_hint(W["tile_hint"]); eng.set_mass_state(pos=W["pos"])
    ty = np.zeros(left.size, np.int32); lmax = np.zeros(left.size); lmin = np.zeros(left.size); rate = np.zeros(left.size)
    if with_link: ty[-1] = 4; lmax[-1] = 0.12; lmin[-1] = 0.08; rate[-1] = 0.01
    eng.set_spring_params(rest=rest, type=ty, l_max=lmax, l_min=lmin, rate=rate)
    eng.set_planes([tb.Plane((tb.capi.C.c_double * 3)(0, 0, 1), 0.0, 0.5, 0.6)]); eng.set_globals((0, 0, -9.81), 1e-4)
    eng.run(400, 0.0); eng.run_async(2000, 0.04); eng.sync()
    us = eng.last_run_ms() * 1e3 / 2000
    print("60^3 lattice%s: %.2f us/step, %.1f G spring-steps/s, kernel %d" % (" + one ACTUATED link" if with_link else "", us, left.size / us / 1e3, eng.info()["kernel"]), flush=True)
    eng.close()
