#!/usr/bin/env python3
"""Generate tests/golden/host_goldens.npz from the REFERENCE's own host-callable code.

Run in the build container (needs oracle/_ref/libtitan_ref_host.so, i.e. `make -C oracle ref`, which needs
/root/reference):   python tests/golden/make_host_goldens.py
The fixture holds seeded inputs and the reference's outputs for: Vec norm/dot/cross, every constraint
applyForce, the Lattice / Beam / Cube builders and Container::rotate.  tests/test_oracle_pin.py replays the
inputs through oracle/titan_oracle.c and demands identical bit patterns.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import refhost  # noqa: E402

LATTICES = [((0, 0, 10), (5, 5, 5), 5, 5, 5, False), ((0.1, -0.2, 3.3), (2, 3, 4), 3, 4, 6, True),
            ((1, 2, 3), (1, 1, 1), 1, 5, 2, False), ((0, 0, 0.55), (0.9, 0.9, 0.9), 10, 10, 10, False),
            ((0, 0, 5), (4, 4, 4), 20, 20, 20, False), ((0, 0, 0), (1, 1, 1), 1, 1, 1, False),
            ((0, 2, 3), (2, 0.5, 0.5), 6, 3, 3, True)]


def vec_cases(rng, n):
    v = rng.normal(size=(n, 3)) * np.exp(rng.normal(size=(n, 1)) * 3)
    v[0] = 0.0
    v[1] = (1e-200, 0, 0)
    v[2] = (3, 4, 0)
    return v


def main():
    rng = np.random.default_rng(12345)
    out = {}
    n = 400
    a, b = vec_cases(rng, n), vec_cases(rng, n)[::-1].copy()
    out["vec_a"], out["vec_b"] = a, b
    out["vec_norm"] = np.array([refhost.norm(x) for x in a])
    out["vec_dot"] = np.array([refhost.dot(x, y) for x, y in zip(a, b)])
    out["vec_cross"] = np.array([refhost.cross(x, y) for x, y in zip(a, b)])

    # contact planes: normal (unnormalised), offset, fk, fs, pos, vel, force
    m = 600
    normal = rng.normal(size=(m, 3))
    normal[:50] = (0, 0, 1)
    normal[50:80] = (0, 0, -2)
    offset = rng.normal(size=m) * 0.5
    fk = np.where(rng.random(m) < 0.3, 0.0, rng.random(m))
    fs = np.where(rng.random(m) < 0.3, 0.0, rng.random(m))
    pos = rng.normal(size=(m, 3))
    vel = rng.normal(size=(m, 3))
    force = rng.normal(size=(m, 3)) * 10
    vel[100:140] = 0.0                       # static-friction branch
    vel[140:160] = normal[140:160] * 0.3     # purely normal velocity -> |v_perp| ~ 0
    force[160:170] = 0.0
    force[170:180] = -0.0
    res = [refhost.contact_plane(normal[i], offset[i], fk[i], fs[i], pos[i], vel[i], force[i]) for i in range(m)]
    out.update(cp_normal=normal, cp_offset=offset, cp_fk=fk, cp_fs=fs, cp_pos=pos, cp_vel=vel, cp_force=force,
               cp_out=np.array([r[0] for r in res]), cp_nhat=np.array([r[1] for r in res]))
    out["lcp_out"] = np.array([refhost.local_contact_plane(normal[i], offset[i], pos[i], vel[i], force[i])
                               for i in range(m)])

    center = rng.normal(size=(m, 3))
    radius = np.abs(rng.normal(size=m)) * 2
    out.update(ball_center=center, ball_radius=radius,
               ball_out=np.array([refhost.ball(center[i], radius[i], pos[i], force[i]) for i in range(m)]))

    fr = rng.random(m)
    r1 = [refhost.constraint_plane(normal[i], fr[i], vel[i], force[i]) for i in range(m)]
    r2 = [refhost.direction(normal[i], fr[i], vel[i], force[i]) for i in range(m)]
    out.update(lc_friction=fr, cplane_vel=np.array([r[0] for r in r1]), cplane_force=np.array([r[1] for r in r1]),
               dir_vel=np.array([r[0] for r in r2]), dir_force=np.array([r[1] for r in r2]))

    for i, (c, d, nx, ny, nz, beam) in enumerate(LATTICES):
        L = refhost.lattice(c, d, nx, ny, nz, beam)
        if L["pos"].shape[0] <= 1000:
            for key in ("pos", "left", "right", "rest", "fixed"):
                out["lat%d_%s" % (i, key)] = L[key]
        else:  # big ones: keep checksums only (bit patterns folded to uint64 sums)
            out["lat%d_sum" % i] = np.array([L["pos"].view(np.uint64).sum(), L["rest"].view(np.uint64).sum(),
                                              (L["left"].astype(np.uint64) * np.arange(1, L["left"].size + 1, dtype=np.uint64)).sum(),
                                              (L["right"].astype(np.uint64) * np.arange(1, L["right"].size + 1, dtype=np.uint64)).sum()],
                                             dtype=np.uint64)
        assert set(L["k"]) <= {10000.0}
    out["lattices"] = np.array([[*c, *d, nx, ny, nz, int(beam)] for (c, d, nx, ny, nz, beam) in LATTICES], dtype=np.float64)

    cu = refhost.cube((0.5, -1, 3), 1.7)
    for key in ("pos", "left", "right", "rest"):
        out["cube_" + key] = cu[key]

    L = refhost.lattice((0, 0, 10), (5, 5, 5), 10, 10, 10)
    masses = 0.1 + 0.01 * (np.arange(1000) % 4)
    out["rot_mass"] = masses
    out["rot_a"] = refhost.rotate(L["pos"], masses, (0, 0, 1), -0.78)
    out["rot_b"] = refhost.rotate(L["pos"], masses, (1, 2, -0.5), 0.5)

    path = os.path.join(HERE, "host_goldens.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
