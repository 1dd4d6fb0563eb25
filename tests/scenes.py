"""Scene construction shared by the tests: an oracle.Scene (plain numpy arrays) and its upload to an Engine
through the C ABI.  Test infrastructure (imports oracle)."""
import numpy as np

import oracle
import titan_b200 as tb


def upload(eng, sc):
    """Upload every field of an oracle.Scene to a titan_b200 Engine."""
    eng.set_topology(sc.nm, sc.left, sc.right)
    eng.set_mass_state(pos=sc.pos, vel=sc.vel, m=sc.m, T=sc.T, extern_force=sc.ext, fixed=sc.fixed, drag=sc.drag,
                       valid=sc.valid)
    eng.set_mass_magnet(rad=sc.rad, stiffness=sc.stiffness, max_mag_force=sc.max_mag, mag_scale=sc.mag_scale)
    eng.set_spring_params(k=sc.k, rest=sc.rest, damping=sc.damping, type=sc.type, omega=sc.omega, l_max=sc.l_max,
                          l_min=sc.l_min, rate=sc.rate)
    eng.set_planes([tb.Plane((tb.capi.C.c_double * 3)(*p.normal), p.offset, p.friction_k, p.friction_s)
                    for p in sc.planes])
    eng.set_balls([tb.Ball((tb.capi.C.c_double * 3)(*b.center), b.radius) for b in sc.balls])
    if sc.local is not None and any(sc.local):
        eng.set_local_constraints([[tb.Local(r.kind, 0, (tb.capi.C.c_double * 3)(*r.vec), r.num) for r in rows]
                                   for rows in sc.local])
    eng.set_globals(sc.g, sc.dt)


def lattice_scene(center, dims, n, k=10000.0, m=0.1, dt=1e-4, g=(0, 0, -9.81), beam=False):
    nx, ny, nz = (n, n, n) if np.isscalar(n) else n
    L = oracle.lattice(center, dims, nx, ny, nz, beam=beam)
    sc = oracle.Scene(L["pos"], L["left"], L["right"], L["rest"], k=k, m=m, dt=dt, g=g)
    sc.fixed[:] = L["fixed"]
    return sc


def c1_readme():
    """README.md:13-18: 5x5x5 lattice at (0,0,10) over the plane z=0, all defaults."""
    sc = lattice_scene((0, 0, 10), (5, 5, 5), 5)
    sc.planes.append(oracle.make_plane((0, 0, 1), 0))
    return sc


def c2_lattice(n=100):
    """n^3 lattice, spacing 0.1, bottom layer 1 mm inside a frictional plane."""
    side = 0.1 * (n - 1)
    sc = lattice_scene((0, 0, side / 2 - 0.001), (side, side, side), n)
    sc.planes.append(oracle.make_plane((0, 0, 1), 0, 0.5, 0.6))
    return sc


def contact_friction():
    sc = lattice_scene((0, 0, 0.499), (1, 1, 1), 6)
    n = np.arange(sc.nm)
    sc.vel[:, 0] = 0.3 + 0.001 * (n % 7)
    sc.vel[:, 1] = -0.2 + 0.002 * (n % 5)
    sc.vel[:, 2] = -0.1
    s = np.arange(sc.ns)
    sc.damping[:] = np.where(s % 3 == 0, 0.05 + 0.01 * (s % 4), 0.0)
    sc.planes.append(oracle.make_plane((0, 0, 2), 0, 0.5, 0.6))
    sc.planes.append(oracle.make_plane((1, 0, 0.2), -0.6, 0.0, 0.3))
    return sc


def actuated():
    sc = lattice_scene((0, 0, 1.0), (0.9, 0.9, 0.9), 4)
    s = np.arange(sc.ns)
    sc.type[:] = s % 6
    sc.omega[:] = 4 * 3.14159265358979323846 * (1 + (s % 3))
    sc.damping[:] = np.where(s % 2 == 1, 0.1, 0.0)
    sc.l_max[:] = sc.rest * 1.001
    sc.l_min[:] = sc.rest * 0.999
    sc.rate[:] = 0.05 + 0.01 * (s % 5)
    passive = sc.type < 4   # the dump program zeroes actuator fields of non-actuated springs
    sc.l_max[passive] = 0; sc.l_min[passive] = 0; sc.rate[passive] = 0
    sc.balls.append(oracle.make_ball((0, 0, -0.2), 0.8))
    sc.planes.append(oracle.make_plane((0, 0, 1), 0, 0.5, 0.6))
    return sc


def local_constraints():
    sc = lattice_scene((0, 0, 2), (1, 1, 1), 4, dt=0.0002, g=(0.1, -0.2, -9.8))
    for n in range(sc.nm):
        sc.vel[n] = (0.05 * ((n % 3) - 1), 0.04 * ((n % 5) - 2), 0.03 * ((n % 7) - 3))
        if n % 11 == 0:
            sc.fixed[n] = 1
        if n % 5 == 1:
            sc.drag[n] = int(2.0 + (n % 3))
        if n % 7 == 2:
            sc.add_local(n, oracle.CONSTRAINT_PLANE, (0.2, 0.1, 1), 0.3)
        if n % 7 == 3:
            sc.add_local(n, oracle.DIRECTION, (1, 0.5, 0.25), 0.2)
        if n % 7 == 4:
            sc.add_local(n, oracle.CONTACT_PLANE, (0, 0.1, 1), 1.7)
        if n % 7 == 5:
            sc.add_local(n, oracle.BALL, (0.1, 0.1, 1.9), 0.6)
        if n % 13 == 6:
            sc.add_local(n, oracle.CONTACT_PLANE, (0, 0, 1), 1.6)
            sc.add_local(n, oracle.DIRECTION, (0, 1, 1), 0.0)
        if n % 4 == 0:
            sc.ext[n] = (0.5, -0.25, 1.0 + 0.1 * n)
        sc.m[n] = 0.1 + 0.01 * (n % 4)
    return sc


def cube_beam():
    cu = oracle.cube((0, 0, 3), 1.0)
    bm = oracle.lattice((0, 2, 3), (2, 0.5, 0.5), 6, 3, 3, beam=True)
    pos = np.vstack([cu["pos"], bm["pos"]])
    left = np.concatenate([cu["left"], bm["left"] + 8])
    right = np.concatenate([cu["right"], bm["right"] + 8])
    rest = np.concatenate([cu["rest"], bm["rest"]])
    sc = oracle.Scene(pos, left, right, rest)
    sc.fixed[8:] = bm["fixed"]
    sc.k[28:] = 5e4
    sc.planes.append(oracle.make_plane((0, 0, 1), 2.45, 0.2, 0.3))
    return sc


def magnets():
    """Magnetic connector masses (RobotLink, src/object.cu:368-384) + the occupancy-grid pass; same scene as
    tests/cpp/titan_api_driver.cu "magnets".  Per group: link (fixed end, free end), a fixed magnet close to the free end, a
    free plain mass beyond the cut-off, a fixed susceptible mass."""
    pos, fixed, m, rad, stiff, maxmag, scale = [], [], [], [], [], [], []
    left, right, k, rest = [], [], [], []
    for g in range(5):
        x0, ln = -0.988 + 0.5 * g, 0.09
        b = len(pos)
        force = 0.004 * (g + 1)
        for p_, fx in (((x0, 0.01, 0.5), 1), ((x0 + 0.03, 0.02, 0.5 + ln), 0)):          # RobotLink ml, mr
            pos.append(p_); fixed.append(fx); m.append(0.05); rad.append(0.015); stiff.append(5000.0)
            maxmag.append(force); scale.append(1.0)
        left.append(b); right.append(b + 1); k.append(2000.0); rest.append(0.08)
        pos.append((x0 + 0.035 + 0.002 * g, 0.015, 0.5 + ln + 0.02 + 0.004 * g))           # fixed magnet
        fixed.append(1); m.append(0.1); rad.append(0.015); stiff.append(4000.0); maxmag.append(0.003); scale.append(0.8)
        pos.append((x0 + 0.03, 0.02, 0.5 + ln + 0.35))                                     # free plain mass
        fixed.append(0); m.append(0.1); rad.append(0.0); stiff.append(1000.0); maxmag.append(0.0); scale.append(0.0)
        pos.append((x0 + 0.03, 0.02 - 0.03, 0.5 + ln - 0.03))                              # fixed susceptible mass
        fixed.append(1); m.append(0.1); rad.append(0.0); stiff.append(1000.0); maxmag.append(0.0); scale.append(0.5)
    sc = oracle.Scene(np.array(pos), np.array(left), np.array(right), np.array(rest), k=np.array(k), m=np.array(m))
    sc.fixed[:] = fixed
    sc.rad[:] = rad; sc.stiffness[:] = stiff; sc.max_mag[:] = maxmag; sc.mag_scale[:] = scale
    sc.magnets = True
    return sc


SCENARIOS = {"c1_readme": c1_readme, "contact_friction": contact_friction, "actuated": actuated,
             "local_constraints": local_constraints, "cube_beam": cube_beam, "magnets": magnets}
# pause() times used by tests/cpp/titan_api_driver.cu for each scenario
PAUSES = {"c1_readme": [0.0, 0.001, 0.01, 0.1], "contact_friction": [0.0, 0.001, 0.01, 0.05],
          "actuated": [0.0, 0.001, 0.01, 0.05], "local_constraints": [0.0, 0.002, 0.01],
          "cube_beam": [0.0, 0.001, 0.02], "magnets": [0.0, 0.001, 0.005]}
# scenarios that were also run through the reference built with -DVERLET (tests/golden/ref_cuda/ref_verlet_<name>.bin)
VERLET_SCENARIOS = ("c1_readme", "contact_friction", "actuated", "local_constraints")
# ... and through the reference built with -DRK2, call sites repaired (tests/golden/ref_cuda/ref_rk2_<name>.bin)
RK2_SCENARIOS = ("c1_readme", "contact_friction", "actuated", "local_constraints")
