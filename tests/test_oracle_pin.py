"""Pins oracle/titan_oracle.c against the reference's OWN host-callable code, bit for bit.

Source of truth: tests/golden/host_goldens.npz, produced by tests/golden/make_host_goldens.py from
oracle/_ref/libtitan_ref_host.so (the reference's vec/mass/spring/object.cu compiled from /root/reference).
When that library is present (build container) a second pass calls it live on fresh random inputs.
"""
import ctypes as C
import os

import numpy as np
import pytest

import oracle
import refhost

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "host_goldens.npz"))
_dp = C.POINTER(C.c_double)


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float64).view(np.uint64)


def same_bits(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    nan = np.isnan(a) & np.isnan(b)
    return bool(np.all((bits(a) == bits(b)) | nan.reshape(bits(a).shape)))


def _d(a):
    return a.ctypes.data_as(_dp)


def o_contact(normal, offset, fk, fs, pos, vel, force):
    p = oracle.make_plane(normal, offset, fk, fs)
    f = np.array(force, dtype=np.float64)
    oracle.lib().to_contact_plane_apply(C.byref(p), _d(np.array(pos, dtype=np.float64)),
                                        _d(np.array(vel, dtype=np.float64)), _d(f))
    return f, np.array(p.normal[:])


def o_ball(center, radius, pos, force):
    b = oracle.make_ball(center, radius)
    f = np.array(force, dtype=np.float64)
    oracle.lib().to_ball_apply(C.byref(b), _d(np.array(pos, dtype=np.float64)), _d(f))
    return f


def o_local(kind, vec, num, vel, force):
    r = oracle.make_local(kind, vec, num)
    v, f = np.array(vel, dtype=np.float64), np.array(force, dtype=np.float64)
    n = np.array(r.vec[:])
    fn = oracle.lib().to_constraint_plane_apply if kind == oracle.CONSTRAINT_PLANE else oracle.lib().to_direction_apply
    fn(_d(n), num, _d(v), _d(f))
    return v, f


def test_vec_math():
    a, b = G["vec_a"], G["vec_b"]
    L = oracle.lib()
    assert same_bits([L.to_norm(_d(x.copy())) for x in a], G["vec_norm"])
    assert same_bits([L.to_dot(_d(x.copy()), _d(y.copy())) for x, y in zip(a, b)], G["vec_dot"])
    out = np.zeros((len(a), 3))
    for i, (x, y) in enumerate(zip(a, b)):
        L.to_cross(_d(x.copy()), _d(y.copy()), _d(out[i]))
    assert same_bits(out, G["vec_cross"])


def test_contact_plane_and_ball():
    n = len(G["cp_offset"])
    res = [o_contact(G["cp_normal"][i], G["cp_offset"][i], G["cp_fk"][i], G["cp_fs"][i], G["cp_pos"][i],
                     G["cp_vel"][i], G["cp_force"][i]) for i in range(n)]
    assert same_bits([r[1] for r in res], G["cp_nhat"])
    assert same_bits([r[0] for r in res], G["cp_out"])
    # local contact plane = frictionless plane built by CudaContactPlane(const Vec&, double)
    lres = [o_contact(G["cp_normal"][i], G["cp_offset"][i], 0.0, 0.0, G["cp_pos"][i], G["cp_vel"][i],
                      G["cp_force"][i])[0] for i in range(n)]
    assert same_bits(lres, G["lcp_out"])
    bres = [o_ball(G["ball_center"][i], G["ball_radius"][i], G["cp_pos"][i], G["cp_force"][i]) for i in range(n)]
    assert same_bits(bres, G["ball_out"])
    # the fixture must exercise every branch
    assert (G["cp_out"] != G["cp_force"]).any()


def test_local_constraints():
    n = len(G["lc_friction"])
    r1 = [o_local(oracle.CONSTRAINT_PLANE, G["cp_normal"][i], G["lc_friction"][i], G["cp_vel"][i], G["cp_force"][i])
          for i in range(n)]
    assert same_bits([r[0] for r in r1], G["cplane_vel"]) and same_bits([r[1] for r in r1], G["cplane_force"])
    r2 = [o_local(oracle.DIRECTION, G["cp_normal"][i], G["lc_friction"][i], G["cp_vel"][i], G["cp_force"][i])
          for i in range(n)]
    assert same_bits([r[0] for r in r2], G["dir_vel"]) and same_bits([r[1] for r in r2], G["dir_force"])


def test_builders_bit_exact():
    for i, row in enumerate(G["lattices"]):
        c, d, (nx, ny, nz, beam) = row[:3], row[3:6], row[6:].astype(int)
        L = oracle.lattice(c, d, nx, ny, nz, beam=bool(beam))
        if "lat%d_pos" % i in G:
            assert same_bits(L["pos"], G["lat%d_pos" % i]) and same_bits(L["rest"], G["lat%d_rest" % i])
            assert (L["left"] == G["lat%d_left" % i]).all() and (L["right"] == G["lat%d_right" % i]).all()
            assert (L["fixed"] == G["lat%d_fixed" % i]).all()
        else:
            w = np.arange(1, L["left"].size + 1, dtype=np.uint64)
            s = np.array([bits(L["pos"]).sum(), bits(L["rest"]).sum(), (L["left"].astype(np.uint64) * w).sum(),
                          (L["right"].astype(np.uint64) * w).sum()], dtype=np.uint64)
            assert (s == G["lat%d_sum" % i]).all()
    cu = oracle.cube((0.5, -1, 3), 1.7)
    assert same_bits(cu["pos"], G["cube_pos"]) and same_bits(cu["rest"], G["cube_rest"])
    assert (cu["left"] == G["cube_left"]).all() and (cu["right"] == G["cube_right"]).all()


def test_lattice_counts_formula():
    # 5^3 -> 125/1036, 10^3 -> 1000/10476, 20^3 -> 8000/93556
    for n, ns in ((5, 1036), (10, 10476), (20, 93556)):
        L = oracle.lattice((0, 0, 0), (1, 1, 1), n, n, n)
        assert L["pos"].shape[0] == n ** 3 and L["left"].shape[0] == ns


def test_rotate():
    L = oracle.lattice((0, 0, 10), (5, 5, 5), 10, 10, 10)
    assert same_bits(oracle.rotate(L["pos"], G["rot_mass"], (0, 0, 1), -0.78), G["rot_a"])
    assert same_bits(oracle.rotate(L["pos"], G["rot_mass"], (1, 2, -0.5), 0.5), G["rot_b"])


@pytest.mark.skipif(not refhost.available(), reason="oracle/_ref/libtitan_ref_host.so not built")
def test_live_against_reference_host_code():
    rng = np.random.default_rng(777)
    for _ in range(300):
        nrm, pos, vel, force = rng.normal(size=(4, 3))
        off, fk, fs = rng.normal(), rng.random() * (rng.random() < 0.7), rng.random() * (rng.random() < 0.7)
        if rng.random() < 0.2:
            vel[:] = 0
        assert same_bits(o_contact(nrm, off, fk, fs, pos, vel, force)[0],
                         refhost.contact_plane(nrm, off, fk, fs, pos, vel, force)[0])
        assert same_bits(o_ball(nrm, abs(off) * 3, pos, force), refhost.ball(nrm, abs(off) * 3, pos, force))
        for kind, fn in ((oracle.CONSTRAINT_PLANE, refhost.constraint_plane), (oracle.DIRECTION, refhost.direction)):
            ov, of = o_local(kind, nrm, fk, vel, force)
            rv, rf = fn(nrm, fk, vel, force)
            assert same_bits(ov, rv) and same_bits(of, rf)
    for (nx, ny, nz, beam) in ((2, 3, 4, True), (7, 1, 3, False), (4, 4, 1, False)):
        o, r = oracle.lattice((1, 2, 3), (3, 2, 1), nx, ny, nz, beam=beam), refhost.lattice((1, 2, 3), (3, 2, 1), nx, ny, nz, beam)
        assert same_bits(o["pos"], r["pos"]) and same_bits(o["rest"], r["rest"])
        assert (o["left"] == r["left"]).all() and (o["right"] == r["right"]).all() and (o["fixed"] == r["fixed"]).all()
