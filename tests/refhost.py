"""ctypes binding of oracle/_ref/libtitan_ref_host.so -- the REFERENCE's own host-callable code
(built by oracle/build_ref.sh from /root/reference).  Test infrastructure; optional: `available()` is
False when the library was never built (then the committed fixtures in tests/golden/ stand in)."""
import ctypes as C
import os

import numpy as np

_PATH = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref",
                     "libtitan_ref_host.so")
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)
_bp = C.POINTER(C.c_ubyte)
_lib = None


def available():
    return os.path.exists(_PATH)


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(_PATH)
        L.ref_norm.restype = C.c_double
        L.ref_norm.argtypes = [_dp]
        L.ref_dot.restype = C.c_double
        L.ref_dot.argtypes = [_dp, _dp]
        L.ref_cross.argtypes = [_dp, _dp, _dp]
        L.ref_contact_plane.argtypes = [_dp, C.c_double, C.c_double, C.c_double, _dp, _dp, _dp, _dp]
        L.ref_local_contact_plane.argtypes = [_dp, C.c_double, _dp, _dp, _dp]
        L.ref_ball.argtypes = [_dp, C.c_double, _dp, _dp]
        L.ref_constraint_plane.argtypes = [_dp, C.c_double, _dp, _dp]
        L.ref_direction.argtypes = [_dp, C.c_double, _dp, _dp]
        L.ref_lattice_counts.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_longlong), C.POINTER(C.c_longlong)]
        L.ref_lattice.argtypes = [_dp, _dp, C.c_int, C.c_int, C.c_int, C.c_int, _dp, _ip, _ip, _dp, _dp, _bp]
        L.ref_cube.argtypes = [_dp, C.c_double, _dp, _ip, _ip, _dp]
        L.ref_rotate.argtypes = [_dp, _dp, C.c_longlong, _dp, C.c_double]
        _lib = L
    return _lib


def _d(a):
    return a.ctypes.data_as(_dp)


def _v(x):
    return np.ascontiguousarray(x, dtype=np.float64).copy()


def norm(v):
    return lib().ref_norm(_d(_v(v)))


def dot(a, b):
    return lib().ref_dot(_d(_v(a)), _d(_v(b)))


def cross(a, b):
    o = np.zeros(3)
    lib().ref_cross(_d(_v(a)), _d(_v(b)), _d(o))
    return o


def contact_plane(normal, offset, fk, fs, pos, vel, force):
    f, n = _v(force), np.zeros(3)
    lib().ref_contact_plane(_d(_v(normal)), offset, fk, fs, _d(_v(pos)), _d(_v(vel)), _d(f), _d(n))
    return f, n


def local_contact_plane(normal, offset, pos, vel, force):
    f = _v(force)
    lib().ref_local_contact_plane(_d(_v(normal)), offset, _d(_v(pos)), _d(_v(vel)), _d(f))
    return f


def ball(center, radius, pos, force):
    f = _v(force)
    lib().ref_ball(_d(_v(center)), radius, _d(_v(pos)), _d(f))
    return f


def constraint_plane(normal, friction, vel, force):
    v, f = _v(vel), _v(force)
    lib().ref_constraint_plane(_d(_v(normal)), friction, _d(v), _d(f))
    return v, f


def direction(tangent, friction, vel, force):
    v, f = _v(vel), _v(force)
    lib().ref_direction(_d(_v(tangent)), friction, _d(v), _d(f))
    return v, f


def lattice(center, dims, nx, ny, nz, beam=False):
    nm, ns = C.c_longlong(), C.c_longlong()
    lib().ref_lattice_counts(nx, ny, nz, C.byref(nm), C.byref(ns))
    nm, ns = nm.value, ns.value
    pos = np.zeros((nm, 3))
    left = np.zeros(ns, np.int32)
    right = np.zeros(ns, np.int32)
    rest, k = np.zeros(ns), np.zeros(ns)
    fixed = np.zeros(nm, np.uint8)
    lib().ref_lattice(_d(_v(center)), _d(_v(dims)), nx, ny, nz, int(beam), _d(pos), left.ctypes.data_as(_ip),
                      right.ctypes.data_as(_ip), _d(rest), _d(k), fixed.ctypes.data_as(_bp))
    return dict(pos=pos, left=left, right=right, rest=rest, k=k, fixed=fixed)


def cube(center, side):
    pos = np.zeros((8, 3))
    left = np.zeros(28, np.int32)
    right = np.zeros(28, np.int32)
    rest = np.zeros(28)
    lib().ref_cube(_d(_v(center)), side, _d(pos), left.ctypes.data_as(_ip), right.ctypes.data_as(_ip), _d(rest))
    return dict(pos=pos, left=left, right=right, rest=rest)


def rotate(pos, m, axis, angle):
    p = np.ascontiguousarray(pos, dtype=np.float64).copy()
    lib().ref_rotate(_d(p), _d(_v(m)), p.shape[0], _d(_v(axis)), angle)
    return p
