"""ctypes binding of the CPU oracle (oracle/titan_oracle.c).  TEST INFRASTRUCTURE ONLY.

Importable from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs and
from nowhere else: the product path (titan_b200/) never imports, links or executes anything in here.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libtitan_oracle.so")

PASSIVE_SOFT, PASSIVE_STIFF, ACTIVE_CONTRACT_THEN_EXPAND, ACTIVE_EXPAND_THEN_CONTRACT, ACTUATED_EXPAND, \
    ACTUATED_CONTRACT = range(6)
CONSTRAINT_PLANE, CONTACT_PLANE, BALL, DIRECTION = range(4)


class Plane(C.Structure):
    _fields_ = [("normal", C.c_double * 3), ("offset", C.c_double), ("friction_k", C.c_double),
                ("friction_s", C.c_double)]


class Ball(C.Structure):
    _fields_ = [("center", C.c_double * 3), ("radius", C.c_double)]


class Local(C.Structure):
    _fields_ = [("kind", C.c_int32), ("pad", C.c_int32), ("vec", C.c_double * 3), ("num", C.c_double)]


_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)
_bp = C.POINTER(C.c_uint8)
_lp = C.POINTER(C.c_int64)


class _Scene(C.Structure):
    _fields_ = [("nm", C.c_int64), ("ns", C.c_int64),
                ("pos", _dp), ("vel", _dp), ("acc", _dp), ("force", _dp), ("ext", _dp),
                ("m", _dp), ("T", _dp),
                ("rad", _dp), ("stiffness", _dp), ("max_mag", _dp), ("mag_scale", _dp),
                ("fixed", _bp), ("valid", _bp), ("drag", _ip),
                ("left", _ip), ("right", _ip), ("type", _ip),
                ("k", _dp), ("rest", _dp), ("damping", _dp), ("omega", _dp), ("l_max", _dp), ("l_min", _dp),
                ("rate", _dp),
                ("n_planes", C.c_int32), ("n_balls", C.c_int32),
                ("planes", C.POINTER(Plane)), ("balls", C.POINTER(Ball)),
                ("lc_ptr", _lp), ("lc", C.POINTER(Local)),
                ("g", C.c_double * 3), ("dt", C.c_double), ("T_sim", C.c_double), ("magnets", C.c_int32),
                ("verlet", C.c_int32), ("bk_pos", _dp), ("bk_vel", _dp)]


def build(force=False):
    """Compile oracle/libtitan_oracle.so with gcc (seconds)."""
    src = os.path.join(_HERE, "titan_oracle.c")
    if (not force and os.path.exists(_LIB_PATH)
            and os.path.getmtime(_LIB_PATH) >= max(os.path.getmtime(src),
                                                    os.path.getmtime(os.path.join(_HERE, "titan_oracle.h")))):
        return _LIB_PATH
    subprocess.check_call(["make", "-s", "-C", _HERE, "oracle"])
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        sp = C.POINTER(_Scene)
        L.to_spring_forces.argtypes = [sp, C.c_double]
        L.to_spring_forces_order.argtypes = [sp, C.c_double, _lp, C.c_int64]
        L.to_mass_update.argtypes = [sp]
        L.to_step.argtypes = [sp, C.c_int64]
        L.to_step_omp.argtypes = [sp, C.c_int64]
        L.to_run_until.argtypes = [sp, C.c_double]
        L.to_run_until.restype = C.c_int64
        L.to_omp_threads.restype = C.c_int
        L.to_norm.argtypes = [_dp]
        L.to_norm.restype = C.c_double
        L.to_dot.argtypes = [_dp, _dp]
        L.to_dot.restype = C.c_double
        L.to_cross.argtypes = [_dp, _dp, _dp]
        L.to_plane_init.argtypes = [C.POINTER(Plane), _dp, C.c_double, C.c_double, C.c_double]
        L.to_contact_plane_apply.argtypes = [C.POINTER(Plane), _dp, _dp, _dp]
        L.to_ball_apply.argtypes = [C.POINTER(Ball), _dp, _dp]
        L.to_constraint_plane_apply.argtypes = [_dp, C.c_double, _dp, _dp]
        L.to_direction_apply.argtypes = [_dp, C.c_double, _dp, _dp]
        L.to_lattice_counts.argtypes = [C.c_int, C.c_int, C.c_int, _lp, _lp]
        L.to_lattice_build.argtypes = [_dp, _dp, C.c_int, C.c_int, C.c_int, C.c_int, _dp, _ip, _ip, _dp, _bp]
        L.to_cube_build.argtypes = [_dp, C.c_double, _dp, _ip, _ip, _dp]
        L.to_rotate.argtypes = [_dp, _dp, C.c_int64, _dp, C.c_double]
        L.to_translate.argtypes = [_dp, C.c_int64, _dp]
        _lib = L
    return _lib


def _d(a):
    return a.ctypes.data_as(_dp)


def _v3(v):
    return (C.c_double * 3)(*[float(x) for x in v])


def make_plane(normal, offset, friction_k=0.0, friction_s=0.0):
    p = Plane()
    lib().to_plane_init(C.byref(p), _v3(normal), offset, friction_k, friction_s)
    return p


def make_ball(center, radius):
    b = Ball()
    b.center = _v3(center)
    b.radius = radius
    return b


def make_local(kind, vec, num):
    """Mass::addConstraint, src/mass.cu:104-122: normal/tangent are normalised, ball centers are not."""
    v = np.asarray(vec, dtype=np.float64)
    if kind != BALL:
        n = lib().to_norm(_d(np.ascontiguousarray(v)))
        v = v / n
    r = Local()
    r.kind = kind
    r.vec = _v3(v)
    r.num = num
    return r


def lattice(center, dims, nx, ny, nz, beam=False):
    """Returns dict(pos[nm,3], left[ns], right[ns], rest[ns], fixed[nm]) exactly as Lattice/Beam build them."""
    L = lib()
    nm, ns = C.c_int64(), C.c_int64()
    L.to_lattice_counts(nx, ny, nz, C.byref(nm), C.byref(ns))
    pos = np.zeros((nm.value, 3))
    left = np.zeros(ns.value, dtype=np.int32)
    right = np.zeros(ns.value, dtype=np.int32)
    rest = np.zeros(ns.value)
    fixed = np.zeros(nm.value, dtype=np.uint8)
    L.to_lattice_build(_v3(center), _v3(dims), nx, ny, nz, int(beam), _d(pos), left.ctypes.data_as(_ip),
                       right.ctypes.data_as(_ip), _d(rest), fixed.ctypes.data_as(_bp))
    return dict(pos=pos, left=left, right=right, rest=rest, fixed=fixed)


def cube(center, side):
    pos = np.zeros((8, 3))
    left = np.zeros(28, dtype=np.int32)
    right = np.zeros(28, dtype=np.int32)
    rest = np.zeros(28)
    lib().to_cube_build(_v3(center), side, _d(pos), left.ctypes.data_as(_ip), right.ctypes.data_as(_ip), _d(rest))
    return dict(pos=pos, left=left, right=right, rest=rest, fixed=np.zeros(8, dtype=np.uint8))


def rotate(pos, m, axis, angle):
    pos = np.ascontiguousarray(pos, dtype=np.float64).copy()
    m = np.ascontiguousarray(m, dtype=np.float64)
    lib().to_rotate(_d(pos), _d(m), pos.shape[0], _v3(axis), angle)
    return pos


class Scene:
    """A simulation state as plain numpy arrays + the oracle step over it.

    Field names follow CUDA_MASS / CUDA_SPRING (include/Titan/mass.h:89-126, spring.h:77-97).
    """

    def __init__(self, pos, left, right, rest, k=10000.0, m=0.1, dt=1e-4, g=(0.0, 0.0, -9.81)):
        nm = int(np.asarray(pos).shape[0])
        ns = int(np.asarray(left).shape[0])
        self.nm, self.ns = nm, ns
        f = np.float64
        self.pos = np.ascontiguousarray(pos, dtype=f).reshape(nm, 3).copy()
        self.vel = np.zeros((nm, 3))
        self.acc = np.zeros((nm, 3))
        self.force = np.zeros((nm, 3))
        self.ext = np.zeros((nm, 3))
        self.m = np.full(nm, m, dtype=f) if np.isscalar(m) else np.ascontiguousarray(m, dtype=f).copy()
        self.T = np.zeros(nm)
        self.rad = np.zeros(nm)
        self.stiffness = np.zeros(nm)
        self.max_mag = np.zeros(nm)
        self.mag_scale = np.zeros(nm)
        self.fixed = np.zeros(nm, dtype=np.uint8)
        self.valid = np.ones(nm, dtype=np.uint8)
        self.drag = np.zeros(nm, dtype=np.int32)
        self.left = np.ascontiguousarray(left, dtype=np.int32).copy()
        self.right = np.ascontiguousarray(right, dtype=np.int32).copy()
        self.type = np.zeros(ns, dtype=np.int32)
        self.k = np.full(ns, k, dtype=f) if np.isscalar(k) else np.ascontiguousarray(k, dtype=f).copy()
        self.rest = np.ascontiguousarray(rest, dtype=f).copy()
        self.damping = np.zeros(ns)
        self.omega = np.zeros(ns)
        self.l_max = np.zeros(ns)
        self.l_min = np.zeros(ns)
        self.rate = np.zeros(ns)
        self.planes = []
        self.balls = []
        self.local = [[] for _ in range(nm)] if nm <= 200000 else None  # per-mass lists of Local
        self.g = tuple(float(x) for x in g)
        self.dt = float(dt)
        self.T_sim = 0.0
        self.magnets = False
        self.verlet = False   # the reference's -DVERLET build
        self.rk2 = False      # ... its -DRK2 build (call sites repaired)

    def add_local(self, mass_index, kind, vec, num):
        self.local[mass_index].append(make_local(kind, vec, num))

    def copy(self):
        import copy
        return copy.deepcopy(self)

    # -- marshalling ------------------------------------------------------------------------------------
    def _c(self):
        s = _Scene()
        s.nm, s.ns = self.nm, self.ns
        for name in ("pos", "vel", "acc", "force", "ext", "m", "T", "rad", "stiffness", "max_mag", "mag_scale",
                     "k", "rest", "damping", "omega", "l_max", "l_min", "rate"):
            a = getattr(self, name)
            assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"], name
            setattr(s, name, _d(a))
        for name in ("fixed", "valid"):
            setattr(s, name, getattr(self, name).ctypes.data_as(_bp))
        for name in ("drag", "left", "right", "type"):
            a = getattr(self, name)
            assert a.dtype == np.int32, name
            setattr(s, name, a.ctypes.data_as(_ip))
        keep = []
        s.n_planes, s.n_balls = len(self.planes), len(self.balls)
        pl = (Plane * max(1, len(self.planes)))(*self.planes)
        bl = (Ball * max(1, len(self.balls)))(*self.balls)
        s.planes, s.balls = pl, bl
        keep += [pl, bl]
        if self.local is not None and any(self.local):
            ptr = np.zeros(self.nm + 1, dtype=np.int64)
            ptr[1:] = np.cumsum([len(x) for x in self.local])
            flat = [r for rows in self.local for r in rows]
            lc = (Local * len(flat))(*flat)
            s.lc_ptr, s.lc = ptr.ctypes.data_as(_lp), lc
            keep += [ptr, lc]
        s.g = _v3(self.g)
        s.dt, s.T_sim, s.magnets = self.dt, self.T_sim, int(self.magnets)
        s.verlet = 2 if getattr(self, "rk2", False) else int(self.verlet)
        if s.verlet == 2:
            bk = np.zeros((2, self.nm, 3))
            s.bk_pos, s.bk_vel = _d(bk[0]), _d(bk[1])
            keep.append(bk)
        return s, keep

    def _run(self, fn, *args):
        s, keep = self._c()
        r = fn(C.byref(s), *args)
        self.T_sim = s.T_sim
        del keep
        return r

    def step(self, nsteps=1):
        self._run(lib().to_step, int(nsteps))

    def step_omp(self, nsteps=1):
        self._run(lib().to_step_omp, int(nsteps))

    def run_until(self, breakpoint):
        return int(self._run(lib().to_run_until, float(breakpoint)))

    def spring_forces(self, t=None, order=None):
        t = self.T_sim if t is None else t
        if order is None:
            self._run(lib().to_spring_forces, float(t))
        else:
            o = np.ascontiguousarray(order, dtype=np.int64)
            self._run(lib().to_spring_forces_order, float(t), o.ctypes.data_as(_lp), o.shape[0])

    def mass_update(self):
        self._run(lib().to_mass_update)


def omp_threads():
    return int(lib().to_omp_threads())
