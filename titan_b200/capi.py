"""ctypes binding of the C ABI in include/titan_b200.h (libtitan_b200.so).

Thin by design: every method maps onto one tb_* entry point with numpy arrays as the plain host buffers.
There is no CPU path: if the library is missing, or no CUDA device is present, creating an Engine raises.
"""
import ctypes as C
import os

import numpy as np

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("TITAN_B200_LIB") or os.path.join(_PKG, "libtitan_b200.so")   # override: A/B builds on one GPU box

MATH_FAST, MATH_STRICT, NO_GRAPH, KERNEL_CSR, INTEGRATOR_VERLET, NO_RESIDENT, INTEGRATOR_RK2 = 0, 1, 2, 4, 8, 16, 32
PASSIVE_SOFT, PASSIVE_STIFF, ACTIVE_CONTRACT_THEN_EXPAND, ACTIVE_EXPAND_THEN_CONTRACT, ACTUATED_EXPAND, \
    ACTUATED_CONTRACT = range(6)
CONSTRAINT_PLANE, CONTACT_PLANE, BALL, DIRECTION = range(4)

EXPORTS = ["tb_abi_version", "tb_device_count", "tb_create", "tb_destroy", "tb_last_error", "tb_set_topology", "tb_set_tile_hint",
           "tb_set_mass_state", "tb_set_mass_magnet", "tb_get_mass_state", "tb_set_spring_params",
           "tb_get_spring_rest", "tb_remove_springs", "tb_compact", "tb_add_masses", "tb_add_springs", "tb_set_planes", "tb_set_balls", "tb_set_local_constraints", "tb_set_globals",
           "tb_run_async", "tb_sync", "tb_steps_done", "tb_wait_steps", "tb_last_run_ms", "tb_get_info",
           "tb_export_positions_f32_dev", "tb_export_indices_dev", "tb_read_device", "tb_alloc_host",
           "tb_static_scratch_max_masses", "tb_get_mass_state_static", "tb_halo_set_ghosts", "tb_halo_prepare",
           "tb_halo_state_offsets", "tb_halo_connect", "tb_halo_set_exports", "tb_halo_push_now"]
EXPORTS += ["tb_create_partitioned", "tb_batch_create", "tb_group_destroy", "tb_group_last_error", "tb_group_size", "tb_group_engine",
            "tb_group_set_topology", "tb_group_set_tile_hint", "tb_group_set_mass_state", "tb_group_set_mass_magnet",
            "tb_group_get_mass_state", "tb_group_set_spring_params", "tb_group_get_spring_rest", "tb_group_remove_springs",
            "tb_group_set_planes", "tb_group_set_balls", "tb_group_set_globals", "tb_group_run_async", "tb_group_sync",
            "tb_group_wait_steps", "tb_group_last_run_ms", "tb_group_steps_done"]
MAX_PEERS = 8
HALO_HANDLE_BYTES = 320


class Plane(C.Structure):
    _fields_ = [("normal", C.c_double * 3), ("offset", C.c_double), ("friction_k", C.c_double),
                ("friction_s", C.c_double)]


class Ball(C.Structure):
    _fields_ = [("center", C.c_double * 3), ("radius", C.c_double)]


class Local(C.Structure):
    _fields_ = [("kind", C.c_int32), ("pad", C.c_int32), ("vec", C.c_double * 3), ("num", C.c_double)]


class Info(C.Structure):
    _fields_ = [("nm", C.c_int64), ("ns", C.c_int64), ("kernel", C.c_int32), ("n_tiles", C.c_int32),
                ("launches", C.c_int64), ("device_bytes", C.c_int64), ("algorithmic_bytes_per_step", C.c_double),
                ("sm_count", C.c_int32), ("graph_steps", C.c_int32), ("threads_per_mass", C.c_int32),
                ("grid", C.c_int32), ("smem_bytes", C.c_int32), ("ell_unique_tiles", C.c_int32),
                ("tiled_bytes_per_step", C.c_int64), ("ns_removed", C.c_int64),
                ("dict_records", C.c_int32), ("dict_split_k", C.c_int32),
                ("layout_uploads", C.c_int32), ("layout_patches", C.c_int32)]


class TitanError(RuntimeError):
    def __init__(self, status, message):
        super().__init__("titan_b200 error %d: %s" % (status, message))
        self.status = status


_lib = None
_vp = C.c_void_p


def lib():
    """Load libtitan_b200.so.  Raises (never falls back) when the CUDA extension has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError("titan_b200: %s is missing -- build it with `python -c 'import __graft_entry__ as g; "
                              "g.build()'` (there is no CPU fallback)" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        L.tb_last_error.restype = C.c_char_p
        L.tb_last_error.argtypes = [_vp]
        L.tb_create.argtypes = [C.c_int, C.c_uint32, C.POINTER(_vp)]
        L.tb_destroy.argtypes = [_vp]
        L.tb_destroy.restype = None
        L.tb_set_topology.argtypes = [_vp, C.c_int64, C.c_int64, _vp, _vp]
        L.tb_set_tile_hint.argtypes = [_vp, _vp]
        L.tb_set_mass_state.argtypes = [_vp, C.c_int64, C.c_int64] + [_vp] * 9
        L.tb_set_mass_magnet.argtypes = [_vp, C.c_int64, C.c_int64] + [_vp] * 5
        L.tb_get_mass_state.argtypes = [_vp, C.c_int64, C.c_int64] + [_vp] * 6
        L.tb_set_spring_params.argtypes = [_vp, C.c_int64, C.c_int64] + [_vp] * 9
        L.tb_get_spring_rest.argtypes = [_vp, C.c_int64, C.c_int64, _vp, _vp]
        L.tb_remove_springs.argtypes = [_vp, C.c_int64, _vp]
        L.tb_add_masses.argtypes = [_vp, C.c_int64]
        L.tb_compact.argtypes = [_vp]
        L.tb_add_springs.argtypes = [_vp, C.c_int64, _vp, _vp]
        L.tb_set_planes.argtypes = [_vp, C.c_int32, _vp]
        L.tb_set_balls.argtypes = [_vp, C.c_int32, _vp]
        L.tb_set_local_constraints.argtypes = [_vp, _vp, _vp]
        L.tb_set_globals.argtypes = [_vp, _vp, C.c_double]
        L.tb_run_async.argtypes = [_vp, C.c_int64, C.c_double]
        L.tb_sync.argtypes = [_vp]
        L.tb_steps_done.argtypes = [_vp, C.POINTER(C.c_int64)]
        L.tb_wait_steps.argtypes = [_vp, C.c_int64]
        L.tb_last_run_ms.argtypes = [_vp, C.POINTER(C.c_double)]
        L.tb_get_info.argtypes = [_vp, C.POINTER(Info)]
        L.tb_export_positions_f32_dev.argtypes = [_vp, C.POINTER(_vp)]
        L.tb_export_indices_dev.argtypes = [_vp, C.POINTER(_vp)]
        L.tb_read_device.argtypes = [_vp, _vp, _vp, C.c_size_t]
        L.tb_alloc_host.argtypes = [_vp, C.c_size_t, C.POINTER(_vp)]
        L.tb_static_scratch_max_masses.restype = C.c_int64
        L.tb_static_scratch_max_masses.argtypes = []
        L.tb_get_mass_state_static.argtypes = [_vp, C.c_int64, C.c_int64, _vp, _vp, _vp]
        L.tb_halo_set_ghosts.argtypes = [_vp, C.c_int64, _vp]
        L.tb_halo_prepare.argtypes = [_vp, _vp]
        L.tb_halo_state_offsets.argtypes = [_vp, C.c_int64, _vp, _vp]
        L.tb_halo_connect.argtypes = [_vp, C.c_int32, _vp, _vp]
        L.tb_halo_set_exports.argtypes = [_vp, C.c_int64, _vp, _vp, _vp]
        L.tb_halo_push_now.argtypes = [_vp]
        L.tb_create_partitioned.argtypes = [C.c_int32, _vp, C.c_uint32, C.POINTER(_vp)]
        L.tb_batch_create.argtypes = [C.c_int32, _vp, C.c_uint32, C.c_int64, _vp, C.POINTER(_vp)]
        L.tb_group_destroy.argtypes = [_vp]
        L.tb_group_destroy.restype = None
        L.tb_group_last_error.argtypes = [_vp]
        L.tb_group_last_error.restype = C.c_char_p
        L.tb_group_size.argtypes = [_vp]
        L.tb_group_engine.argtypes = [_vp, C.c_int32]
        L.tb_group_engine.restype = _vp
        L.tb_group_set_topology.argtypes = [_vp, C.c_int64, C.c_int64, _vp, _vp, _vp]
        L.tb_group_set_tile_hint.argtypes = [_vp, _vp]
        L.tb_group_set_mass_state.argtypes = [_vp, C.c_int64, C.c_int64] + [_vp] * 9
        L.tb_group_set_mass_magnet.argtypes = [_vp, C.c_int64, C.c_int64] + [_vp] * 5
        L.tb_group_get_mass_state.argtypes = [_vp, C.c_int64, C.c_int64] + [_vp] * 6
        L.tb_group_set_spring_params.argtypes = [_vp, C.c_int64, C.c_int64] + [_vp] * 9
        L.tb_group_get_spring_rest.argtypes = [_vp, C.c_int64, C.c_int64, _vp, _vp]
        L.tb_group_remove_springs.argtypes = [_vp, C.c_int64, _vp]
        L.tb_group_set_planes.argtypes = [_vp, C.c_int32, _vp]
        L.tb_group_set_balls.argtypes = [_vp, C.c_int32, _vp]
        L.tb_group_set_globals.argtypes = [_vp, _vp, C.c_double]
        L.tb_group_run_async.argtypes = [_vp, C.c_int64, C.c_double]
        L.tb_group_sync.argtypes = [_vp]
        L.tb_group_wait_steps.argtypes = [_vp, C.c_int64]
        L.tb_group_last_run_ms.argtypes = [_vp, C.POINTER(C.c_double)]
        L.tb_group_steps_done.argtypes = [_vp, C.POINTER(C.c_int64)]
        _lib = L
    return _lib


def device_count():
    return int(lib().tb_device_count())


def _arr(a, dtype, shape=None):
    if a is None:
        return None, None
    x = np.ascontiguousarray(a, dtype=dtype)
    if shape is not None:
        x = x.reshape(shape)
    return x, x.ctypes.data_as(_vp)


class Engine:
    """One tb_engine: one CUDA device, one stream."""

    def __init__(self, device=0, flags=MATH_FAST):
        self._L = lib()
        h = _vp()
        rc = self._L.tb_create(int(device), int(flags), C.byref(h))
        if rc != 0:
            raise TitanError(rc, self._L.tb_last_error(None).decode())
        self._h = h
        self.nm = self.ns = 0

    def close(self):
        if getattr(self, "_h", None):
            self._L.tb_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != 0:
            raise TitanError(rc, self._L.tb_last_error(self._h).decode())

    # ---- uploads ---------------------------------------------------------------------------------------
    def set_topology(self, nm, left, right):
        l, lp = _arr(left, np.int32)
        r, rp = _arr(right, np.int32)
        ns = 0 if l is None else l.shape[0]
        self._ck(self._L.tb_set_topology(self._h, int(nm), int(ns), lp, rp))
        self.nm, self.ns = int(nm), int(ns)

    def set_tile_hint(self, tile_of_mass):
        if tile_of_mass is None:
            self._ck(self._L.tb_set_tile_hint(self._h, None))
            return
        a, ap = _arr(tile_of_mass, np.int32, (self.nm,))
        self._ck(self._L.tb_set_tile_hint(self._h, ap))

    def _sel(self, n_total, first, count, idx):
        if idx is not None:
            i, ip = _arr(idx, np.int32)
            return 0, i.shape[0], i, ip
        if count is None:
            count = n_total - first
        return first, count, None, None

    def set_mass_state(self, first=0, count=None, idx=None, pos=None, vel=None, m=None, T=None, extern_force=None,
                       fixed=None, drag=None, valid=None):
        first, count, keep, ip = self._sel(self.nm, first, count, idx)
        a = [_arr(pos, np.float64, (count, 3)), _arr(vel, np.float64, (count, 3)), _arr(m, np.float64, (count,)),
             _arr(T, np.float64, (count,)), _arr(extern_force, np.float64, (count, 3)),
             _arr(fixed, np.uint8, (count,)), _arr(drag, np.int32, (count,)), _arr(valid, np.uint8, (count,))]
        self._ck(self._L.tb_set_mass_state(self._h, first, count, ip, *[p for _, p in a]))

    def set_mass_magnet(self, first=0, count=None, idx=None, rad=None, stiffness=None, max_mag_force=None,
                        mag_scale=None):
        first, count, keep, ip = self._sel(self.nm, first, count, idx)
        a = [_arr(x, np.float64, (count,)) for x in (rad, stiffness, max_mag_force, mag_scale)]
        self._ck(self._L.tb_set_mass_magnet(self._h, first, count, ip, *[p for _, p in a]))

    def set_spring_params(self, first=0, count=None, idx=None, k=None, rest=None, damping=None, type=None,
                          omega=None, l_max=None, l_min=None, rate=None):
        first, count, keep, ip = self._sel(self.ns, first, count, idx)
        d = lambda x: _arr(x, np.float64, (count,))  # noqa: E731
        a = [d(k), d(rest), d(damping), _arr(type, np.int32, (count,)), d(omega), d(l_max), d(l_min), d(rate)]
        self._ck(self._L.tb_set_spring_params(self._h, first, count, ip, *[p for _, p in a]))

    def set_planes(self, planes):
        arr = (Plane * max(1, len(planes)))(*planes)
        self._ck(self._L.tb_set_planes(self._h, len(planes), C.cast(arr, _vp)))

    def set_balls(self, balls):
        arr = (Ball * max(1, len(balls)))(*balls)
        self._ck(self._L.tb_set_balls(self._h, len(balls), C.cast(arr, _vp)))

    def set_local_constraints(self, per_mass):
        """per_mass: list (len nm) of lists of Local records, or None to clear."""
        if not per_mass or not any(per_mass):
            self._ck(self._L.tb_set_local_constraints(self._h, None, None))
            return
        ptr = np.zeros(self.nm + 1, dtype=np.int64)
        ptr[1:] = np.cumsum([len(x) for x in per_mass])
        flat = [r for rows in per_mass for r in rows]
        recs = (Local * len(flat))(*flat)
        self._ck(self._L.tb_set_local_constraints(self._h, ptr.ctypes.data_as(_vp), C.cast(recs, _vp)))

    def set_globals(self, g=None, dt=1e-4):
        gp = None if g is None else (C.c_double * 3)(*[float(x) for x in g])
        self._ck(self._L.tb_set_globals(self._h, None if gp is None else C.cast(gp, _vp), float(dt)))

    # ---- stepping --------------------------------------------------------------------------------------
    def run_async(self, nsteps, t0=0.0):
        self._ck(self._L.tb_run_async(self._h, int(nsteps), float(t0)))

    def sync(self):
        self._ck(self._L.tb_sync(self._h))

    def run(self, nsteps, t0=0.0):
        self.run_async(nsteps, t0)
        self.sync()

    def steps_done(self):
        n = C.c_int64()
        self._ck(self._L.tb_steps_done(self._h, C.byref(n)))
        return n.value

    def last_run_ms(self):
        ms = C.c_double()
        self._ck(self._L.tb_last_run_ms(self._h, C.byref(ms)))
        return ms.value

    def info(self):
        i = Info()
        self._ck(self._L.tb_get_info(self._h, C.byref(i)))
        return {f: getattr(i, f) for f, _ in Info._fields_}

    # ---- downloads -------------------------------------------------------------------------------------
    def get_mass_state(self, first=0, count=None, idx=None, fields=("pos", "vel", "acc", "T", "extern_force"),
                       out=None):
        """Download the selected fields; `out` may hold preallocated (e.g. pinned) float64 arrays to fill."""
        first, count, keep, ip = self._sel(self.nm, first, count, idx)
        out = dict(out or {})
        ptrs = []
        for name in ("pos", "vel", "acc", "T", "extern_force"):
            if name in fields:
                shape = (count,) if name == "T" else (count, 3)
                if name not in out:
                    out[name] = np.zeros(shape)
                a = out[name]
                assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"] and a.size == int(np.prod(shape)), name
                ptrs.append(out[name].ctypes.data_as(_vp))
            else:
                ptrs.append(None)
        self._ck(self._L.tb_get_mass_state(self._h, first, count, ip, *ptrs))
        return out

    def get_spring_rest(self, first=0, count=None, idx=None):
        first, count, keep, ip = self._sel(self.ns, first, count, idx)
        out = np.zeros(count)
        self._ck(self._L.tb_get_spring_rest(self._h, first, count, ip, out.ctypes.data_as(_vp)))
        return out

    def remove_springs(self, idx):
        """Turn springs into tombstones (no force from now on; indices of the others unchanged)."""
        a, ap = _arr(idx, np.int32)
        self._ck(self._L.tb_remove_springs(self._h, a.shape[0], ap))

    def compact(self):
        self._ck(self._L.tb_compact(self._h))

    def add_masses(self, n):
        """Grow by n masses (defaults); returns the index of the first new one."""
        first = self.nm
        self._ck(self._L.tb_add_masses(self._h, int(n)))
        self.nm += int(n)
        return first

    def add_springs(self, left, right):
        """Grow by len(left) springs (default parameters); returns the index of the first new one."""
        l, lp = _arr(left, np.int32)
        r, rp = _arr(right, np.int32)
        first = self.ns
        self._ck(self._L.tb_add_springs(self._h, l.shape[0], lp, rp))
        self.ns += int(l.shape[0])
        return first

    def alloc_pinned(self, shape, dtype=np.float64):
        """A numpy array over page-locked host memory (tb_alloc_host); freed with the engine."""
        n = int(np.prod(shape)) * np.dtype(dtype).itemsize
        p = _vp()
        self._ck(self._L.tb_alloc_host(self._h, C.c_size_t(n), C.byref(p)))
        buf = (C.c_char * n).from_address(p.value)
        return np.frombuffer(buf, dtype=dtype).reshape(shape)

    # ---- partitioned simulation (halo exchange over peer memory) ----------------------------------------
    def halo_set_ghosts(self, mass_idx):
        a, ap = _arr(mass_idx, np.int32)
        self._ck(self._L.tb_halo_set_ghosts(self._h, a.shape[0], ap))

    def halo_prepare(self):
        """Builds the device layout; returns this engine's handle (bytes) for its peers."""
        buf = C.create_string_buffer(HALO_HANDLE_BYTES)
        self._ck(self._L.tb_halo_prepare(self._h, C.cast(buf, _vp)))
        return buf.raw

    def halo_state_offsets(self, mass_idx):
        a, ap = _arr(mass_idx, np.int32)
        out = np.zeros(a.shape[0], dtype=np.int64)
        self._ck(self._L.tb_halo_state_offsets(self._h, a.shape[0], ap, out.ctypes.data_as(_vp)))
        return out

    def halo_connect(self, handles, flag_slot_at_peer):
        blob = C.create_string_buffer(b"".join(handles), HALO_HANDLE_BYTES * max(1, len(handles)))
        slots, sp = _arr(np.asarray(flag_slot_at_peer, dtype=np.int32).reshape(-1), np.int32)
        self._ck(self._L.tb_halo_connect(self._h, len(handles), C.cast(blob, _vp), sp))

    def halo_set_exports(self, mass_idx, peer_idx, remote_offset):
        a, ap = _arr(mass_idx, np.int32)
        b, bp = _arr(peer_idx, np.int32)
        c, cp = _arr(remote_offset, np.int64)
        self._ck(self._L.tb_halo_set_exports(self._h, a.shape[0], ap, bp, cp))

    def halo_push_now(self):
        self._ck(self._L.tb_halo_push_now(self._h))

    def export_positions_f32_dev(self):
        p = _vp()
        self._ck(self._L.tb_export_positions_f32_dev(self._h, C.byref(p)))
        return p.value

    def export_indices_dev(self):
        p = _vp()
        self._ck(self._L.tb_export_indices_dev(self._h, C.byref(p)))
        return p.value

    def read_device(self, dev_ptr, shape, dtype):
        """Host copy of a device buffer the engine handed out (tb_read_device)."""
        out = np.empty(shape, dtype=dtype)
        self._ck(self._L.tb_read_device(self._h, out.ctypes.data_as(_vp), _vp(dev_ptr), C.c_size_t(out.nbytes)))
        return out


class _GroupLib:
    """Routes Engine's tb_xxx calls to their tb_group_xxx namesakes (same signatures, global indices)."""

    def __init__(self, L):
        self._L = L

    def __getattr__(self, name):
        if name.startswith("tb_") and hasattr(self._L, "tb_group_" + name[3:]):
            return getattr(self._L, "tb_group_" + name[3:])
        raise AttributeError("%s has no tb_group_ counterpart" % name)


class Group(Engine):
    """One tb_group: several engines in this process behind one handle (tb_create_partitioned / tb_batch_create).
    Same methods as Engine, with global mass / spring indices."""

    def __init__(self, devices, flags=MATH_FAST, batch_sizes=None):
        L = lib()
        dev, dp = _arr(devices, np.int32)
        h = _vp()
        if batch_sizes is None:
            rc = L.tb_create_partitioned(dev.shape[0], dp, int(flags), C.byref(h))
        else:
            bs, bp = _arr(batch_sizes, np.int64)
            rc = L.tb_batch_create(dev.shape[0], dp, int(flags), bs.shape[0], bp, C.byref(h))
        if rc != 0:
            raise TitanError(rc, L.tb_last_error(None).decode())
        self._raw = L
        self._L = _GroupLib(L)
        self._h = h
        self.nm = self.ns = 0

    def set_topology(self, nm, left, right, owner=None):
        l, lp = _arr(left, np.int32)
        r, rp = _arr(right, np.int32)
        o, op = _arr(owner, np.int32)
        self._ck(self._raw.tb_group_set_topology(self._h, int(nm), l.shape[0], lp, rp, op))
        self.nm, self.ns = int(nm), int(l.shape[0])

    def engines(self):
        return [self._raw.tb_group_engine(self._h, r) for r in range(self._raw.tb_group_size(self._h))]

    def info(self):
        """per-engine tb_info dicts"""
        out = []
        for e in self.engines():
            i = Info()
            self._raw.tb_get_info(_vp(e), C.byref(i))
            out.append({f: getattr(i, f) for f, _ in Info._fields_})
        return out

    def alloc_pinned(self, shape, dtype=np.float64):
        n = int(np.prod(shape)) * np.dtype(dtype).itemsize
        p = _vp()
        rc = self._raw.tb_alloc_host(_vp(self.engines()[0]), C.c_size_t(n), C.byref(p))
        if rc != 0:
            raise TitanError(rc, "tb_alloc_host failed")
        return np.frombuffer((C.c_char * n).from_address(p.value), dtype=dtype).reshape(shape)

    def set_local_constraints(self, rows):
        raise NotImplementedError("per-mass local constraints are not routed through tb_group")
