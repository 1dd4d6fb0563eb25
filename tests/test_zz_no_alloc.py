"""tb_get_mass_state_static: the download entry point that allocates nothing (its gather scratch is a file-scope
`static __device__` array inside libtitan_b200.so), against the ordinary tb_get_mass_state on the same engine.

Collected last on purpose.  Bit-exact: both entry points run the same gather kernel on
the same device arrays, only the scratch buffer differs.  Two calls in a row (the scratch keeps the previous call's
contents: nothing in it has to start at zero, every byte a call reads back it has written first).
"""
import ctypes as C

import numpy as np
import pytest

import scenes
import titan_b200 as tb

pytestmark = pytest.mark.gpu

_vp = C.c_void_p


def static_get(eng, first, count, idx, want_pos=True, want_vel=True):
    L = tb.lib()
    pos = np.full((count, 3), np.nan)
    vel = np.full((count, 3), np.nan)
    ia = None if idx is None else np.ascontiguousarray(idx, dtype=np.int32)
    rc = L.tb_get_mass_state_static(eng._h, C.c_int64(first), C.c_int64(count), None if ia is None else ia.ctypes.data_as(_vp),
                                    pos.ctypes.data_as(_vp) if want_pos else None, vel.ctypes.data_as(_vp) if want_vel else None)
    return rc, pos, vel


def same_bits(a, b):
    return bool((np.ascontiguousarray(a).view(np.uint64) == np.ascontiguousarray(b).view(np.uint64)).all())


def test_static_scratch_download_matches_the_allocating_one_twice():
    L = tb.lib()
    cap = int(L.tb_static_scratch_max_masses())
    assert cap >= 1024
    sc = scenes.c2_lattice(20)                       # 8,000 masses: streaming tiled kernel, state lives tile-blocked
    eng = tb.Engine(0, tb.MATH_FAST)
    scenes.upload(eng, sc)
    t = 0.0
    rng = np.random.RandomState(12345)
    for call in range(2):                            # two calls in a row, the state moves in between
        eng.run(25, t)
        for _ in range(25):
            t += sc.dt
        # whole range
        ref = eng.get_mass_state(fields=("pos", "vel"))
        rc, pos, vel = static_get(eng, 0, sc.nm, None)
        assert rc == 0, eng._err() if hasattr(eng, "_err") else rc
        assert same_bits(pos, ref["pos"]) and same_bits(vel, ref["vel"]), call
        # scattered indices (a container download), pos only / vel only
        idx = rng.permutation(sc.nm)[:777].astype(np.int32)
        ref = eng.get_mass_state(idx=idx, fields=("pos", "vel"))
        rc, pos, _ = static_get(eng, 0, idx.size, idx, want_vel=False)
        assert rc == 0 and same_bits(pos, ref["pos"]), call
        rc, _, vel = static_get(eng, 0, idx.size, idx, want_pos=False)
        assert rc == 0 and same_bits(vel, ref["vel"]), call
        # a sub-range
        ref = eng.get_mass_state(first=100, count=300, fields=("pos",))
        rc, pos, _ = static_get(eng, 100, 300, None, want_vel=False)
        assert rc == 0 and same_bits(pos, ref["pos"]), call
    eng.close()


def test_static_scratch_rejects_more_than_its_maximum():
    L = tb.lib()
    cap = int(L.tb_static_scratch_max_masses())
    n = 48
    sc = scenes.c2_lattice(n)                        # 110,592 masses > the 65,536 the static scratch holds
    assert sc.nm > cap
    eng = tb.Engine(0, tb.MATH_FAST)
    scenes.upload(eng, sc)
    rc, _, _ = static_get(eng, 0, sc.nm, None)
    assert rc == 1                                   # TB_ERR_INVALID
    msg = L.tb_last_error(eng._h).decode()
    assert "static scratch" in msg and str(cap) in msg, msg
    # ... while a window of the same system that fits goes through
    ref = eng.get_mass_state(first=sc.nm - cap, count=cap, fields=("pos",))
    rc, pos, _ = static_get(eng, sc.nm - cap, cap, None, want_vel=False)
    assert rc == 0 and same_bits(pos, ref["pos"])
    eng.close()
