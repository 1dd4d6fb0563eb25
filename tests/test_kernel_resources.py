"""Resource usage of the step kernel variants the BASELINE configs run on, read from the built library with cuobjdump (no GPU
needed).  A stack frame in the plain dictionary kernel means some per-thread vector went to local memory -- which is how a
reference parameter of an out-of-line helper once cost 20 % on every config without any test noticing
(profiles/r3z_regression_found.md).  Register counts above 128 would drop the kernel from 3 resident CTAs per SM to 2."""
import os
import re
import shutil
import subprocess

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "..", "titan_b200", "libtitan_b200.so")


def resource_usage():
    exe = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(exe):
        pytest.skip("cuobjdump not found")
    out = subprocess.run([exe, "-res-usage", LIB], capture_output=True, text=True, check=True).stdout
    table = {}
    for name, reg, stack in re.findall(r"Function (\S+):\s*\n\s*REG:(\d+) STACK:(\d+)", out):
        table[name] = (int(reg), int(stack))
    return table


def variant(table, damping, active, tpm, cs, halo, dict_mode, arith="Fast"):
    key = "tiled_step_kernelINS_%d%sELb%dELb%dELi%dELi%dELb%dELi%dE" % (len(arith), arith, damping, active, tpm, cs, halo, dict_mode)
    hits = [v for k, v in table.items() if key in k]
    assert len(hits) == 1, (key, len(hits))
    return hits[0]


def test_library_is_built():
    assert os.path.exists(LIB), "run python -c 'import __graft_entry__ as g; g.build()'"


@pytest.mark.parametrize("dict_mode", [1, 2], ids=["full-dictionary (C2, C4)", "split-dictionary (C5, per-spring k)"])
@pytest.mark.parametrize("active", [0, 1], ids=["passive", "ACTIVE_* springs (C3)"])
def test_dictionary_kernels_keep_three_ctas_per_sm(dict_mode, active):
    reg, stack = variant(resource_usage(), 0, active, 1, 384, 0, dict_mode)
    assert reg <= 128, reg
    if not active:
        assert stack == 0, stack          # nothing in local memory: force, pos, vel and the incidence words live in registers
    else:
        assert stack <= 96, stack         # the sine-table variant spills a little outside the spring loop (measured 160 G on C3)


def test_halo_variant_matches_the_plain_one():
    table = resource_usage()
    reg, stack = variant(table, 0, 0, 1, 384, 1, 1)
    assert reg <= 128 and stack <= 16, (reg, stack)


def test_plain_record_kernel_has_no_stack_frame():
    reg, stack = variant(resource_usage(), 0, 0, 1, 384, 0, 0)
    assert stack == 0 and reg <= 168, (reg, stack)
