"""In-tree build of the native libraries (nvcc cross-compiles sm_100a without a GPU)."""
import os
import subprocess

_PKG = os.path.dirname(os.path.abspath(__file__))


def _make(directory, *targets):
    env = dict(os.environ)
    env.pop("CC", None)   # the image exports CC/CXX=/opt/gcc/..., which lacks OpenMP specs; use distro gcc
    env.pop("CXX", None)
    subprocess.check_call(["make", "-s", "-j8", "-C", directory, *targets], env=env)


def build_engine():
    _make(os.path.join(_PKG, "csrc"))
    return os.path.join(_PKG, "libtitan_b200.so")


def build_host():
    host = os.path.join(_PKG, "host")
    if os.path.exists(os.path.join(host, "Makefile")):
        _make(host)
    return os.path.join(_PKG, "libTitan.so")


def build_programs():
    """User programs of the facade (API driver, C++ tests): tests/cpp/bin/."""
    d = os.path.join(os.path.dirname(_PKG), "tests", "cpp")
    if os.path.exists(os.path.join(d, "Makefile")):
        _make(d)
    return os.path.join(d, "bin")


def build_all():
    build_engine()
    build_host()
    build_programs()
