"""pytest configuration: `-m "not gpu"` runs on the CPU-only container, `-m gpu` on a B200."""
import importlib
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _ensure_built():
    pkg = os.path.join(ROOT, "raytracing-weekend_b200")
    if not os.path.exists(os.path.join(pkg, "librtw.so")):
        subprocess.check_call(["make", "-C", pkg, "librtw.so"])
    if not os.path.exists(os.path.join(ROOT, "oracle", "liboracle.so")):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "liboracle.so"])


@pytest.fixture(scope="session")
def rtw():
    _ensure_built()
    return importlib.import_module("raytracing-weekend_b200")


@pytest.fixture(scope="session")
def orc():
    _ensure_built()
    from oracle import oracle_py
    return oracle_py


@pytest.fixture(scope="session")
def emu():
    import emu_py
    emu_py.lib()
    return emu_py


@pytest.fixture(scope="session")
def director(rtw):
    """One context on cuda:0 for the whole GPU session; fails loudly without a device."""
    d = rtw.Director(0)
    yield d
    d.destroy()
