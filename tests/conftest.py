"""pytest configuration: the `gpu` marker, repo root on sys.path, and session-wide builds of the checker
libraries (oracle/, tests/hostsim) and of the product library (coolfluid_b200/libcfb200.so)."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _make(directory, *targets):
    subprocess.check_call(["make", "-s", "-C", directory] + list(targets))


@pytest.fixture(scope="session")
def oracle():
    """The CPU oracle (test infrastructure)."""
    from oracle import binding
    have_ref_tree = os.path.isdir("/root/reference/src/MathTools")
    _make(os.path.join(ROOT, "oracle"), "liboracle.so", "liboracle_fast.so", *(["ref"] if have_ref_tree else []))
    return binding


@pytest.fixture(scope="session")
def product_lib():
    """libcfb200.so; built here when nvcc is available, otherwise the prebuilt in-tree file is used."""
    from coolfluid_b200 import lib
    from coolfluid_b200 import build
    build.ensure()           # rebuilds when the sources changed since the last build: never a stale binary
    return lib.load()


@pytest.fixture(scope="session")
def hostsim_lib():
    """TEST ONLY: the product's kernel functors compiled for the host (tests/hostsim)."""
    from coolfluid_b200 import lib
    d = os.path.join(ROOT, "tests", "hostsim")
    _make(d)
    return lib.load(os.path.join(d, "libhostsim.so"), "hostsim_")


def has_gpu():
    try:
        from coolfluid_b200 import lib
        return lib.load().device_count() > 0
    except Exception:
        return False
