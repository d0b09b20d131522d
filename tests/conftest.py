import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import __graft_entry__ as entry  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu under gpurun)")


@pytest.fixture(scope="session")
def pkg():
    if not os.path.exists(os.path.join(entry.PKG_DIR, "libgsb200.so")):
        entry.build()
    return entry.load_package()


@pytest.fixture(scope="session")
def sim(pkg):
    return pkg.simulation


@pytest.fixture(scope="session")
def orc():
    o = entry.load_oracle()
    o.build()
    return o


# examples/settings-files.toml:1-9 (the reference's own physics for every BASELINE config)
PHYS = dict(Du=0.2, Dv=0.1, F=0.02, k=0.048, dt=1.0)
# Settings() defaults, Structs.jl:8-13 (the unit-Simulation.jl:19-25 case)
PHYS_DEFAULT = dict(Du=0.05, Dv=0.1, F=0.04, k=0.0, dt=0.2)
