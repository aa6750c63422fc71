import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
GOLDEN = os.path.join(ROOT, "tests", "golden")
# the library honours its development A-B switches (MOCOB200_FD_LOCAL, _KERNEL, ...) only with this set (kernels.cuh: devEnv)
os.environ.setdefault("MOCOB200_DEVELOPER_SWITCHES", "1")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def pkg():
    """The product's host harness (ctypes over libmocob200.so); builds the library if it is missing."""
    import __graft_entry__ as ge

    if not os.path.exists(os.path.join(ROOT, "opensim-moco_b200", "libmocob200.so")):
        ge.build()
    return ge.package()


@pytest.fixture(scope="session")
def orc(pkg):
    """The CPU oracle (test infrastructure)."""
    import oracle_py

    oracle_py.lib()
    return oracle_py


def golden(name):
    return os.path.join(GOLDEN, name)


FIXTURES = ["sliding_mass", "sliding_mass_trap", "hanging_muscle", "hanging_muscle_rigidtendon", "pendulum2",
            "pendulum3_implicit", "gait2d_tracking", "gait2d_prediction", "pendulum_cb_lower", "pendulum_cb_upper_trap",
            "pendulum_cb_equality", "pendulum_cb_both",
            # kinematic constraints: multipliers, slacks, velocity correction
            "point_mass_coupler", "point_mass_coupler_implicit", "double_pendulum_coupler", "double_pendulum_coupler_noderiv_trap",
            # prescribed kinematics (PositionMotion, MocoInverse on the 2D gait model) and its free-coordinate twin
            "gait2d_inverse", "gait2d_inverse_trap_central", "gait2d_inverse_free",
            # function-based joint driven by one coordinate (the knee of the full-body models) on a model small enough for both
            # kernel families: the reference's velocity-product term of that case (tests/test_reference_dynamics_pins.py)
            "knee_chain", "knee_chain_implicit_trap"]
