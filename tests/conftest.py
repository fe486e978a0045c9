import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "modern-rzip_b200"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def gpu_ctx():
    import mrz_py
    ctx = mrz_py.Context(0, 1 << 28, poison=1)   # every work buffer starts as 0xA5: a read-before-write cannot hide
    yield ctx
    ctx.close()


@pytest.fixture(scope="session")
def sim_ctx():
    """The same kernel sources compiled for the CPU emulator (tests/cusim) — no GPU needed."""
    import subprocess
    import mrz_py
    d = os.path.join(ROOT, "tests", "cusim")
    subprocess.check_call(["make", "-C", d, "libmrz_sim.so"], stdout=subprocess.DEVNULL)
    # a cluster of 2 CTAs x 64 threads per automaton (MRZ_SIM_GEOMETRY=CLxT overrides); every work buffer starts poisoned (0xA5), not zeroed
    lib = mrz_py.bind(os.path.join(d, "libmrz_sim.so"))
    lib.mrz_sim_selfcheck(1)    # per batch: hash_count and the sweep ring are compared with the table itself (parity_common.check asserts)
    cl, t = (int(x) for x in os.environ.get("MRZ_SIM_GEOMETRY", "2x64").split("x"))   # soak runs: another cluster shape
    ctx = mrz_py.Context(0, 1 << 26, library=lib, cluster_size=cl, threads=t, poison=1)
    yield ctx
    ctx.close()
