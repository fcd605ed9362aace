import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _device_count():
    """CUDA devices as the product library sees them (0 on a CPU-only box, or when it cannot load)."""
    try:
        import ctypes
        import fvad_b200
        n = ctypes.c_int32(0)
        rc = fvad_b200.load().capi.lib().fvad_device_count(ctypes.byref(n))
        return int(n.value) if rc == 0 else 0
    except Exception:  # noqa: BLE001
        # library missing or unloadable: on a GPU box the gpu tests must then FAIL loudly, not skip
        return 1 if os.path.exists("/dev/nvidia0") else 0


def pytest_collection_modifyitems(config, items):
    # Without a CUDA device the `gpu` tests are skipped, not failed: the library has no CPU path
    # (every compute entry point returns FVAD_ERR_NO_DEVICE), so they could only error.
    gpu_items = [it for it in items if it.get_closest_marker("gpu")]
    if not gpu_items or _device_count() > 0:
        return
    skip = pytest.mark.skip(reason="no CUDA device: libfvad_cuda has no CPU path")
    for it in gpu_items:
        it.add_marker(skip)


@pytest.fixture(scope="session")
def fv():
    import fvad_b200
    return fvad_b200.load()


@pytest.fixture(scope="session")
def po():
    from oracle import pyoracle
    pyoracle.lib()
    return pyoracle


@pytest.fixture(scope="session")
def wd(fv):
    return fv.weights.synthetic(seed=1234)
