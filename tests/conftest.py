import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session", autouse=True)
def _native_built():
    """Build the native libraries if they are missing (nvcc/g++ cross-compile without a GPU)."""
    from cekirdekler_b200 import _build

    lib = ROOT / "cekirdekler_b200" / "lib"
    if not (lib / "libKutuphaneCL.so").exists() or not (lib / "libcekirdekler_host.so").exists():
        _build.build_shim()
        _build.build_host()
    if not (ROOT / "oracle" / "liboracle.so").exists():
        _build.build_oracle()
    yield


def gpu_count() -> int:
    from cekirdekler_b200._native import shim

    return shim().ckDeviceCount()


@pytest.fixture(scope="session")
def gpus():
    n = gpu_count()
    if n <= 0:
        pytest.skip("no CUDA device")
    return n
