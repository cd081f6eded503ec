import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
for p in (str(ROOT), str(ROOT / "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


def pytest_collection_modifyitems(config, items):
    """Without a CUDA device every gpu-marked test is skipped (liborbx has no CPU fallback, and the C++ drop-in
    aborts the interpreter when a call fails), whatever -m / -k selected."""
    gpu_items = [it for it in items if it.get_closest_marker("gpu")]
    if not gpu_items:
        return
    try:
        import orbslam2_b200
        have = orbslam2_b200.device_count() > 0
    except Exception:
        have = False
    if not have:
        skip = pytest.mark.skip(reason="no CUDA device")
        for it in gpu_items:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def pkg():
    import orbslam2_b200
    return orbslam2_b200


@pytest.fixture(scope="session")
def oracle():
    import oracle_lib
    oracle_lib.lib()
    return oracle_lib
