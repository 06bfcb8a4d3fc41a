import os
import sys

# The CPU suite runs many short torch ops from many Python threads (threaded simulators, transports).  Every such thread owns an
# OpenMP team; with the default active wait policy the idle teams spin and starve whoever runs next - the same test then takes 5-10x
# longer late in the session than alone.  Must be set before torch (libgomp) is loaded; inherited by the example subprocesses.
os.environ.setdefault("OMP_WAIT_POLICY", "PASSIVE")
os.environ.setdefault("GOMP_SPINCOUNT", "0")

import pytest  # noqa: E402

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "multigpu: needs >= 2 CUDA devices")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
        ngpu = torch.cuda.device_count() if has_gpu else 0
    except Exception:
        has_gpu, ngpu = False, 0
    skip_gpu = pytest.mark.skip(reason="no CUDA device")
    skip_multi = pytest.mark.skip(reason="needs >= 2 CUDA devices")
    for item in items:
        if "gpu" in item.keywords and not has_gpu:
            item.add_marker(skip_gpu)
        if "multigpu" in item.keywords and ngpu < 2:
            item.add_marker(skip_multi)


@pytest.fixture(scope="session")
def native_lib():
    from fedml_b200.native.build import build
    build()
    from fedml_b200.ops import native
    return native.lib()


@pytest.fixture(autouse=True)
def _run_in_tmp_cwd(tmp_path, monkeypatch):
    """The reference-compatible defaults write ./log and ./model_file_cache: keep them out of the source tree."""
    monkeypatch.chdir(tmp_path)
