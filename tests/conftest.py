import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: test needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "multigpu: test needs >= 2 CUDA devices")
    config.addinivalue_line("markers", "slow: long-running integration test")


def _n_gpus():
    try:
        import torch
        return torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:
        return 0


def pytest_collection_modifyitems(config, items):
    n = _n_gpus()
    skip_gpu = pytest.mark.skip(reason="no CUDA device")
    skip_multi = pytest.mark.skip(reason="needs >= 2 CUDA devices")
    for item in items:
        if "gpu" in item.keywords and n == 0:
            item.add_marker(skip_gpu)
        if "multigpu" in item.keywords and n < 2:
            item.add_marker(skip_multi)


def pytest_sessionfinish(session, exitstatus):
    session.config._dtf_exitstatus = int(exitstatus)


def pytest_unconfigure(config):
    """A test that leaves a non-daemon thread behind (a queue runner, a supervisor service, an RPC
    client blocked on a dead peer) would keep the interpreter alive after the summary line and look
    like a hung test run. Name the culprits and exit with pytest's own status instead."""
    import threading
    stuck = [t for t in threading.enumerate()
             if t is not threading.main_thread() and t.is_alive() and not t.daemon]
    if stuck:
        sys.stderr.write("conftest: non-daemon threads still alive at exit: %s\n" % [t.name for t in stuck])
        sys.stderr.flush()
        sys.stdout.flush()
        os._exit(getattr(config, "_dtf_exitstatus", 0))
