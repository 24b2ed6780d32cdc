import os
import socket
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    import torch

    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def run_dist(module: str, args, nproc: int, timeout: int = 600, env_extra=None) -> str:
    """Run ``python -m torch.distributed.run -m <module> <args>`` on one node; returns combined output."""
    env = dict(os.environ)
    env["PYTHONPATH"] = str(ROOT) + os.pathsep + env.get("PYTHONPATH", "")
    env.setdefault("OMP_NUM_THREADS", "2")
    env.update(env_extra or {})
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}", "--master-addr",
           "127.0.0.1", "--master-port", str(free_port()), "-m", module] + [str(a) for a in args]
    proc = subprocess.run(cmd, cwd=str(ROOT), env=env, capture_output=True, text=True, timeout=timeout)
    out = proc.stdout + proc.stderr
    assert proc.returncode == 0, f"{' '.join(cmd)} failed (rc={proc.returncode}):\n{out[-4000:]}"
    assert "PASS" in out, out[-4000:]
    return out


@pytest.fixture
def dist_runner():
    return run_dist


def gpu_count() -> int:
    import torch

    return torch.cuda.device_count() if torch.cuda.is_available() else 0
