"""Multi-GPU parity (`-m gpu`, needs >= 2 GPUs on the box; skipped otherwise).

One process per GPU over NCCL, launched exactly as the driver launches bench.py.
Each case renders a frame with the boxes sharded over P ranks and compares it
with the oracle run with the same P and box->rank map: rank order and pixel
ownership bit-exact, float RGBA within 1e-4, bytes within 1 LSB, sample count
exact.  The worker is tests/mp_frame_worker.py."""
import os
import socket
import subprocess
import sys
from pathlib import Path

import pytest

pytestmark = pytest.mark.gpu

ROOT = Path(__file__).resolve().parent.parent


def _gpu_count():
    import torch

    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _run(nranks, case, exchange=None):
    if _gpu_count() < nranks:
        pytest.skip(f"needs {nranks} GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nranks}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           str(ROOT / "tests" / "mp_frame_worker.py"), case]
    env = dict(os.environ, OMP_NUM_THREADS="1")
    if exchange is not None:
        env["AMRVR_EXCHANGE"] = exchange  # "nccl": pack + alltoallv instead of peer-memory loads
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    if p.returncode != 0 and "EADDRINUSE" in p.stderr:  # the port found free was taken meanwhile: once more
        cmd[cmd.index("--master-port") + 1] = str(_free_port())
        p = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
    tail = (p.stdout + p.stderr)[-3000:]
    assert p.returncode == 0, tail
    assert "OK" in p.stdout, tail


@pytest.mark.parametrize("case", ["c4", "c3_aa4", "c2prime", "c1_few_boxes", "c5"])
def test_two_ranks(case):
    _run(2, case)


@pytest.mark.parametrize("case", ["c4", "c3_aa4"])
def test_two_ranks_nccl_exchange(case):
    """The same frames with the peer-memory path switched off."""
    _run(2, case, exchange="nccl")


@pytest.mark.parametrize("case", ["c4", "c2prime"])
def test_four_ranks(case):
    _run(4, case)


@pytest.mark.parametrize("case", ["c4", "c3_aa4", "c1_few_boxes", "c5"])
def test_eight_ranks(case):
    _run(8, case)


def test_render_under_torchrun(tmp_path):
    """render() without a context on 2 ranks (Context.from_env) against the
    same plotfile rendered by one rank; worker: tests/mp_render_worker.py."""
    if _gpu_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           str(ROOT / "tests" / "mp_render_worker.py"), str(tmp_path)]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600,
                       env=dict(os.environ, OMP_NUM_THREADS="1"), cwd=ROOT)
    tail = (p.stdout + p.stderr)[-3000:]
    assert p.returncode == 0, tail
    assert "OK" in p.stdout, tail
