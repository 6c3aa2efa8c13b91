"""CPU multi-process tests (gloo, world_size 2 and 3): host-side logic of the
N>1 path and the launch contract of bench.py's reference arm."""
import json
import os
import socket
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _torchrun(nproc, script, *args, timeout=600):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()), str(script), *args]
    env = dict(os.environ, OMP_NUM_THREADS="1", CUDA_VISIBLE_DEVICES="")
    return subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env, cwd=ROOT)


@pytest.mark.parametrize("world", [2, 3])
def test_host_side_plan_is_consistent_across_ranks(lib, oracle, world):
    p = _torchrun(world, ROOT / "tests" / "mp_cpu_worker.py")
    assert p.returncode == 0, (p.stdout + p.stderr)[-3000:]
    assert "mp_cpu_worker OK" in p.stdout


def test_reference_arm_under_torchrun_prints_one_line(lib, oracle):
    """bench.py --impl reference at N=2: rank 0 alone runs and prints."""
    p = _torchrun(2, ROOT / "bench.py", "--impl", "reference", "--gpus", "2", "--steps", "1",
                  "--warmup", "0", "--workload", "C2")
    assert p.returncode == 0, (p.stdout + p.stderr)[-3000:]
    lines = [json.loads(l) for l in p.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    line = lines[0]
    assert line["impl"] == "reference" and line["value"] > 0 and line["n_gpus"] == 2
    assert line["cpu_baseline"]["kind"] in ("reference", "port")
    assert line["e2e"]["h2d_bytes_per_step"] == 0
