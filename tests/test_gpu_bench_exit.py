"""bench.py must print ONE JSON line and EXIT -- with one rank and, launched
the way the driver launches it, with two (round 1's two-rank run printed its
line and then sat in communicator teardown until it was killed)."""
import json
import os
import socket
import subprocess
import sys

import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _one_json_line(stdout):
    lines = [ln for ln in stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, lines
    return json.loads(lines[0])


def test_bench_one_rank_small_config_exits():
    if not torch.cuda.is_available():
        pytest.fail("needs a CUDA device")
    p = subprocess.run([sys.executable, "bench.py", "--config", "C1", "--steps", "5", "--warmup", "3",
                        "--e2e-steps", "2", "--no-cpu-baseline"], cwd=ROOT, capture_output=True, text=True,
                       timeout=240)
    assert p.returncode == 0, p.stderr[-2000:]
    line = _one_json_line(p.stdout)
    assert line["n_gpus"] == 1 and line["value"] > 0 and line["gpu_launches"] > 0
    assert line["e2e"]["value"] > 0 and line["e2e"]["h2d_bytes_per_step"] > 0


def test_bench_two_ranks_exits():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", str(port), "bench.py", "--gpus", "2",
                        "--config", "C2", "--steps", "5", "--warmup", "3", "--e2e-steps", "2", "--no-cpu-baseline"],
                       cwd=ROOT, capture_output=True, text=True, timeout=240)
    assert p.returncode == 0, p.stderr[-2000:]
    line = _one_json_line(p.stdout)
    assert line["n_gpus"] == 2 and line["value"] > 0 and line["gpu_launches"] > 0
    assert line["e2e"]["value"] > 0
