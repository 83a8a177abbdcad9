"""The DLRM-small example (reference examples/cpp/DLRM/dlrm.cc with our
embedding operator) runs end to end and learns on its fixed synthetic batch."""
import importlib.util
import os

import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_dlrm_small_end_to_end(capsys):
    if not torch.cuda.is_available():
        pytest.fail("needs a CUDA device")
    spec = importlib.util.spec_from_file_location("dlrm_small", os.path.join(ROOT, "examples", "dlrm_small.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    first, last = mod.main(["-e", "2", "--iterations", "10", "--arch-embedding-size", "-".join(["20000"] * 8)])
    out = capsys.readouterr().out
    assert "THROUGHPUT" in out and "samples/s" in out          # the reference's result line (dlrm.cc:197)
    assert last == last and last < first                       # finite, and the fixed batch is being fitted


def test_dlrm_small_sharded_with_the_reference_strategy_file(tmp_path):
    """The same example under torchrun with -ll:gpu 2 and a strategy file written
    by the reference's own generator rule (dlrm_strategy.py:26-46): tables are
    placed by the file, both exchanges run, and the loss goes down."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import socket
    import subprocess
    import sys
    import dlrm_flexflow_b200 as ff
    path = str(tmp_path / "dlrm_strategy_8embs_2gpus.pb")
    ff.Strategy.dlrm_simple(2, 8).save(path)
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", str(port),
                        os.path.join(ROOT, "examples", "dlrm_small.py"), "-ll:gpu", "2", "--import-strategy", path,
                        "-e", "2", "--iterations", "10", "--arch-embedding-size", "-".join(["20000"] * 8)],
                       cwd=ROOT, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr[-3000:]
    assert "THROUGHPUT" in p.stdout
    first, last = [float(x) for x in p.stdout.split("loss: first")[1].split()[0:3:2]]
    assert last < first
