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
