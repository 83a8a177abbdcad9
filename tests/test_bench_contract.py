"""CPU checks of bench.py's output contract: the reference arm runs here (it is
the CPU path), prints exactly one JSON line on stdout with the required keys;
and the source of our arm carries every key the driver reads."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

BASE_KEYS = ["metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
             "vs_baseline", "dtype", "data", "config"]


def test_reference_arm_prints_one_json_line():
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--config", "C1",
                        "--steps", "2", "--warmup", "1"], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [l for l in p.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, lines
    d = json.loads(lines[0])
    for k in BASE_KEYS + ["impl", "cpu_baseline", "e2e"]:
        assert k in d, k
    assert d["impl"] == "reference" and d["unit"] == "samples/s" and d["higher_is_better"] is True
    assert d["value"] > 0 and d["vs_baseline"] is None and d["dtype"] == "f32" and d["data"] == "synthetic"
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1
    assert d["cpu_baseline"]["value"] == d["value"] and "capped" in d["cpu_baseline"]["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in d["config"] and "model" not in d["config"]


def test_reference_arm_other_ranks_exit_silently():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2",
                        "--config", "C1", "--steps", "1", "--warmup", "1"], capture_output=True, text=True,
                       timeout=120, env=env)
    assert p.returncode == 0 and p.stdout.strip() == ""


def test_our_arm_emits_the_contract_keys():
    src = open(os.path.join(ROOT, "bench.py")).read()
    for k in BASE_KEYS + ["roofline", "cpu_baseline", "e2e", "clocks", "gpu_launches", "h2d_bytes_per_step",
                          "d2h_bytes_per_step", "sm_mhz", "sm_max_mhz", "reasons", "bound", "achieved", "peak",
                          "frac", "traffic", "cores", "kind", "sample"]:
        assert '"%s"' % k in src, k
    # the oracle is only reached from the CPU legs
    assert src.count("from oracle import oracle") == 1
