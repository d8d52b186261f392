"""bench.py on a GPU: one JSON line with every key of the measurement contract (small batch)."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("extra", [[], ["--graph"]])
def test_bench_line_has_the_contract_keys(native_lib, extra):
    env = dict(os.environ, PYTHONPATH=ROOT)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "3", "--warmup", "3", "--batch", "256",
                        "--no-cpu-baseline"] + extra, capture_output=True, text=True, env=env, timeout=600)
    assert r.returncode == 0, r.stderr[-3000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, r.stdout
    d = json.loads(lines[0])
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "e2e", "gpu_launches", "clocks", "roofline"):
        assert k in d, k
    assert d["n_gpus"] == 1 and d["steps"] == 3 and d["warmup"] == 3 and d["value"] > 0 and d["vs_baseline"] is None
    assert d["value"] == pytest.approx(256 / d["ms_per_step"] * 1e3, rel=1e-6)
    assert d["e2e"]["h2d_bytes_per_step"] > 0 and d["e2e"]["d2h_bytes_per_step"] > 0 and d["e2e"]["value"] > 0
    assert d["gpu_launches"] > 0 and "workload" in d["config"]
    roof = d["roofline"]
    assert roof["bound"] in ("hbm", "tensor") and roof["peak"] > 0 and 0 < roof["frac"] < 1.5
    assert roof["achieved"] == pytest.approx(roof["frac"] * roof["peak"], rel=1e-6)
