"""The bench.py contract on a GPU box, at a reduced trajectory length: one JSON line with the keys the driver
reads, the roofline / cpu_baseline / e2e objects, and numbers that are consistent with each other."""
import json
import os
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _run(*extra):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--knots", "101", "--steps", "2", "--warmup", "1", *extra],
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    return json.loads(r.stdout.strip().splitlines()[-1])


def test_bench_line_has_the_contract_keys(built_libs):
    d = _run("--ref-units", "32")
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
              "dtype", "data", "config", "e2e", "gpu_launches", "clocks", "roofline", "cpu_baseline"):
        assert k in d, k
    assert d["unit"] == "knot-evals/s" and d["dtype"] == "f64" and d["n_gpus"] == 1 and d["vs_baseline"] is None
    assert "workload" in d["config"] and d["config"]["units"] == 100 and d["gpu_launches"] > 0
    assert abs(d["value"] - 100 / (d["ms_per_step"] * 1e-3)) < 1e-6 * d["value"]
    rf = d["roofline"]
    assert rf["bound"] == "tensor" and rf["unit"] == "TFLOP/s" and abs(rf["frac"] - rf["achieved"] / rf["peak"]) < 1e-12
    assert 0.2 < rf["frac"] < 1.0
    e = d["e2e"]
    assert e["value"] > 0 and e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] > 0 and e["value"] < d["value"]
    assert 0.0 < e["frac_of_host_write_peak"] < 1.5 and e["host_threads"] >= 1
    c = d["cpu_baseline"]
    assert c["kind"] == "port" and c["cores"] >= 1 and c["value"] > 0 and "sample" in c
    m = d["callback_mix"]
    assert m["evaluator_ms_per_ipopt_iteration"] < m["evaluator_ms_per_ipopt_iteration_recomputing"]
    p = e["problem_callbacks"]              # the five Ipopt callbacks of the whole problem, one C call each
    assert "note" not in p, p
    assert p["m"] == 100 * 8192 + 100 * 24 + 100 and p["nnz_jac"] > 100 * 8192 and p["ms_per_ipopt_iteration"] > 0


def test_reference_arm_line(built_libs):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--knots", "41", "--steps", "1",
                        "--warmup", "0", "--ref-units", "16"], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stderr[-2000:]
    d = json.loads(r.stdout.strip().splitlines()[-1])
    assert d["impl"] == "reference" and d["unit"] == "knot-evals/s" and d["value"] > 0
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": "knot-evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
