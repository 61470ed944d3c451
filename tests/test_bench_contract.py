"""bench.py prints the one JSON line the driver reads, with every key of the measurement contract (small workload)."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args):
    res = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True, timeout=600,
                         cwd=ROOT)
    assert res.returncode == 0, res.stderr[-2000:]
    lines = [ln for ln in res.stdout.strip().splitlines() if ln.startswith("{")]
    assert len(lines) == 1, res.stdout[-2000:]
    return json.loads(lines[0])


@pytest.mark.gpu
def test_bench_line_has_the_contract_keys():
    line = _run("--images", "4096", "--steps", "3", "--warmup", "3", "--cpu-sample", "256", "--no-other-configs")
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
              "dtype", "data", "config", "e2e", "gpu_launches", "clocks", "roofline", "cpu_baseline"):
        assert k in line, k
    assert line["n_gpus"] == 1 and line["steps"] == 3 and line["higher_is_better"] is True and line["scaling"] == "weak"
    assert line["value"] > 0 and line["ms_per_step"] > 0 and line["gpu_launches"] > 0 and line["vs_baseline"] is None
    assert "workload" in line["config"] and line["data"] == "synthetic"
    e2e = line["e2e"]
    assert e2e["value"] > 0 and e2e["h2d_bytes_per_step"] == 2 * 4096 * 5880 and e2e["d2h_bytes_per_step"] > 0
    roof = line["roofline"]
    assert roof["bound"] == "hbm" and roof["unit"] == "GB/s" and roof["peak"] > 0
    assert roof["frac"] == pytest.approx(roof["achieved"] / roof["peak"])
    cpu = line["cpu_baseline"]
    assert cpu["kind"] == "port" and cpu["value"] > 0 and cpu["cores"] >= 1 and cpu["sample"]
    assert {"sm_mhz", "sm_max_mhz", "reasons"} <= set(line["clocks"])
    chk = line["result_check"]["sensitive"]
    assert all(v is True for v in chk["agree"].values()) and len(chk["sha256"]) >= 16
    hl = line["head_linear_loss"]
    assert hl["roofline"]["bound"] == "tensor" and hl["loss_parts_agree_with_two_calls"] is True
    assert hl["fp32_operands"]["loss_parts_agree_with_two_calls_rtol_1e-5"] is True


@pytest.mark.gpu
def test_reference_arm_line():
    line = _run("--impl", "reference", "--steps", "2", "--warmup", "1", "--cpu-sample", "256")
    assert line["impl"] == "reference" and line["value"] > 0 and line["unit"] == "images/s"
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["value"] == line["value"]
    assert line["e2e"] == {"value": line["value"], "unit": line["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
