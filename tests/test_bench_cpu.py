"""bench.py's contract, as far as it can be checked without a GPU: the reference arm prints exactly one JSON line on
stdout with the keys the driver reads, and the committed round result (profiles/bench_r02.json) carries the
roofline, CPU-baseline and end-to-end objects, for the headline workload and for the target box."""
import json
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent

BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
             "vs_baseline", "dtype", "data", "config", "e2e"}


def test_reference_arm_prints_one_json_line():
    r = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, r.stdout
    d = json.loads(lines[0])
    assert d["impl"] == "reference"
    if "unavailable" in d:
        return
    assert BASE_KEYS <= set(d)
    assert d["metric"] == "site_pair_interactions_per_s" and d["unit"] == "site-pairs/s" and d["higher_is_better"] is True
    assert d["value"] > 1e6 and d["dtype"] == "f64" and "workload" in d["config"]
    assert d["e2e"]["value"] == d["value"] and d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("reference", "port") and cb["cores"] >= 1 and cb["sample"]


def test_committed_round_result_has_the_contract_keys():
    d = json.loads((ROOT / "profiles" / "bench_r02.json").read_text())
    assert BASE_KEYS | {"gpu_launches", "clocks", "roofline", "cpu_baseline"} <= set(d)
    assert d["n_gpus"] == 1 and d["warmup"] >= 3 and d["gpu_launches"] >= d["steps"]
    rf = d["roofline"]
    assert {"bound", "achieved", "peak", "unit", "frac", "traffic"} <= set(rf)
    assert abs(rf["frac"] - rf["achieved"] / rf["peak"]) < 1e-9
    e = d["e2e"]
    assert e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] > 0 and 0 < e["value"] <= d["value"]
    assert not set(d["clocks"]["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    assert d["oracle_check"]["ok"] and d["oracle_check"]["all_ranks_ok"]
    # the box north_star's target is quoted on rides in the same line
    t = d["target_box"]
    assert "240 molecules" in t["workload"] and t["chains_per_gpu"] == 512 and t["oracle_check"]["ok"]
    assert abs(t["roofline"]["frac"] - t["roofline"]["achieved"] / t["roofline"]["peak"]) < 1e-9
    assert 0 < t["e2e"]["value"] <= t["value"] and t["gpu_launches"] >= d["steps"]
    for r in (rf, t["roofline"]):
        assert r["dram_bytes_per_move"] > 0 and r["l2_bytes_per_move"] > 0
