"""bench.py's JSON contract, checked on CPU through the reference arm (the oracle port on a tiny
workload) — the GPU arm prints the same keys plus roofline / clocks / gpu_launches."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_one_contract_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "lenet",
                          "--steps", "2", "--warmup", "1"], capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "images/s" and d["value"] > 0 and d["higher_is_better"] is True
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["config"]["workload"].startswith("lenet") and d["steps"] == 2 and d["warmup"] == 1


def test_non_zero_ranks_of_the_reference_arm_exit_quietly():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"],
                         capture_output=True, text=True, timeout=120, cwd=ROOT, env=env)
    assert out.returncode == 0 and out.stdout.strip() == ""


def test_committed_bench_line_carries_the_contract_keys_and_feeds_the_roofline_table():
    """The last GPU bench line committed under profiles/ has every key the driver reads, and tools/roofline_table.py
    turns its per-layer event times into the per-layer roofline table (fractions of the peaks the line itself names)."""
    path = os.path.join(ROOT, "profiles", "r2_bench_caffenet_tf32_final.json")
    d = json.loads(open(path).read().strip().splitlines()[-1])
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
              "dtype", "data", "config", "e2e", "e2e_u8", "gpu_launches", "clocks", "roofline", "roofline_hbm", "parity", "cpu_baseline"):
        assert k in d, k
    assert d["dtype"] == "tf32" and d["n_gpus"] == 1 and d["gpu_launches"] > 0 and "workload" in d["config"]
    assert d["e2e"]["h2d_bytes_per_step"] == 256 * (3 * 227 * 227 + 1) * 4 and d["e2e"]["value"] < d["value"] * 1.02
    r = d["roofline"]
    assert r["bound"] == "tensor" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9 and r["traffic"] > 0
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1
    assert d["parity"]["ok"] is True and d["clocks"]["samples"] > 0 and not d["clocks"]["reasons"]
    assert all(row["frac"] <= 1.0 for row in d["roofline_hbm"]["per_kernel"]) and 0 < d["roofline_hbm"]["frac"] <= 1.0
    assert d["e2e_u8"]["h2d_bytes_per_step"] == 256 * (3 * 227 * 227) + 256 * 4
    out = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "roofline_table.py"), path], capture_output=True, text=True,
                         timeout=60, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-1000:]
    rows = [l for l in out.stdout.splitlines() if l.startswith("| conv") or l.startswith("| fc") or l.startswith("| pool") or l.startswith("| norm")]
    assert len(rows) == 8 + 5
    for l in rows:
        cells = [c.strip() for c in l.strip("|").split("|")]
        frac = float(cells[4])
        assert 0.0 < frac < 1.2, l
