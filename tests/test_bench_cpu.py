"""bench.py's reference arm (`--impl reference`: the oracle's restatement of the
reference's two-pass Bloom algorithm on the host cores) keeps the JSON contract the
driver reads, alone and under torchrun (rank 0 prints, the other ranks exit 0 without
work).  Tiny sample here; the driver runs it at full size."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEYS = ["impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step",
        "higher_is_better", "scaling", "vs_baseline", "dtype", "data", "config",
        "cpu_baseline", "e2e"]


def check(line, n_gpus):
    d = json.loads(line)
    for key in KEYS:
        assert key in d, key
    assert d["impl"] == "reference" and d["n_gpus"] == n_gpus
    assert d["unit"] == "Gbp/s" and d["higher_is_better"] is True
    assert d["value"] > 0 and d["ms_per_step"] > 0 and d["steps"] == 1
    assert d["vs_baseline"] is None and d["data"] == "synthetic"
    # N > 1: the GPU arm's job of that N (one cluster per GPU), the same config on both arms
    assert d["config"]["workload"].startswith("brucella17") and d["config"]["anchor_size"] == 20
    assert (n_gpus == 1) == (d["config"]["workload"] == "brucella17")
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"]
    assert "sample" in cb and cb["unit"] == d["unit"]
    e = d["e2e"]
    assert e["value"] == d["value"] and e["unit"] == d["unit"]
    assert e["h2d_bytes_per_step"] == 0 and e["d2h_bytes_per_step"] == 0


def test_reference_arm_json_line():
    out = subprocess.run(
        [sys.executable, "bench.py", "--impl", "reference", "--cpu-scale", "0.02",
         "--steps", "1", "--warmup", "0"],
        cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, lines   # ONE JSON line on stdout
    check(lines[0], 1)


def test_reference_arm_under_torchrun_prints_once():
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    out = subprocess.run(
        [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
         "--master-addr", "127.0.0.1", "--master-port", "29711", "bench.py", "--impl",
         "reference", "--gpus", "2", "--cpu-scale", "0.02", "--steps", "1", "--warmup", "0"],
        cwd=ROOT, capture_output=True, text=True, timeout=900, env=env)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.strip().startswith("{")]
    assert len(lines) == 1, lines
    check(lines[0], 2)
