"""bench.py's JSON contract, checked on the arm that runs without a GPU: `--impl reference` (the CPU restatement of the
reference's path on the host cores) prints ONE line with the metric / unit / config of the B200 arm, `impl`, a
`cpu_baseline` describing the run and an `e2e` object with zero copy bytes; under torchrun only rank 0 prints."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def run_bench(*args, env=None):
    e = dict(os.environ)
    e.update(env or {})
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True, env=e, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    return [ln for ln in out.stdout.splitlines() if ln.strip()]


def test_reference_arm_prints_one_contract_line():
    lines = run_bench("--impl", "reference", "--worlds", "8", "--entities", "256", "--steps", "5", "--warmup", "1")
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "batched world-steps/sec" and d["unit"] == "world-steps/s"
    assert d["higher_is_better"] is True and d["scaling"] == "weak" and d["vs_baseline"] is None and d["dtype"] == "f32"
    assert d["value"] > 0 and d["steps"] == 5 and d["n_gpus"] == 1 and "workload" in d["config"]
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and "8 worlds x 256 entities" in cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}


def test_reference_arm_other_ranks_stay_silent():
    assert run_bench("--impl", "reference", "--gpus", "2", "--worlds", "4", "--entities", "64", "--steps", "3", env={"RANK": "1", "WORLD_SIZE": "2"}) == []
