"""bench.py's contract, the part that can be checked without a GPU: the reference arm
(`--impl reference` = the looper port on the host cores) prints one JSON line with the agreed
keys, and ranks other than 0 stay silent."""
from __future__ import annotations

import json
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def _run(env_extra=None):
    env = dict(os.environ, TQDM_DISABLE="1", OMP_NUM_THREADS="4")
    env.update(env_extra or {})
    return subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference",
                           "--steps", "1", "--warmup", "1"], capture_output=True, text=True,
                          env=env, timeout=600)


def test_reference_arm_prints_the_contract_line():
    proc = _run()
    assert proc.returncode == 0, proc.stderr[-2000:]
    lines = [ln for ln in proc.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    line = json.loads(lines[0])
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step",
                "higher_is_better", "scaling", "vs_baseline", "dtype", "data", "config", "e2e",
                "cpu_baseline"):
        assert key in line, key
    assert line["impl"] == "reference" and line["higher_is_better"] is True
    assert line["unit"] == "samples/s" and line["value"] > 0
    assert line["vs_baseline"] is None                      # BASELINE.md publishes no number
    assert "workload" in line["config"] and "model" not in line["config"]
    assert line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1
    assert line["cpu_baseline"]["value"] == line["value"] == line["e2e"]["value"]
    assert line["e2e"]["h2d_bytes_per_step"] == line["e2e"]["d2h_bytes_per_step"] == 0
    # same --steps and the same config.workload string as our arm prints (the driver compares)
    assert line["steps"] == 1 and line["cpu_baseline"]["steps"] == 1
    sys.path.insert(0, str(ROOT))
    import bench
    assert line["config"]["workload"] == bench.WORKLOADS["resnet18"]["label"]
    assert line["dtype"] == bench.WORKLOADS["resnet18"]["dtype"]


def test_reference_arm_is_silent_on_other_ranks():
    proc = _run({"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"})
    assert proc.returncode == 0, proc.stderr[-2000:]
    assert not [ln for ln in proc.stdout.splitlines() if ln.startswith("{")]
