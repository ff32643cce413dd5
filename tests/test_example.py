"""examples/mnist_synthetic.py -- the reference's example script, corrected and
fed synthetic data -- is executed as a file.

* on the UNMODIFIED reference, when /root/reference is present (build container; CPU): the
  script is written against `import rocket` only, so putting the reference (and the test
  shims for its three missing dependencies) on PYTHONPATH is all it takes;
* on this package, on a GPU: the script's own user-written `Accuracy(rocket.Metric)` -- which
  goes through Meter's gather path and reads a ratio back every eval step -- must report the
  same counts as the same graph run with the device-side metric (rocket_b200.metrics.Accuracy).
"""
from __future__ import annotations

import importlib.util
import json
import os
import subprocess
import sys
from pathlib import Path

import pytest
import torch

import oracle

ROOT = Path(__file__).resolve().parent.parent
SCRIPT = ROOT / "examples" / "mnist_synthetic.py"


def _run(args, extra_path=(), env_extra=None, timeout=600):
    env = dict(os.environ, TQDM_DISABLE="1")
    env["PYTHONPATH"] = os.pathsep.join([*map(str, extra_path), env.get("PYTHONPATH", "")])
    env.update(env_extra or {})
    return subprocess.run([sys.executable, str(SCRIPT), *args], capture_output=True, text=True,
                          env=env, timeout=timeout, cwd=str(ROOT))


@pytest.mark.skipif(not oracle.reference_available(), reason="needs /root/reference")
def test_example_file_runs_unchanged_on_the_reference(tmp_path):
    out = tmp_path / "ref.json"
    proc = _run(["--epochs", "1", "--samples", "2048", "--batch-size", "256", "--json", str(out)],
                extra_path=(oracle.SHIMS, oracle.REFERENCE_ROOT),
                env_extra={"ACCELERATE_USE_CPU": "1", "OMP_NUM_THREADS": "4"})
    assert proc.returncode == 0, proc.stderr[-3000:]
    rec = json.loads(out.read_text())
    assert rec["package"] == "rocket"                       # the reference answered the import
    assert rec["history"] and rec["history"][0][1] == 512     # one eval pass over 2048 // 4


@pytest.mark.gpu
def test_example_file_on_this_package_user_metric_equals_device_metric(tmp_path):
    out = tmp_path / "b200.json"
    args = ["--epochs", "2", "--samples", "4096", "--batch-size", "512", "--deterministic"]
    proc = _run([*args, "--json", str(out)])
    assert proc.returncode == 0, proc.stderr[-3000:]
    rec = json.loads(out.read_text())
    assert rec["package"] == "rocket_b200"
    assert len(rec["history"]) == 2 and all(t == 1024 for _, t in rec["history"])

    # the same graph in this process with the device-side metric (integer atomics, no gather)
    torch.backends.cudnn.deterministic = True
    torch.backends.cudnn.benchmark = False
    torch.backends.cudnn.allow_tf32 = False
    had = {k: v for k, v in sys.modules.items() if k == "rocket" or k.startswith("rocket.")}
    try:
        spec = importlib.util.spec_from_file_location("mnist_synthetic_example", SCRIPT)
        example = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(example)
        from rocket_b200.metrics import Accuracy
        launcher, net, accuracy = example.build(2, 4096, 512, accuracy=Accuracy())
        launcher.launch()
    finally:
        for k in [k for k in sys.modules if k == "rocket" or k.startswith("rocket.")]:
            if k not in had:
                del sys.modules[k]
    assert [(p, t) for _, p, t in accuracy.history] == [tuple(x) for x in rec["history"]]
    checksum = float(sum(p.detach().double().sum() for p in net.parameters()))
    assert checksum == rec["checksum"]                      # same training, bit for bit
