"""Diagnostic (torchrun, 2 GPUs): the ws2 capsule graph with Module(graph=True); dumps stacks if stuck."""
import faulthandler, os, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
os.environ["TQDM_DISABLE"] = "1"
faulthandler.dump_traceback_later(int(os.environ.get("DUMP_AFTER", "50")), exit=True)
import torch
import rocket_b200 as rk
import workloads, dist_worker
from rocket_b200.metrics import Accuracy
from rocket_b200.tracking import MemoryTracker
case = sys.argv[1] if len(sys.argv) > 1 else "ws2_adamw_clip"
kw = dict(dist_worker.WS2_CASES[case])
for a in sys.argv[2:]:
    k, v = a.split("="); kw[k] = int(v)
tracker = MemoryTracker("memory") if os.environ.get("TRACKER", "1") == "1" else None
launcher, h = workloads.build_mnist_graph(rk, tracker_backend=tracker, clip_via="kwarg" if kw["max_norm"] else "capsule",
                                          accuracy_cls=Accuracy, graph=True, **kw)
launcher.launch()
print("rank", os.environ.get("RANK"), "done", h["accuracy"].history, flush=True)
