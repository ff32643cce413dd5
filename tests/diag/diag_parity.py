"""Diagnostic (GPU): how far apart are the arms after N steps of the mnist-shaped graph?"""
import os, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
os.environ["TQDM_DISABLE"] = "1"
import torch
import workloads
from oracle import looper_port
import rocket_b200 as rk
from rocket_b200.metrics import Accuracy

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
torch.backends.cudnn.deterministic = True
torch.backends.cudnn.benchmark = False

def rel(a, b):
    num = sum(float((a[k] - b[k]).pow(2).sum()) for k in a)
    den = sum(float(b[k].pow(2).sum()) for k in a)
    worst = max(float((a[k] - b[k]).norm() / (b[k].norm() + 1e-30)) for k in a)
    return (num / den) ** 0.5, worst

CASES = {
    "adamw_accum2": dict(optimizer="adamw", max_norm=None, accum=2, epochs=2, n_train=64 * 25),
    "adamw_clip": dict(optimizer="adamw", max_norm=0.5, accum=2, epochs=2, n_train=64 * 25),
    "sgd_clip": dict(optimizer="sgd", max_norm=1.0, accum=1, epochs=1, n_train=64 * 20),
    "adam_l2": dict(optimizer="adam", max_norm=None, accum=3, epochs=1, n_train=64 * 20),
}
for name, kw in CASES.items():
    t1 = looper_port.run_mnist_port(device="cuda", workloads=workloads, **kw)["params"]
    t2 = looper_port.run_mnist_port(device="cuda", workloads=workloads, **kw)["params"]
    o1 = looper_port.run_mnist_port(device="cuda", workloads=workloads, optimizer_impl="oracle", **kw)["params"]
    launcher, h = workloads.build_mnist_graph(rk, accuracy_cls=Accuracy, clip_via="kwarg", **kw)
    launcher.launch()
    e = {k: p.detach().cpu() for k, p in h["net"].named_parameters()}
    print(name, "torch-cuda twice", rel(t1, t2), "| engine vs oracle-arm", rel(e, o1),
          "| engine vs torch-cuda", rel(e, t1), "| oracle-arm vs torch-cuda", rel(o1, t1), flush=True)
