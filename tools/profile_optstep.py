"""Small driver for ncu: a few launches of K1 + K2 on one of the configs' tensor lists.

    python tools/profile_optstep.py [--layout flat|resnet18|gpt2|gpt2x8] [--n 268435456] [--iters 4]
                                    [--variant V]

`flat` is one tensor of --n elements (default 2^28 = 4 GiB of optimizer state, far larger than
L2).  For lists that fit partly in L2 a 512 MB buffer is rewritten before every iteration, so
each profiled launch sees what it sees inside a training step (cold p/m/v).
"""
import argparse
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
sys.path.insert(0, str(Path(__file__).resolve().parent))
import torch  # noqa: E402

from rocket_b200 import _native as N  # noqa: E402
from rocket_b200.backend import NativeBackend  # noqa: E402
from bench_step_variants import build, layout  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--layout", default="flat")
ap.add_argument("--n", type=int, default=1 << 28)
ap.add_argument("--iters", type=int, default=4)
ap.add_argument("--variant", type=int, default=-1)
args = ap.parse_args()

dev = torch.device("cuda", 0)
be = NativeBackend(dev)
sizes = [args.n] if args.layout == "flat" else layout(args.layout)
n = sum(sizes)
flat, specs = build(sizes, dev)
plan = be.make_plan(specs)
N.load().rk_step_variant_override(args.variant)
flush = torch.empty(128 << 20, device=dev) if n < (1 << 27) else None
hp = dict(lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=1e-2, decoupled=True)
ts = []
for i in range(args.iters):
    flat["g"].normal_(0, 1e-3)
    if flush is not None:
        flush.fill_(float(i))
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    e[0].record()
    norm = be.l2norm(plan, 1.0)
    e[1].record()
    be.adam_step(plan, [dict(hp, step=i + 1)], norm, 1.0, 1.0, True)
    e[2].record()
    torch.cuda.synchronize()
    ts.append((e[0].elapsed_time(e[1]), e[1].elapsed_time(e[2])))
tn, tk = ts[-1]
print(f"layout={args.layout} n={n} norm {tn:.4f} ms = {4*n/tn/1e6:.1f} GB/s | "
      f"step {tk:.4f} ms = {32*n/tk/1e6:.1f} GB/s")
