"""Small driver for ncu: a few launches of K1 + K2 on a flat tensor (default 2^28 params =
4 GiB of optimizer state, far larger than L2).  Usage: python tools/profile_optstep.py [n] [iters]"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from rocket_b200.backend import NativeBackend, TensorSpec

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 28
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 4
dev = torch.device("cuda", 0)
be = NativeBackend(dev)
p = torch.randn(n, device=dev) * 0.02
g = torch.randn(n, device=dev) * 1e-3
m = torch.zeros(n, device=dev)
v = torch.zeros(n, device=dev)
plan = be.make_plan([TensorSpec(p=p, g=g, m=m, v=v, group=0)])
hp = dict(lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=1e-2, decoupled=True)
ts = []
for i in range(iters):
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    e[0].record()
    norm = be.l2norm(plan, 1.0)
    e[1].record()
    be.adam_step(plan, [dict(hp, step=i + 1)], norm, 1.0, 1.0, True)
    e[2].record()
    torch.cuda.synchronize()
    ts.append((e[0].elapsed_time(e[1]), e[1].elapsed_time(e[2])))
tn, tk = ts[-1]
print(f"n={n} norm {tn:.4f} ms = {4*n/tn/1e6:.1f} GB/s | step {tk:.4f} ms = {32*n/tk/1e6:.1f} GB/s")
