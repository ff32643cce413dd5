"""Diagnostic (GPU): GPU-busy time vs wall time per ResNet-18 step through the capsule API."""
import os, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
os.environ["TQDM_DISABLE"] = "1"
import torch
from torch.profiler import ProfilerActivity, profile
import rocket_b200 as rk
import workloads, bench
torch.backends.cudnn.benchmark = True
net = bench.make_resnet18()
opt = torch.optim.AdamW(net.parameters(), lr=1e-3, weight_decay=1e-2)
batch = [t.cuda() for t in bench.synthetic_batch(0)]
class Resident(rk.Capsule):
    def __init__(self): super().__init__(priority=2000)
    def launch(self, attrs=None):
        attrs.batch = list(batch); attrs.looper.terminate = False
N = 30
prof_holder = {}
class Prof(rk.Capsule):
    def __init__(self): super().__init__(priority=0); self.i = 0
    def launch(self, attrs=None):
        self.i += 1
        if self.i == 10:
            torch.cuda.synchronize()
            prof_holder["p"] = profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]); prof_holder["p"].__enter__()
        if self.i == 10 + N:
            torch.cuda.synchronize(); prof_holder["p"].__exit__(None, None, None)
looper = rk.Looper([Resident(), rk.Module(net, capsules=[rk.Loss(objective=workloads.CrossEntropy()), rk.Optimizer(opt, max_grad_norm=1.0)]), Prof()], tag="d", repeats=10 + N)
rk.Launcher([looper], num_epochs=1, destroy_process_group_after_launch=False).launch()
p = prof_holder["p"]
evs = [e for e in p.events() if e.device_type == torch.autograd.DeviceType.CUDA]
busy = sum(e.device_time for e in evs) / 1e3 / N
t0 = min(e.time_range.start for e in evs); t1 = max(e.time_range.end for e in evs)
print(f"GPU kernels: {len(evs)/N:.0f}/step, busy {busy:.3f} ms/step, span {(t1-t0)/1e3/N:.3f} ms/step")
top = {}
for e in evs: top[e.name[:60]] = top.get(e.name[:60], 0) + e.device_time
for k, v in sorted(top.items(), key=lambda kv: -kv[1])[:8]: print(f"  {v/1e3/N:.3f} ms/step  {k}")
