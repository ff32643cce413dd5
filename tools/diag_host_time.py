"""Diagnostic (GPU): where does host time go per Looper iteration? Times every capsule's launch."""
import os, sys, time, collections
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
os.environ["TQDM_DISABLE"] = "1"
import torch
import rocket_b200 as rk
import workloads, bench

workers = int(sys.argv[1]) if len(sys.argv) > 1 else 0
acc = collections.defaultdict(float); cnt = collections.Counter()
def wrap(cls, name="launch"):
    orig = getattr(cls, name)
    def timed(self, attrs=None):
        t = time.perf_counter(); r = orig(self, attrs); dt = time.perf_counter() - t
        acc[cls.__name__] += dt; cnt[cls.__name__] += 1; return r
    setattr(cls, name, timed)
for c in (rk.Dataset, rk.Module, rk.Loss, rk.Optimizer):
    wrap(c)
from rocket_b200 import staging
wrap2 = lambda cls, name: (lambda orig: setattr(cls, name, (lambda self, *a, **k: _t(cls.__name__ + "." + name, orig, self, *a, **k))))(getattr(cls, name))
def _t(key, orig, *a, **k):
    t = time.perf_counter(); r = orig(*a, **k); acc[key] += time.perf_counter() - t; cnt[key] += 1; return r
for cls, name in ((staging.PinnedRingStager, "submit"), (staging.PinnedRingStager, "collect"),
                  (staging.PinnedRingStager, "release"), (staging._Reader, "get")):
    wrap2(cls, name)
SKIP = 30
class Reset(rk.Capsule):
    def __init__(self): super().__init__(priority=1); self.i = 0
    def launch(self, attrs=None):
        self.i += 1
        if self.i == SKIP: acc.clear(); cnt.clear()
torch.backends.cudnn.benchmark = True
N = 120
class PoolDataset(torch.utils.data.Dataset):
    def __init__(self, pool, length):
        g = torch.Generator().manual_seed(99)
        self.x = torch.randn(pool, 3, 32, 32, generator=g); self.y = torch.randint(0, 10, (pool,), generator=g); self.length = length
    def __len__(self): return self.length
    def __getitem__(self, i):
        j = i % self.x.shape[0]; return self.x[j], self.y[j]
net = bench.make_resnet18()
opt = torch.optim.AdamW(net.parameters(), lr=1e-3, weight_decay=1e-2)
ds = PoolDataset(16 * 256, N * 256)
kw = dict(num_workers=workers, prefetch_factor=4) if workers else {}
looper = rk.Looper([rk.Dataset(ds, batch_size=256, shuffle=False, **kw),
                    rk.Module(net, capsules=[rk.Loss(objective=workloads.CrossEntropy()), rk.Optimizer(opt, max_grad_norm=1.0)]), Reset()], tag="diag")
t0 = time.perf_counter()
rk.Launcher([looper], num_epochs=1, destroy_process_group_after_launch=False).launch()
torch.cuda.synchronize()
print("workers", workers, "total", round((time.perf_counter() - t0) * 1e3 / N, 3), "ms/iter incl. setup")
for k in acc: print(f"  {k:10s} {acc[k] / cnt[k] * 1e3:.3f} ms/iter (Module includes Loss+Optimizer)")
