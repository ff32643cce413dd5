"""Where the norm kernel's time goes: the whole kernel (level 1 + ticket + level 2 in the last
CTA) against level 2 alone (rk_l2norm_finalize: one CTA over the same partials)."""
import statistics
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
sys.path.insert(0, str(Path(__file__).resolve().parent))
import torch  # noqa: E402

from rocket_b200.backend import NativeBackend  # noqa: E402
from bench_step_variants import build, layout  # noqa: E402

dev = torch.device("cuda", 0)
be = NativeBackend(dev)
flush = torch.empty(128 << 20, device=dev)
for name in (sys.argv[1].split(",") if len(sys.argv) > 1 else ["resnet18", "gpt2", "gpt2x8", "flat1e9"]):
    sizes = layout(name)
    flat, specs = build(sizes, dev)
    flat["g"].normal_(0, 1e-3)
    plan = be.make_plan(specs)
    full, fin = [], []
    for i in range(12):
        flush.fill_(float(i))
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        e[0].record()
        be.l2norm(plan, 1.0)
        e[1].record()
        be.l2norm_finalize(plan)
        e[2].record()
        torch.cuda.synchronize()
        if i >= 2:
            full.append(e[0].elapsed_time(e[1]))
            fin.append(e[1].elapsed_time(e[2]))
    n = sum(sizes)
    print(f"{name}: tensors {len(sizes)} chunks {plan.n_chunks_norm} chunk {plan.chunk_norm} | "
          f"norm kernel {statistics.mean(full)*1e3:.1f} us ({4*n/statistics.mean(full)/1e6:.0f} GB/s) | "
          f"level 2 alone {statistics.mean(fin)*1e3:.1f} us", flush=True)
    del flat, specs, plan
    torch.cuda.empty_cache()
