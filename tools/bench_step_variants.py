"""GPU tuning sweep: every code shape of the fused AdamW step (K2) on the tensor lists the
configs use, timed with CUDA events.

    python tools/bench_step_variants.py [--layouts flat1e9,gpt2x8,gpt2,resnet18] [--variants 4,9,...]
                                        [--iters 12] [--out gpurun_out/step_variants.jsonl]

Layouts: `flat1e9` one 1e9-element tensor; `gpt2x8` GPT-2-small's 148 shapes x 8
plus a pad tensor up to 1e9 elements (the realistic configs[4] list); `gpt2` 124M; `resnet18`
the 62 tensors of torchvision's resnet18(num_classes=10).  Lists under 2^27 elements fit partly
in L2, so a 512 MB buffer is rewritten between iterations and the norm kernel runs right before
the step, as in a real step (gradients L2-warm, everything else cold).
GB/s = 32 B/param (AdamW 28 + grad zero-fill 4) / event time of the step kernel alone.
"""
from __future__ import annotations

import argparse
import json
import statistics
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch  # noqa: E402

from rocket_b200 import _native as N  # noqa: E402
from rocket_b200.backend import NativeBackend, TensorSpec  # noqa: E402


def gpt2_shapes():
    s = [50257 * 768, 1024 * 768]
    for _ in range(12):
        s += [768, 768, 768 * 2304, 2304, 768 * 768, 768, 768, 768, 768 * 3072, 3072,
              3072 * 768, 768]
    return s + [768, 768]


def resnet18_shapes():
    import torchvision
    return [p.numel() for p in torchvision.models.resnet18(num_classes=10).parameters()]


def layout(name: str):
    if name == "flat1e9":
        return [1_000_000_000]
    if name == "gpt2":
        return gpt2_shapes()
    if name == "gpt2x8":
        s = gpt2_shapes() * 8
        return s + [1_000_000_000 - sum(s)]
    if name == "resnet18":
        return resnet18_shapes()
    raise SystemExit(f"unknown layout {name}")


def build(sizes, dev):
    """Four flat streams with 256-byte aligned segments, like rocket_b200/flat.py."""
    offs, off = [], 0
    for n in sizes:
        offs.append(off)
        off += (n + 63) // 64 * 64
    flat = {k: torch.zeros(off, device=dev) for k in "pgmv"}
    flat["p"].normal_(0, 0.02)
    specs = [TensorSpec(*[flat[k][o:o + n] for k in "pgmv"], group=0)
             for o, n in zip(offs, sizes)]
    return flat, specs


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--layouts", default="flat1e9,gpt2x8,gpt2,resnet18")
    ap.add_argument("--variants", default="4,9,30,31,33,20,21,22,23,24")
    ap.add_argument("--iters", type=int, default=12)
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    be = NativeBackend(dev)
    lib = N.load()
    hp = dict(lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=1e-2, decoupled=True)
    flush = torch.empty(128 << 20, device=dev)            # 512 MB > 126 MB of L2
    rows = []
    for name in args.layouts.split(","):
        sizes = layout(name)
        n = sum(sizes)
        flat, specs = build(sizes, dev)
        plan = be.make_plan(specs)
        small = n < (1 << 27)
        for v in [int(x) for x in args.variants.split(",")]:
            lib.rk_step_variant_override(v)
            tn, ts = [], []
            for i in range(args.iters + 3):
                flat["g"].normal_(0, 1e-3)
                if small:
                    flush.fill_(float(i))
                e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
                e[0].record()
                norm = be.l2norm(plan, 1.0)
                e[1].record()
                be.adam_step(plan, [dict(hp, step=i + 1)], norm, 1.0, 1.0, True)
                e[2].record()
                torch.cuda.synchronize()
                if i >= 3:
                    tn.append(e[0].elapsed_time(e[1]))
                    ts.append(e[1].elapsed_time(e[2]))
            # the pair with no event between the two launches (an event record in between
            # would turn a programmatic dependent launch back into a full dependency)
            tp = []
            for i in range(args.iters + 3):
                flat["g"].normal_(0, 1e-3)
                if small:
                    flush.fill_(float(i))
                e = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
                e[0].record()
                norm = be.l2norm(plan, 1.0)
                be.adam_step(plan, [dict(hp, step=i + 1)], norm, 1.0, 1.0, True)
                e[1].record()
                torch.cuda.synchronize()
                if i >= 3:
                    tp.append(e[0].elapsed_time(e[1]))
            t_step, t_norm = statistics.mean(ts), statistics.mean(tn)
            row = {"layout": name, "params": n, "tensors": len(sizes), "variant": v,
                   "step_ms": round(t_step, 5), "step_ms_min": round(min(ts), 5),
                   "norm_ms": round(t_norm, 5),
                   "step_gbs": round(32 * n / t_step / 1e6, 1),
                   "step_gbs_best": round(32 * n / min(ts) / 1e6, 1),
                   "full_gbs_36B": round(36 * n / (t_step + t_norm) / 1e6, 1),
                   "pair_ms": round(statistics.mean(tp), 5),
                   "pair_gbs_36B": round(36 * n / statistics.mean(tp) / 1e6, 1)}
            rows.append(row)
            print(json.dumps(row), flush=True)
        lib.rk_step_variant_override(-1)
        del flat, specs, plan
        torch.cuda.empty_cache()
    if args.out:
        Path(args.out).parent.mkdir(parents=True, exist_ok=True)
        Path(args.out).write_text("".join(json.dumps(r) + "\n" for r in rows))


if __name__ == "__main__":
    main()
