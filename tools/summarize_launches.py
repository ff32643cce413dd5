"""Turn an ncu launch list (`ncu --metrics gpu__time_duration.sum --csv --log-file X ...`) of a
bench.py run into per-kernel shares of a steady-state step.

    python tools/summarize_launches.py gpurun_out/launches_eager.csv [n_steps] > profiles/...csv

A step is the span between two consecutive launches of the fused step kernel (rk_step_kernel);
the LAST `n_steps` complete spans of the first leg with that many identical spans are used.
"""
from __future__ import annotations

import csv
import sys
from collections import defaultdict


def main() -> None:
    path = sys.argv[1]
    n_steps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    rows = []
    with open(path, newline="") as f:
        lines = [ln for ln in f if ln.startswith('"')]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        try:
            ns = float(r["Metric Value"].replace(",", ""))
        except ValueError:
            continue
        rows.append((r["Kernel Name"], ns))
    marks = [i for i, (k, _) in enumerate(rows) if "rk_step_kernel" in k or "rk_step_bulk_kernel" in k]
    spans = [(a + 1, b + 1) for a, b in zip(marks, marks[1:])]       # (after step i, through step i+1]
    sizes = [b - a for a, b in spans]
    # steady state: the last run of n_steps spans with the same number of launches
    best = None
    for end in range(len(spans), n_steps - 1, -1):
        if len(set(sizes[end - n_steps:end])) == 1:
            best = spans[end - n_steps:end]
            break
    if best is None:
        raise SystemExit(f"no {n_steps} identical consecutive steps found: {sizes}")
    per = defaultdict(lambda: [0, 0.0])
    total = 0.0
    for a, b in best:
        for k, ns in rows[a:b]:
            per[k][0] += 1
            per[k][1] += ns
            total += ns
    launches = sum(v[0] for v in per.values())
    print(f"# {launches} launches in {n_steps} steps ({launches / n_steps:.0f}/step), "
          f"device time {total / n_steps / 1e6:.3f} ms/step")
    print("share_pct,launches_per_step,avg_us,kernel")
    for k, (n, ns) in sorted(per.items(), key=lambda kv: -kv[1][1]):
        print(f"{100 * ns / total:.2f},{n / n_steps:.1f},{ns / n / 1e3:.2f},{k[:110]}")


if __name__ == "__main__":
    main()
