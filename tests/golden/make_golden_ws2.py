"""Golden vectors for the 2-rank data-parallel path: the UNMODIFIED reference under real torch
DDP over gloo (CPU), two processes.  Build container only.

    python tests/golden/make_golden_ws2.py
"""
import sys
from pathlib import Path

import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT / "tests"))
import dist_worker  # noqa: E402

if __name__ == "__main__":
    here = str(Path(__file__).resolve().parent)
    only = set(sys.argv[1:])              # optional: regenerate just the named cases
    for i, case in enumerate(dist_worker.ALL_CASES):
        if only and case not in only:
            continue
        mp.spawn(dist_worker.reference_worker, args=(2, 29610 + i, case, here), nprocs=2, join=True)
        print("wrote", case)
