"""Tuning sweep (GPU): GB/s of the fused AdamW kernel per RK_STEP_VARIANT, one process each."""
import os, subprocess, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
n = sys.argv[1] if len(sys.argv) > 1 else str(1 << 30)
for v in (sys.argv[2].split(",") if len(sys.argv) > 2 else ["0", "1", "2", "4", "9", "12"]):
    env = dict(os.environ, RK_STEP_VARIANT=v)
    out = subprocess.run([sys.executable, str(ROOT / "tools" / "profile_optstep.py"), n, "8"],
                         env=env, capture_output=True, text=True)
    print("variant", v, out.stdout.strip() or out.stderr[-300:], flush=True)
