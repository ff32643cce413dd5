"""Freeze the reference's public signatures (build container only: needs /root/reference).

    python tests/golden/make_api_signatures.py   ->  tests/golden/api_signatures.json
"""
import inspect
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
import oracle  # noqa: E402

CLASSES = ["Capsule", "Dispatcher", "Launcher", "Looper", "Dataset", "Module", "Loss",
           "Optimizer", "Scheduler", "Meter", "Metric", "Tracker", "Checkpointer"]


def describe(cls) -> dict:
    out = {}
    for name, member in inspect.getmembers(cls):
        if name.startswith("__") and name != "__init__":
            continue
        if inspect.isfunction(member) or inspect.ismethod(member):
            sig = inspect.signature(member)
            out[name] = [[p.name, str(p.kind), None if p.default is p.empty else repr(p.default)]
                         for p in sig.parameters.values()]
    return out


def main():
    rk = oracle.import_reference()
    data = {c: describe(getattr(rk, c)) for c in CLASSES}
    from rocket.core.capsule import Events
    data["Events"] = {e.name: e.value for e in Events}
    import rocket.utils.torch as rt
    import rocket.utils.collections as rc
    data["utils.torch"] = {n: [p.name for p in inspect.signature(getattr(rt, n)).parameters.values()]
                           for n in ("torch_collate", "torch_move", "register_move_hook",
                                     "register_default_move_hook", "move")}
    data["utils.collections"] = {n: [p.name for p in inspect.signature(getattr(rc, n)).parameters.values()]
                                 for n in ("apply_to_collection", "is_collection")}
    path = Path(__file__).resolve().parent / "api_signatures.json"
    path.write_text(json.dumps(data, indent=1, sort_keys=True))
    print("wrote", path, sum(len(v) for v in data.values()), "entries")


if __name__ == "__main__":
    main()
