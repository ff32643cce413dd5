#!/usr/bin/env python
"""bench.py -- rocket's per-iteration training step on B200: samples/sec + fused-step GB/s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Headline workload (BASELINE.json configs[1]): ResNet-18 (num_classes=10) on synthetic
CIFAR-10-shaped 3x32x32 batches, bs=256 per GPU, AdamW(lr=1e-3, wd=1e-2) + global-norm clip
1.0, true fp32 (TF32 off, as in the parity tests), driven through the unchanged capsule API
(Launcher > Looper > Dataset/Module/Loss/Optimizer).  One "step" is one Looper iteration.

Prints ONE JSON line (rank 0):
  value         train samples/s, whole job, batch already resident in HBM when timing starts
  e2e           same metric through the public API with HOST batches: pinned-ring H2D staging
                and the loss read back to the host are inside the timed region
  roofline      the fused AdamW step kernel: algorithmic 32 B/param / CUDA-event duration
                measured live in the timed region, against MEASURED_PEAKS.json
  eager         the same workload without the additive Module(graph=True) keyword: what an
                unchanged reference script gets
  torch_gpu     baseline B: stock torch on the same GPU(s), no capsules: AdamW(fused=True) +
                clip_grad_norm_(foreach=True) + zero_grad(set_to_none=True), DDP at N > 1
  resnet50      configs[2] (ResNet-50, 3x224x224, bs 256/GPU, bf16 autocast), value + e2e
  gpt2          configs[3] (GPT-2 small, seq 1024, micro-batch 8, accum 4), at 8 GPUs
  optstep       configs[4] microbench (fused norm+AdamW+zero over 1e9 fp32 params): one flat
                tensor and the realistic list (GPT-2's 148 shapes x 8 + pad), and stock torch
                (clip_grad_norm_ foreach + AdamW fused) on the same tensors; at N > 1 the
                data-parallel step (fused NVLink exchange + step) replicated and sharded
  parity_check  N > 1 only, before anything is timed: the exchange kernel against an fp64 host
                sum and the oracle norm, the sharded step against the replicated one bit for
                bit, and (after the timed run) replicas bit-identical; any mismatch exits 1
  cpu_baseline  the looper port (torch.optim foreach=False) on this box's host cores,
                bounded sample

--impl reference times that CPU port alone, with all host threads (the reference itself is
pure Python + torch and is not present on the GPU box; oracle/looper_port.py restates it).
"""
from __future__ import annotations

import argparse
import importlib.util
import json
import os
import statistics
import subprocess
import sys
import threading
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
os.environ.setdefault("TQDM_DISABLE", "1")

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

BATCH = 256
METRIC = "train samples/sec (ResNet-18 CIFAR-shaped bs=256/GPU, AdamW+clip) + fused-step HBM GB/s"
ALGO_BYTES_STEP = 32      # AdamW update 28 B + grad zero-fill 4 B per fp32 param (K2)
ALGO_BYTES_NORM = 4       # K1
STEP_KERNEL = "rk_step_bulk_kernel<AdamOp,AdamTable,16>"   # what the library picks by default


def measured_traffic(kernel: str, workload: str):
    """DRAM bytes per launch of `kernel` inside `workload`'s step, from the committed ncu
    --set full capture of exactly that kernel (profiles/r2_roofline.json); None if there is no
    capture of it on record (or a tuning override is active: the capture would not apply)."""
    if os.environ.get("RK_STEP_VARIANT"):
        return None
    try:
        rec = json.loads((ROOT / "profiles" / "r2_roofline.json").read_text())
        for e in rec["captures"]:
            if e["kernel"] == kernel and e["workload"] == workload:
                return e["dram_bytes_per_launch"]
    except Exception:
        pass
    return None


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons, sampled during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self._stop, self._t = index, [], threading.Event(), None

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(
                    ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                     "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([c.strip() for c in out.strip().split(",")])
            except Exception:
                pass
            self._stop.wait(0.2)

    def start(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()

    def stop(self) -> dict:
        self._stop.set()
        if self._t:
            self._t.join(timeout=6)
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 6
                          for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": statistics.median(sm) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.rows)}


# ---- workloads -----------------------------------------------------------------------------------
def make_resnet18():
    import torchvision
    import workloads
    torch.manual_seed(0)
    return workloads.BatchModel(torchvision.models.resnet18(num_classes=10))


def make_resnet50():
    import torchvision
    import workloads
    torch.manual_seed(0)
    return workloads.BatchModel(torchvision.models.resnet50())


def make_gpt2():
    from transformers import GPT2Config, GPT2LMHeadModel
    torch.manual_seed(0)
    lm = GPT2LMHeadModel(GPT2Config(resid_pdrop=0.0, embd_pdrop=0.0, attn_pdrop=0.0))

    class GPT(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.lm = lm

        def forward(self, batch):
            ids = batch[0]
            logits = self.lm(input_ids=ids).logits
            return ids, ids, logits      # (inputs, targets, logits) like the image workloads
    return GPT()


class NextTokenLoss(torch.nn.Module):
    def forward(self, batch):
        ids, logits = batch[0], batch[2]
        return torch.nn.functional.cross_entropy(
            logits[:, :-1].reshape(-1, logits.shape[-1]).float(), ids[:, 1:].reshape(-1))


# `resnet18` is the driver's headline workload (configs[1]); `resnet50` is configs[2]
# (ImageNet-shaped, bf16 autocast, SGD momentum), `gpt2` configs[3].
WORKLOADS = {
    "gpt2": dict(
        net=make_gpt2, shape=(1024,), classes=50257, mixed_precision="bf16", max_norm=1.0,
        batch=8, accum=4, tokens=True, algo=ALGO_BYTES_STEP,
        opt=lambda ps: torch.optim.AdamW(ps, lr=6e-4, betas=(0.9, 0.95), weight_decay=0.1),
        label="configs[3]: GPT-2 small 124M (GPT2LMHeadModel), synthetic tokens seq 1024, "
              "micro-batch 8/GPU, grad-accum 4, AdamW(6e-4,(0.9,0.95),wd 0.1)+clip 1.0, bf16 "
              "autocast, capsule API", dtype="bf16 autocast / f32 master"),
    "resnet18": dict(
        net=make_resnet18, shape=(3, 32, 32), classes=10, mixed_precision=None, max_norm=1.0,
        algo=ALGO_BYTES_STEP,
        opt=lambda ps: torch.optim.AdamW(ps, lr=1e-3, weight_decay=1e-2),
        label="configs[1]: ResNet-18(num_classes=10), synthetic 3x32x32, bs=256/GPU, "
              "AdamW(lr=1e-3,wd=1e-2)+clip 1.0, fp32, capsule API", dtype="f32"),
    "resnet50": dict(
        net=make_resnet50, shape=(3, 224, 224), classes=1000, mixed_precision="bf16",
        max_norm=None, algo=24,                                   # SGD+momentum: 20 + 4
        opt=lambda ps: torch.optim.SGD(ps, lr=0.1, momentum=0.9, weight_decay=1e-4),
        label="configs[2]: ResNet-50, synthetic 3x224x224, bs=256/GPU, bf16 autocast, "
              "SGD(momentum 0.9, wd 1e-4), capsule API", dtype="bf16 autocast / f32 master"),
}


def set_precision(wl, tf32_conv: bool = False) -> None:
    """The line whose dtype says f32 computes in true fp32: TF32 off for convolutions and
    matmuls, which is what every parity test runs (tests/test_engine_gpu.py) and what the
    reference arm's host CPU computes.  (torch's own default leaves cuDNN TF32 convolutions
    on; `tf32_conv=True` is that setting, reported separately as `tf32_conv`.)  The bf16
    workloads autocast; whatever stays fp32 there keeps torch's defaults."""
    fp32 = wl["mixed_precision"] is None and not tf32_conv
    torch.backends.cudnn.allow_tf32 = not fp32
    torch.backends.cuda.matmul.allow_tf32 = wl["mixed_precision"] is not None
    torch.backends.cudnn.benchmark = True


def synthetic_batch(wl, rank: int, n: int | None = None):
    n = n or wl.get("batch", BATCH)
    g = torch.Generator().manual_seed(1234 + rank)
    if wl.get("tokens"):
        ids = torch.randint(0, wl["classes"], (n, *wl["shape"]), generator=g)
        return ids, ids[:, 0].clone()
    return (torch.randn(n, *wl["shape"], generator=g),
            torch.randint(0, wl["classes"], (n,), generator=g))


def env_ranks():
    return (int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)),
            int(os.environ.get("LOCAL_RANK", 0)))


def run_capsule(wl, *, graph: bool, steps: int, warmup: int, e2e: bool, loader_workers: int = 0,
                tf32_conv: bool = False):
    """One Launcher.launch() of the capsule graph on `wl`: `warmup` untimed iterations, then
    exactly `steps` timed ones between two barrier + device-sync fences (CUDA events, max over
    ranks).  e2e=False: one batch resident in HBM; e2e=True: host batches through
    rocket.Dataset + the pinned-ring stager, every step's loss read back to the host."""
    import rocket_b200 as rk
    import workloads
    from rocket_b200 import _native
    from rocket_b200.backend import NativeBackend
    from rocket_b200.engine import Engine

    rank, world, local = env_ranks()
    W, K = warmup, steps
    accum = wl.get("accum", 1)
    TAIL = 8 if graph else 0        # synchronised replays after the timed region (kernel timing)
    if graph:
        # Module(graph=True) runs GRAPH_WARMUP eager iterations per input signature and then
        # captures: with a short --warmup the capture would land inside the timed region.
        # Extra untimed steps keep it out; the line reports the warm-up actually done.
        W = max(W, (rk.Module.GRAPH_WARMUP + 2) * accum)
    set_precision(wl, tf32_conv)

    class TimedBackend(NativeBackend):
        """Brackets each engine kernel launch with CUDA events on its own stream.  While a
        step is being captured into a CUDA graph the events are `external` ones, i.e. event
        record nodes of the graph: every replay re-records them and the host samples the pair
        whenever it is complete and consistent."""
        def __init__(self, device):
            super().__init__(device)
            self.timing = False
            self.engine = None
            self.ev = {"norm": [], "step": []}
            self.graph_ev = {}
            self.graph_ms = {"norm": [], "step": []}

        def sample_graph_events(self):
            for key, (s, e) in self.graph_ev.items():
                try:
                    if e.query():
                        dt = s.elapsed_time(e)
                        if 0.0 < dt < 20.0:          # a torn pair (replay in flight) is negative
                            self.graph_ms.setdefault(key, []).append(dt)
                except Exception:
                    pass

        def _timed(self, key, fn, *a, **k):
            gp = getattr(self.engine, "graph_pass", None)
            if gp is not None and gp.mode == "capture":
                s = torch.cuda.Event(enable_timing=True, external=True)
                e = torch.cuda.Event(enable_timing=True, external=True)
                s.record()
                out = fn(*a, **k)
                e.record()
                self.graph_ev[key] = (s, e)
                return out
            if not self.timing:
                return fn(*a, **k)
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            out = fn(*a, **k)
            e.record()
            self.ev[key].append((s, e))
            return out

        def l2norm(self, *a, **k):
            return self._timed("norm", super().l2norm, *a, **k)

        def adam_step(self, *a, **k):
            return self._timed("step", super().adam_step, *a, **k)

        def sgd_step(self, *a, **k):
            return self._timed("step", super().sgd_step, *a, **k)

        def step_devtable(self, *a, **k):
            return self._timed("step", super().step_devtable, *a, **k)

        def l2norm_finalize(self, *a, **k):
            return self._timed("norm", super().l2norm_finalize, *a, **k)

    holder = {}

    def factory(**kw):
        dev = torch.device("cuda", local)
        torch.cuda.set_device(dev)
        eng = Engine(backend=TimedBackend(dev), device=dev, **kw)
        eng.backend.engine = eng
        holder["engine"] = eng
        return eng

    rk.Launcher.engine_factory = staticmethod(factory)
    from rocket_b200.comm import P2PGradSync
    P2PGradSync.graph_timing = True      # event pair around the exchange inside captured steps

    class ResidentBatch(rk.Capsule):
        """Hands the Looper one batch that already lives in HBM (the `value` leg)."""
        def __init__(self, batch):
            super().__init__(priority=2000)
            self._host, self._dev = batch, None

        def setup(self, attrs=None):
            super().setup(attrs)
            self._dev = [t.to(self._accelerator.device) for t in self._host]

        def launch(self, attrs=None):
            attrs.batch = list(self._dev)
            attrs.looper.terminate = False

    class Probe(rk.Capsule):
        """Runs last in every iteration; opens the timed region after W steps and closes it
        after W+K, with a barrier + device sync on both sides (CUDA events, device time)."""
        def __init__(self, read_loss: bool):
            super().__init__(priority=0)
            self.i, self.read_loss = 0, read_loss
            self.t0 = torch.cuda.Event(enable_timing=True)
            self.t1 = torch.cuda.Event(enable_timing=True)
            self.ms = self.launches = None
            self.losses = []
            self.exchange = None
            self.sync = None
            self.sampler = ClockSampler(local)

        def _fence(self):
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
                torch.cuda.synchronize()

        def launch(self, attrs=None):
            self.i += 1
            be = self._accelerator.backend
            if self.i > W and self.read_loss and attrs.looper.state.loss is not None:
                self.losses.append(attrs.looper.state.loss)
            if self.i == W:
                self._fence()
                self.launch0 = _native.launch_count() + self._accelerator.replayed_kernel_launches
                be.timing = True
                be.graph_ms = {"norm": [], "step": []}
                for sync in self._accelerator._grad_sync.values():
                    if hasattr(sync, "profile"):
                        sync.profile, self.sync = True, sync
                        sync.graph_ms = []
                if rank == 0:
                    self.sampler.start()
                self.t0.record()
            elif self.i > W + K:
                # graph replays cannot be bracketed from the host while they are queued: right
                # after the timed region a few replays are synchronised one by one so that the
                # in-graph event pairs of the engine kernels can be read
                torch.cuda.synchronize()
                be.sample_graph_events()
                if self.sync is not None:
                    self.sync.sample_graph_events()
            elif self.i == W + K:
                if self.read_loss:            # the host must hold every step's loss
                    self.host_losses = [float(v) for v in self.losses]
                self.t1.record()
                self._fence()
                be.sample_graph_events()
                be.timing = False
                if self.sync is not None:
                    self.sync.sample_graph_events()
                    self.sync.profile = False
                    self.exchange = [(nb, a.elapsed_time(b)) for nb, a, b in self.sync.timings]
                self.launches = _native.launch_count() \
                    + self._accelerator.replayed_kernel_launches - self.launch0
                self.clocks = self.sampler.stop() if rank == 0 else None
                self.ms = self.t0.elapsed_time(self.t1)

    class PoolDataset(torch.utils.data.Dataset):
        """(W+K)*world batches of host samples drawn from a fixed synthetic pool."""
        def __init__(self, pool: int, length: int):
            g = torch.Generator().manual_seed(99)
            if wl.get("tokens"):
                self.x = torch.randint(0, wl["classes"], (pool, *wl["shape"]), generator=g)
                self.y = self.x[:, 0].clone()
            else:
                self.x = torch.randn(pool, *wl["shape"], generator=g)
                self.y = torch.randint(0, wl["classes"], (pool,), generator=g)
            self.length = length

        def __len__(self):
            return self.length

        def __getitem__(self, i):
            j = i % self.x.shape[0]
            return self.x[j], self.y[j]

    B = wl.get("batch", BATCH)
    probe = Probe(read_loss=e2e)
    if e2e:
        pool = 16 * B if wl["shape"][-1] <= 64 or wl.get("tokens") else 2 * B
        host_ds = PoolDataset(pool=pool, length=(W + K + TAIL) * B * world)
        # an ordinary input pipeline: DataLoader kwargs pass straight through rocket.Dataset
        data = rk.Dataset(host_ds, batch_size=B, shuffle=False, num_workers=loader_workers,
                          prefetch_factor=4 if loader_workers else None)
    else:
        data = ResidentBatch(synthetic_batch(wl, rank))
    net = wl["net"]()
    opt = wl["opt"](net.parameters())
    objective = NextTokenLoss() if wl.get("tokens") else workloads.CrossEntropy()
    inner = [rk.Loss(objective=objective), rk.Optimizer(opt, max_grad_norm=wl["max_norm"])]
    looper = rk.Looper([data, rk.Module(net, capsules=inner, graph=bool(graph)), probe],
                       tag="bench", repeats=W + K + TAIL)
    launcher = rk.Launcher([looper], num_epochs=1, mixed_precision=wl["mixed_precision"],
                           gradient_accumulation_steps=accum,
                           destroy_process_group_after_launch=False)
    launcher.launch()

    be = holder["engine"].backend
    params = list(opt.param_groups[0]["params"])
    n_params = sum(p.numel() for p in params)
    step_ms = [s.elapsed_time(e) for s, e in be.ev["step"]] + be.graph_ms.get("step", [])
    norm_ms = [s.elapsed_time(e) for s, e in be.ev["norm"]] + be.graph_ms.get("norm", [])
    ms = torch.tensor([probe.ms], device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())
    res = {"ms_total": ms_total, "ms_per_step": ms_total / K, "steps": K, "warmup": W,
           "value": world * B * K / (ms_total / 1e3), "launches": probe.launches,
           "clocks": probe.clocks, "step_ms": step_ms, "norm_ms": norm_ms,
           "n_params": n_params, "batch": B, "accum": accum,
           "grad_sync": (holder["engine"].grad_sync_kinds or ["none"])[0],
           "exchange": probe.exchange,
           "exchange_graph_ms": list(getattr(probe.sync, "graph_ms", []) or []),
           "exchange_graph_bytes": getattr(probe.sync, "graph_bytes", 0)}
    if world > 1:
        import parity_checks
        # parameters only: BatchNorm running statistics are rank-local between the broadcasts
        # in front of each forward (as under DDP), and do not feed the training-mode forward
        res["replicas_identical"] = parity_checks.replicas_identical(
            params, torch.device("cuda", local))
    if e2e:
        sample_bytes = 8 * wl["shape"][0] if wl.get("tokens") else \
            4 * wl["shape"][0] * wl["shape"][1] * wl["shape"][2]
        res["h2d_bytes_per_step"] = B * (sample_bytes + 8)
        res["losses_read"] = len(getattr(probe, "host_losses", []))
    del launcher, looper, net, opt, holder
    torch.cuda.empty_cache()
    return res


def exchange_summary(res, world: int):
    """K6 as timed inside the run: eager mode brackets every bucket's kernel with events on
    the side stream; graph mode records external event pairs inside the captured step."""
    K = res["steps"]
    if res.get("exchange"):
        nbytes = sum(nb for nb, _ in res["exchange"])
        ms_x = sum(t for _, t in res["exchange"])
        per_step_ms, per_step_bytes = ms_x / K, nbytes // K
        launches = round(len(res["exchange"]) / K, 2)
        how = "CUDA events around every bucket's kernel on the exchange stream (eager step)"
    elif res.get("exchange_graph_ms"):
        per_step_ms = statistics.mean(res["exchange_graph_ms"])
        per_step_bytes = res["exchange_graph_bytes"]
        launches = None
        how = ("external CUDA event pairs around every bucket's kernel inside the replayed graph "
               "(launched from the gradient hooks during capture: they run beside the rest of "
               "backward), summed; read for the last replay of the timed region and 8 "
               "synchronised replays after it")
    else:
        return None
    bus = 2 * (world - 1) / world * per_step_bytes / (per_step_ms * 1e-3) / 1e9
    return {"kernel": "rk_allreduce_norm_kernel (NVLink all-reduce + 1/W + norm partials)",
            "launches_per_step": launches, "bytes_per_step": per_step_bytes,
            "kernel_ms_per_step": round(per_step_ms, 4), "bus_gbs": round(bus, 1),
            "nvlink_peak_gbs": 770.0, "frac": round(bus / 770.0, 4), "timing": how,
            "note": "bus GB/s = 2(W-1)/W x bytes / time = bytes crossing NVLink per direction "
                    "per GPU; peak = measured peer-copy bandwidth (B200_PROFILING.md); the "
                    "span includes waiting for the slowest rank's backward"}


def capsule_object(wl, res, res_e2e, peak):
    """A secondary workload's numbers, in the same terms as the headline line."""
    k2 = statistics.mean(res["step_ms"]) if res["step_ms"] else None
    out = {"workload": wl["label"], "dtype": wl["dtype"], "value": round(res["value"], 1),
           "unit": "samples/s", "ms_per_step": round(res["ms_per_step"], 4),
           "steps": res["steps"], "warmup": res["warmup"], "params": res["n_params"],
           "global_batch": res["batch"] * res["accum"] * env_ranks()[1],
           "grad_sync": res["grad_sync"]}
    if k2:
        gbs = wl["algo"] * res["n_params"] / (k2 * 1e-3) / 1e9
        out["step_kernel_ms"] = round(k2, 5)
        out["step_kernel_gbs"] = round(gbs, 1)
        out["step_kernel_frac_of_measured_peak"] = round(gbs / peak, 4)
        out["algorithmic_bytes_per_param"] = wl["algo"]
    if res["norm_ms"]:
        out["norm_kernel_ms"] = round(statistics.mean(res["norm_ms"]), 5)
    if res_e2e is not None:
        out["e2e"] = {"value": round(res_e2e["value"], 1), "unit": "samples/s",
                      "ms_per_step": round(res_e2e["ms_per_step"], 4),
                      "h2d_bytes_per_step": res_e2e["h2d_bytes_per_step"],
                      "d2h_bytes_per_step": 8,
                      # K5: what staging a host batch costs that the step does not hide
                      "exposed_ms_per_step": round(res_e2e["ms_per_step"] - res["ms_per_step"], 4)}
    x = exchange_summary(res, env_ranks()[1])
    if x:
        out["exchange"] = x
    if "replicas_identical" in res:
        out["replicas_bit_identical"] = res["replicas_identical"]
    return out


def torch_gpu_baseline(wl, steps: int, warmup: int, tf32_conv: bool = False):
    """Baseline B: the same step with stock torch on the same GPU(s)
    and no capsules -- torch.optim.AdamW(fused=True) + clip_grad_norm_(foreach=True) +
    zero_grad(set_to_none=True), DistributedDataParallel at N > 1.  That is the strongest path
    the reference's Optimizer capsule reaches on a GPU today (rocket/core/optimizer.py:129-130
    with a fused torch optimizer).  Batch resident in HBM, eager, same precision settings."""
    import workloads
    rank, world, local = env_ranks()
    dev = torch.device("cuda", local)
    set_precision(wl, tf32_conv)
    net = wl["net"]().to(dev)
    model = net
    if world > 1:
        model = torch.nn.parallel.DistributedDataParallel(net, device_ids=[local])
    if wl["max_norm"] is not None or "AdamW" in wl["label"]:
        opt = torch.optim.AdamW(net.parameters(), lr=1e-3, weight_decay=1e-2, fused=True)
    else:
        opt = torch.optim.SGD(net.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4,
                              fused=True)
    obj = workloads.CrossEntropy()
    batch = [t.to(dev) for t in synthetic_batch(wl, rank)]
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ac = torch.autocast("cuda", dtype=torch.bfloat16) if wl["mixed_precision"] == "bf16" \
        else torch.autocast("cuda", enabled=False)
    model.train()
    for i in range(warmup + steps):
        if i == warmup:
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
                torch.cuda.synchronize()
            t0.record()
        with ac:
            loss = obj(model(batch))
        loss.backward()
        if wl["max_norm"] is not None:
            torch.nn.utils.clip_grad_norm_(net.parameters(), wl["max_norm"], foreach=True)
        opt.step()
        opt.zero_grad(set_to_none=True)
    t1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([t0.elapsed_time(t1)], device=dev)
    if world > 1:
        dist.barrier()
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())
    B = wl.get("batch", BATCH)
    del model, net, opt, batch
    torch.cuda.empty_cache()
    return {"value": round(world * B * steps / (ms_total / 1e3), 1), "unit": "samples/s",
            "ms_per_step": round(ms_total / steps, 4), "steps": steps, "warmup": warmup,
            "what": "stock torch, no capsules: torch.optim.AdamW(fused=True) + "
                    "clip_grad_norm_(foreach=True) + zero_grad(set_to_none=True)"
                    + (", DistributedDataParallel" if world > 1 else "")
                    + "; batch resident in HBM, eager launches, same precision flags"}


def parity_check(rank: int, world: int, local: int):
    """N > 1, before anything is timed (the driver's GPU-test box has one GPU, so this is
    where the 2/4/8-GPU parity evidence is produced).  Any mismatch raises."""
    import parity_checks
    from rocket_b200.backend import NativeBackend
    dev = torch.device("cuda", local)
    be = NativeBackend(dev)
    out = {"world": world}
    out["exchange_p2p"] = parity_checks.exchange_check(be, rank, world, dev, multicast=False)
    mc = parity_checks.exchange_check(be, rank, world, dev, multicast=True)
    out["exchange_multimem"] = mc if mc is not None else "no NVSwitch multicast mapping on this box"
    out["sharded_step_adamw_p2p"] = parity_checks.sharded_step_check(
        be, rank, world, dev, multicast=False, kind="adamw", compact=True)
    out["sharded_step_sgd_p2p"] = parity_checks.sharded_step_check(
        be, rank, world, dev, multicast=False, kind="sgd", compact=True)
    mcs = parity_checks.sharded_step_check(be, rank, world, dev, multicast=True, kind="adamw",
                                           compact=True)
    out["sharded_step_adamw_multimem"] = mcs if mcs is not None else \
        "no NVSwitch multicast mapping on this box"
    out["all_true"] = True          # reached only if no check raised
    return out


def run_ours(args) -> dict:
    rank, world, local = env_ranks()
    if os.environ.get("RK_SWITCH"):
        sys.setswitchinterval(float(os.environ["RK_SWITCH"]))
    assert args.warmup >= 3, "timing rules: at least 3 warm-up steps"
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1 and not dist.is_initialized():
        dist.init_process_group("nccl", device_id=dev)
    wl = WORKLOADS[args.workload]
    peak, peak_src = peaks()
    extras = args.workload == "resnet18" and not args.lean

    parity = None
    if world > 1 and not args.no_parity:
        parity = parity_check(rank, world, local)

    # ---- headline: inputs resident in HBM, then end to end from host batches ---------------------
    res = run_capsule(wl, graph=args.graph, steps=args.steps, warmup=args.warmup, e2e=False)
    res2 = run_capsule(wl, graph=args.graph, steps=args.steps, warmup=args.warmup, e2e=True,
                       loader_workers=args.loader_workers)
    K, W = res["steps"], res["warmup"]
    k2_ms = statistics.mean(res["step_ms"]) if res["step_ms"] else float("nan")
    algo = wl["algo"]
    achieved = algo * res["n_params"] / (k2_ms * 1e-3) / 1e9
    kernel = STEP_KERNEL if algo == ALGO_BYTES_STEP else STEP_KERNEL.replace("Adam", "Sgd")
    out = {
        "metric": METRIC, "value": round(res["value"], 1), "unit": "samples/s", "n_gpus": world,
        "steps": K, "warmup": W, "ms_per_step": round(res["ms_per_step"], 4),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": wl["dtype"],
        "data": "synthetic",
        "config": {"workload": wl["label"],
                   "global_batch": res["batch"] * world * res["accum"],
                   "parallelism": f"dp{world}",
                   "grad_sync": res["grad_sync"],
                   "cuda_graph_step": bool(args.graph),
                   "tf32": wl["mixed_precision"] is not None,     # off for the f32 workload
                   "l2": "no explicit flush: per-step working set (45 MB x4 optimizer streams "
                         "+ activations of bs 256) exceeds the 126 MB L2",
                   "params": res["n_params"],
                   "accelerate_installed": importlib.util.find_spec("accelerate") is not None},
        "e2e": {"value": round(res2["value"], 1), "unit": "samples/s",
                "h2d_bytes_per_step": res2["h2d_bytes_per_step"], "d2h_bytes_per_step": 8,
                "ms_per_step": round(res2["ms_per_step"], 4),
                "exposed_ms_per_step": round(res2["ms_per_step"] - res["ms_per_step"], 4),
                "loader_workers": args.loader_workers},
        "gpu_launches": res["launches"],
        "clocks": res["clocks"],
        "roofline": {"bound": "hbm", "kernel": kernel + " (fused AdamW+clip+zero)",
                     "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                     "frac": round(achieved / peak, 4),
                     "traffic": measured_traffic(kernel, args.workload),
                     "peak_source": peak_src, "algorithmic_bytes_per_param": algo,
                     "launch_ms": round(k2_ms, 5), "launches_timed": len(res["step_ms"]),
                     "timing": ("external CUDA events recorded inside the replayed graph; read for "
                                "the last replay of the timed region and for 8 synchronised "
                                "replays right after it") if args.graph else
                               "CUDA events around every launch of the timed region",
                     "norm_kernel_ms": round(statistics.mean(res["norm_ms"]), 5)
                     if res["norm_ms"] else None,
                     "frac_of_nominal_8000": round(achieved / 8000.0, 4)},
    }
    if world > 1:
        x = exchange_summary(res, world)
        if x:
            out["exchange"] = x
        ok = bool(res.get("replicas_identical")) and bool(res2.get("replicas_identical"))
        if parity is not None:
            parity["replicas_bit_identical_after_timed_runs"] = ok
            parity["all_true"] = parity["all_true"] and ok
            out["parity_check"] = parity
        if not ok:
            if rank == 0:
                print(json.dumps({"parity_check": "FAILED: replicas differ after the timed run",
                                  "resident": res.get("replicas_identical"),
                                  "e2e": res2.get("replicas_identical")}), flush=True)
            sys.exit(1)

    if extras:
        # what an unchanged reference script gets: no Module(graph=True)
        if args.graph:
            eg = run_capsule(wl, graph=False, steps=min(K, 50), warmup=max(5, args.warmup // 2),
                             e2e=False)
            out["eager"] = {"value": round(eg["value"], 1), "unit": "samples/s",
                            "ms_per_step": round(eg["ms_per_step"], 4), "steps": eg["steps"],
                            "cuda_graph_step": False,
                            "note": "same workload without the additive Module(graph=True) "
                                    "keyword (resident batch)"}
            xe = exchange_summary(eg, world)
            if xe:
                out["eager"]["exchange"] = xe
        out["torch_gpu"] = torch_gpu_baseline(wl, steps=min(K, 50), warmup=max(5, args.warmup // 2))
        # torch's default precision for this model (cuDNN convolutions in TF32): not the
        # parity-tested setting and not what the CPU reference arm computes, so not the headline
        tf = run_capsule(wl, graph=args.graph, steps=K, warmup=args.warmup, e2e=False,
                         tf32_conv=True)
        tfe = run_capsule(wl, graph=args.graph, steps=K, warmup=args.warmup, e2e=True,
                          loader_workers=args.loader_workers, tf32_conv=True)
        k2t = statistics.mean(tf["step_ms"]) if tf["step_ms"] else float("nan")
        out["tf32_conv"] = {
            "dtype": "f32 params and optimizer / TF32 cuDNN convolutions (torch default)",
            "value": round(tf["value"], 1), "unit": "samples/s",
            "ms_per_step": round(tf["ms_per_step"], 4),
            "e2e": {"value": round(tfe["value"], 1), "ms_per_step": round(tfe["ms_per_step"], 4)},
            "step_kernel_ms": round(k2t, 5),
            "step_kernel_gbs": round(algo * tf["n_params"] / (k2t * 1e-3) / 1e9, 1)}
        # the same two comparisons at this precision: unchanged script, and stock torch
        tfe_eager = run_capsule(wl, graph=False, steps=min(K, 50), warmup=max(5, args.warmup // 2),
                                e2e=False, tf32_conv=True)
        out["tf32_conv"]["eager"] = {"value": round(tfe_eager["value"], 1),
                                     "ms_per_step": round(tfe_eager["ms_per_step"], 4)}
        tft = torch_gpu_baseline(wl, steps=min(K, 50), warmup=max(5, args.warmup // 2),
                                 tf32_conv=True)
        out["tf32_conv"]["torch_gpu"] = {"value": tft["value"], "ms_per_step": tft["ms_per_step"]}
        # configs[2]: the workload the north_star puts the >= 7x scaling target on
        w50 = WORKLOADS["resnet50"]
        r50 = run_capsule(w50, graph=False, steps=args.r50_steps, warmup=5, e2e=False)
        r50e = run_capsule(w50, graph=False, steps=args.r50_steps, warmup=5, e2e=True,
                           loader_workers=args.loader_workers)
        out["resnet50"] = capsule_object(w50, r50, r50e, peak)
        # configs[3]: 8 GPUs (or on request)
        if world == 8 or os.environ.get("ROCKET_B200_BENCH_GPT2") == "1":
            wg = WORKLOADS["gpt2"]
            g2 = run_capsule(wg, graph=False, steps=16, warmup=8, e2e=False)
            out["gpt2"] = capsule_object(wg, g2, None, peak)
            out["gpt2"]["tokens_per_s"] = round(g2["value"] * 1024, 1)
            out["gpt2"]["optimizer_steps_timed"] = 16 // wg["accum"]
    if world > 1 and not args.no_optstep and out["config"]["grad_sync"].startswith("P2PGradSync"):
        out["optstep"] = optstep_microbench_dp(peak, rank, world, local)
    if world == 1 and not args.no_optstep:
        out["optstep"] = optstep_microbench(peak)
    if rank == 0 and world == 1 and not args.no_cpu and args.workload == "resnet18":
        out["cpu_baseline"] = cpu_port(budget_s=args.cpu_budget, threads=None)
    return out


def gpt2_shapes():
    s = [50257 * 768, 1024 * 768]
    for _ in range(12):
        s += [768, 768, 768 * 2304, 2304, 768 * 768, 768, 768, 768, 768 * 3072, 3072,
              3072 * 768, 768]
    return s + [768, 768]


def optstep_microbench(peak: float, n: int = 1_000_000_000, steps: int = 20, warmup: int = 5):
    """configs[4]: fused norm + AdamW + clip + zero_grad over 1e9 fp32 params (16 GB of state:
    far larger than L2), timed per kernel with CUDA events, on the two tensor lists SURVEY
    8(d) names: (i) one flat tensor, (ii) GPT-2-small's 148 shapes x 8 plus a pad tensor.  Then
    stock torch on the flat tensor: clip_grad_norm_(foreach=True) + AdamW(fused=True)."""
    from rocket_b200.backend import NativeBackend, TensorSpec
    dev = torch.device("cuda", torch.cuda.current_device())
    be = NativeBackend(dev)
    hp = dict(lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=1e-2, decoupled=True)

    def ours(sizes):
        offs, off = [], 0
        for k in sizes:
            offs.append(off)
            off += (k + 63) // 64 * 64                       # 256-byte aligned, like flat.py
        flat = {k: torch.zeros(off, device=dev) for k in "pgmv"}
        flat["p"].normal_(0, 0.02)
        plan = be.make_plan([TensorSpec(*[flat[k][o:o + m] for k in "pgmv"], group=0)
                             for o, m in zip(offs, sizes)])
        tn, ts, tp = [], [], []
        for i in range(warmup + steps):
            flat["g"].normal_(0, 1e-3)                          # fresh grads: clipping is active
            e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            e[0].record()
            norm = be.l2norm(plan, 1.0)
            e[1].record()
            be.adam_step(plan, [dict(hp, step=i + 1)], norm, 1.0, 1.0, True)
            e[2].record()
            torch.cuda.synchronize()
            if i >= warmup:
                tn.append(e[0].elapsed_time(e[1]))
                ts.append(e[1].elapsed_time(e[2]))
        for i in range(3 + steps // 2):      # the pair back to back (no event between them:
            flat["g"].normal_(0, 1e-3)       # the step kernel starts as a programmatic dependent)
            e = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
            e[0].record()
            norm = be.l2norm(plan, 1.0)
            be.adam_step(plan, [dict(hp, step=warmup + steps + i + 1)], norm, 1.0, 1.0, True)
            e[1].record()
            torch.cuda.synchronize()
            if i >= 3:
                tp.append(e[0].elapsed_time(e[1]))
        del flat, plan
        torch.cuda.empty_cache()
        t_norm, t_step, t_pair = statistics.mean(tn), statistics.mean(ts), statistics.mean(tp)
        total = sum(sizes)
        gbs_step = ALGO_BYTES_STEP * total / (t_step * 1e-3) / 1e9
        gbs_full = (ALGO_BYTES_STEP + ALGO_BYTES_NORM) * total / (t_pair * 1e-3) / 1e9
        return {"tensors": len(sizes), "step_kernel_ms": round(t_step, 4),
                "norm_kernel_ms": round(t_norm, 4), "pair_ms": round(t_pair, 4),
                "step_kernel_gbs": round(gbs_step, 1),
                "norm_kernel_gbs": round(ALGO_BYTES_NORM * total / (t_norm * 1e-3) / 1e9, 1),
                "full_step_gbs_36B": round(gbs_full, 1),
                "step_kernel_frac_of_measured_peak": round(gbs_step / peak, 4),
                "frac_of_measured_peak": round(gbs_full / peak, 4),
                "frac_of_nominal_8000": round(gbs_full / 8000.0, 4)}

    flat_res = ours([n])
    lst = gpt2_shapes() * 8
    lst = lst + [n - sum(lst)]
    list_res = ours(lst)

    # stock torch on the flat tensor: what the reference's Optimizer capsule reaches today with
    # a fused torch optimizer, plus the clipping pass (28 + 12 B/param of traffic)
    p = torch.nn.Parameter(torch.randn(n, device=dev) * 0.02)
    p.grad = torch.empty(n, device=dev)
    topt = torch.optim.AdamW([p], lr=1e-3, weight_decay=1e-2, fused=True)
    tt = []
    for i in range(3 + steps // 2):
        p.grad.normal_(0, 1e-3)
        e = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        e[0].record()
        torch.nn.utils.clip_grad_norm_([p], 1.0, foreach=True)
        topt.step()
        p.grad.zero_()                      # persistent gradients, like ours (4 B/param more)
        e[1].record()
        torch.cuda.synchronize()
        if i >= 3:
            tt.append(e[0].elapsed_time(e[1]))
    del p, topt
    torch.cuda.empty_cache()
    t_torch = statistics.mean(tt)
    out = {"params": n, "layout": "one flat fp32 tensor", "steps": steps}
    out.update(flat_res)
    out["realistic_list"] = dict(list_res, layout="GPT-2-small's 148 shapes x 8 + one pad "
                                                  "tensor, 256-byte aligned segments")
    out["torch_fused_ms"] = round(t_torch, 4)
    out["torch_fused_what"] = ("clip_grad_norm_(foreach=True) + torch.optim.AdamW(fused=True)"
                               ".step() + grad.zero_() on the same flat tensor")
    out["speedup_vs_torch_fused"] = round(t_torch / flat_res["pair_ms"], 3)
    return out


def optstep_microbench_dp(peak: float, rank: int, world: int, local: int,
                          n: int = 1_000_000_000, steps: int = 10, warmup: int = 3):
    """configs[4] at N > 1: every rank holds a full replica of 1e9 fp32 params; per step the
    gradients (different per rank) are averaged by the fused NVLink exchange kernel, which
    also yields the norm partials, then finalize + fused AdamW + clip + zero_grad.  Timed on
    the device with CUDA events, max over ranks."""
    from rocket_b200 import _native as N
    from rocket_b200.backend import NativeBackend, TensorSpec
    from rocket_b200.comm import alloc_arena, map_arena
    dev = torch.device("cuda", local)
    be = NativeBackend(dev)
    sig = N.RK_COMM_SIGNAL_BYTES
    scratch = -(-be.scratch_bytes([n]) // 256) * 256
    # Sharded arm at every world size (validated bit for bit against the replicated step by
    # parity_check above, on this very run, before anything is timed).
    run_sharded = os.environ.get("ROCKET_B200_BENCH_SHARDED", "1") == "1"
    arena = alloc_arena(sig + scratch + (8 if run_sharded else 4) * n, dev)  # [signals|scratch|g|p]
    g = arena[sig + scratch:sig + scratch + 4 * n].view(torch.float32)
    p = arena[sig + scratch + 4 * n:].view(torch.float32) if run_sharded \
        else torch.empty(n, device=dev)
    torch.manual_seed(5)
    p.normal_(0, 0.02)
    m = torch.zeros(n, device=dev)
    v = torch.zeros(n, device=dev)
    plan = be.make_plan([TensorSpec(p=p, g=g, m=m, v=v, group=0)], scratch=arena[sig:sig + scratch])

    class Eng:
        device, process_index, num_processes = dev, rank, world
    ptrs, mc_ptr, keep = map_arena(arena, Eng)
    comm = be.comm_create(rank, world, ptrs, sig, mc_ptr)
    dist.barrier()
    use_mc = bool(mc_ptr) and world >= 4
    hp = dict(lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=1e-2, decoupled=True)
    stream = torch.cuda.current_stream().cuda_stream
    tx, ts = [], []
    for i in range(warmup + steps):
        torch.manual_seed(1000 * i + rank)
        g.normal_(0, 1e-3)
        torch.cuda.synchronize()
        dist.barrier()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        e[0].record()
        be.comm_allreduce_norm(comm, plan, 0, plan.n_chunks_norm, 0, 74 if use_mc else 0, stream,
                               use_mc)
        norm = be.l2norm_finalize(plan)
        e[1].record()
        be.adam_step(plan, [dict(hp, step=i + 1)], norm, 1.0, 1.0, True)
        e[2].record()
        torch.cuda.synchronize()
        if i >= warmup:
            tx.append(e[0].elapsed_time(e[1]))
            ts.append(e[1].elapsed_time(e[2]))
    # the same step with the optimizer sharded: reduce-scatter (+ norm), then every rank
    # updates the 1/W of the chunks it owns and writes the new parameters to all ranks
    # (P2P loads / stores at every W, like the engine: comm.P2PGradSync)
    table = torch.zeros(N.RK_STEP_TABLE_BYTES, dtype=torch.uint8, device=dev)
    host_table = torch.zeros(N.RK_STEP_TABLE_BYTES, dtype=torch.uint8).pin_memory()
    del m, v                                   # the sharded step keeps 1/W of the moments
    torch.cuda.empty_cache()
    shard = -(-plan.n_chunks_norm // world) * plan.chunk_norm
    sm, sv = torch.zeros(shard, device=dev), torch.zeros(shard, device=dev)
    sx, ss = [], []
    for i in range(warmup + steps if run_sharded else 0):
        torch.manual_seed(1000 * i + rank)
        g.normal_(0, 1e-3)
        be.step_table_write("adamw", [dict(hp, step=i + 1)], host_table)
        table.copy_(host_table, non_blocking=True)
        torch.cuda.synchronize()
        dist.barrier()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        e[0].record()
        be.comm_reduce_scatter_norm(comm, plan, 0, plan.n_chunks_norm, 0, 0, stream, False)
        norm = be.l2norm_finalize(plan)
        e[1].record()
        be.comm_sharded_step(comm, plan, "adamw", table.data_ptr(), norm, 1.0, 0, False, sm, sv)
        e[2].record()
        torch.cuda.synchronize()
        if i >= warmup:
            sx.append(e[0].elapsed_time(e[1]))
            ss.append(e[1].elapsed_time(e[2]))
    t = torch.tensor([statistics.mean(tx), statistics.mean(ts), statistics.mean(sx or [0.0]),
                      statistics.mean(ss or [0.0])], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    t_x, t_s, s_x, s_s = (float(x) for x in t)
    dist.barrier()
    be.comm_destroy(comm)
    del keep, p, sm, sv, g, plan, arena
    torch.cuda.ipc_collect()
    torch.cuda.empty_cache()
    bus = 2 * (world - 1) / world * 4 * n / (t_x * 1e-3) / 1e9
    return {"params": n, "layout": "one flat fp32 tensor, full replica per rank", "steps": steps,
            "exchange": "NVSwitch multimem" if use_mc else "NVLink P2P",
            "exchange_plus_norm_ms": round(t_x, 4), "exchange_bus_gbs": round(bus, 1),
            "step_kernel_ms": round(t_s, 4),
            "step_kernel_gbs": round(ALGO_BYTES_STEP * n / (t_s * 1e-3) / 1e9, 1),
            "ms_per_optimizer_step": round(t_x + t_s, 4),
            "sharded": {"moment_bytes_per_gpu": 8 * shard,
                        "reduce_scatter_plus_norm_ms": round(s_x, 4),
                        "sharded_step_ms": round(s_s, 4),
                        "ms_per_optimizer_step": round(s_x + s_s, 4)} if run_sharded else
            "not run at this world size by default (ROCKET_B200_BENCH_SHARDED=1 runs it)",
            "step_kernel_frac_of_measured_peak":
                round(ALGO_BYTES_STEP * n / (t_s * 1e-3) / 1e9 / peak, 4)}


def cpu_port(budget_s: float, threads, steps=None, warmup=1):
    """The looper port on the host cores: ResNet-18 bs=256, AdamW(foreach=False) + clip."""
    import workloads
    from oracle import looper_port
    if threads:
        torch.set_num_threads(threads)
    wl = WORKLOADS["resnet18"]
    net = make_resnet18()
    opt = torch.optim.AdamW(net.parameters(), lr=1e-3, weight_decay=1e-2, foreach=False)
    batch = synthetic_batch(wl, 0)
    obj = workloads.CrossEntropy()
    if steps is None:                      # bounded sample: as many steps as fit the budget
        t1 = looper_port.time_steps(net, obj, opt, batch, steps=1, warmup=1, max_norm=1.0,
                                    device="cpu")
        steps = max(3, min(50, int(budget_s / max(t1, 1e-3))))
    sec = looper_port.time_steps(net, obj, opt, batch, steps=steps, warmup=warmup, max_norm=1.0,
                                 device="cpu")
    return {"value": round(BATCH / sec, 2), "unit": "samples/s", "cores": torch.get_num_threads(),
            "kind": "port", "ms_per_step": round(sec * 1e3, 2), "steps": steps,
            "sample": f"{steps} steps of the same workload (ResNet-18 bs=256, fp32) on CPU, "
                      f"os.cpu_count()={os.cpu_count()}"}


def run_reference(args) -> dict:
    """The reference arm: the reference's own arithmetic for this step (looper port: torch.optim
    foreach=False + clip_grad_norm_) on this box's host cores, all threads, the same --steps
    and --warmup and the same config as our arm.  One process whatever N is: the reference's
    CPU path has no multi-GPU meaning, so at N > 1 the driver's ratio compares N GPUs with one
    host (rank 0 runs it, the others exit)."""
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return {}
    wl = WORKLOADS["resnet18"]
    base = cpu_port(budget_s=0, threads=os.cpu_count(), steps=max(1, args.steps),
                    warmup=max(1, min(args.warmup, 3)))
    return {"impl": "reference", "metric": METRIC, "value": base["value"], "unit": "samples/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": base["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": wl["dtype"], "data": "synthetic",
            "config": {"workload": wl["label"], "global_batch": BATCH, "parallelism": "dp1",
                       "params": 11181642},
            "runs_on": "host CPU cores: reference arithmetic (torch.optim foreach=False + "
                       "clip_grad_norm_) through oracle/looper_port.py; warm-up capped at 3 steps",
            "cpu_baseline": base,
            "e2e": {"value": base["value"], "unit": "samples/s", "h2d_bytes_per_step": 0,
                    "d2h_bytes_per_step": 0}}


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-optstep", action="store_true", help="skip the 1e9-param microbench")
    ap.add_argument("--cpu-budget", type=float, default=15.0)
    ap.add_argument("--workload", default="resnet18", choices=list(WORKLOADS))
    ap.add_argument("--graph", type=int, default=None,
                    help="Module(graph=...): replay the training step from a CUDA graph (1) or "
                         "run it eagerly (0)")
    ap.add_argument("--lean", action="store_true",
                    help="headline legs only: no eager / torch_gpu / resnet50 / gpt2 objects")
    ap.add_argument("--no-parity", action="store_true", help="skip the N > 1 parity_check")
    ap.add_argument("--r50-steps", type=int, default=20, help="timed steps of the resnet50 object")
    ap.add_argument("--loader-workers", type=int, default=0,
                    help="DataLoader worker processes of the e2e leg (0 = load in-process)")
    args = ap.parse_args()
    if args.graph is None:          # the host-bound headline workload replays its step from a graph
        args.graph = 1 if args.workload == "resnet18" else 0
    if args.impl == "reference":
        out = run_reference(args)
        if out:
            print(json.dumps(out), flush=True)
        return
    try:
        out = run_ours(args)
    except AssertionError as e:          # a parity check failed: say so and fail the run
        if int(os.environ.get("RANK", 0)) == 0:
            print(json.dumps({"parity_check": f"FAILED: {e!r}"}), flush=True)
        raise
    if int(os.environ.get("RANK", 0)) == 0:
        print(json.dumps(out), flush=True)
    if dist.is_available() and dist.is_initialized():
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
