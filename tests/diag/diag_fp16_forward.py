"""Diagnostic (GPU): where do the engine's and the port's fp16 forward passes part ways?"""
import os, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
os.environ["TQDM_DISABLE"] = "1"
import torch
import torch.nn.functional as F
import workloads
import rocket_b200 as rk

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
torch.backends.cudnn.deterministic = True
torch.backends.cudnn.benchmark = False

for dtype in (torch.float16, torch.bfloat16):
    torch.manual_seed(0)
    net = workloads.LeNet().cuda()
    ds = workloads.synthetic_classification(64 * 2, (1, 28, 28), 10, 1, 128.0)
    x, y = ds.tensors[0][:64].cuda(), ds.tensors[1][:64].cuda()

    def fwd(xx, nested=False, to32=False):
        with torch.autocast("cuda", dtype=dtype):
            if nested:
                with torch.autocast("cuda", dtype=dtype):
                    out = net([xx, y])[2]
            else:
                out = net([xx, y])[2]
            if to32:
                out = out.float()
            return out, F.cross_entropy(out, y)

    a, la = fwd(x)
    b, lb = fwd(x, nested=True, to32=True)
    buf = torch.empty(x.numel() * 4 + 4096, dtype=torch.uint8, device="cuda")
    for off in (0, 16, 256, 1024):
        xv = buf[off:off + x.numel() * 4].view(torch.float32).view(x.shape)
        xv.copy_(x)
        c, lc = fwd(xv)
        print(dtype, "offset", off, "logit diff", float((a.float() - c.float()).abs().max()),
              "loss", float(la), float(lc))
    print(dtype, "nested+fp32 out: logit diff", float((a.float() - b).abs().max()),
          "loss", float(la), float(lb))
    # twice the same call
    a2, la2 = fwd(x)
    print(dtype, "same call twice: logit diff", float((a.float() - a2.float()).abs().max()))

    # through the engine: grab the logits the Loss capsule sees at the first iteration
    seen = {}

    class Probe(rk.Capsule):
        def __init__(self):
            super().__init__(priority=1150)

        def launch(self, attrs=None):
            if "logits" not in seen:
                seen["logits"] = attrs.batch[2].detach().clone()
                seen["x_ptr"] = attrs.batch[0].data_ptr()
                seen["x_eq"] = bool(torch.equal(attrs.batch[0], x))
                seen["grad"] = torch.is_grad_enabled()
                seen["stream"] = torch.cuda.current_stream().cuda_stream

    torch.manual_seed(0)
    net2 = workloads.LeNet()
    looper = rk.Looper([rk.Dataset(ds, batch_size=64, shuffle=False),
                        rk.Module(net2, capsules=[Probe(), rk.Loss(objective=workloads.CrossEntropy())])],
                       tag="t")
    rk.Launcher([looper], num_epochs=1, destroy_process_group_after_launch=False,
                mixed_precision="fp16" if dtype == torch.float16 else "bf16").launch()
    e = seen["logits"]
    print(dtype, "engine: x equal", seen["x_eq"], "x_ptr % 512 =", seen["x_ptr"] % 512,
          "stream", seen["stream"], "logits dtype", e.dtype,
          "diff vs plain", float((a.float() - e.float()).abs().max()),
          "weights equal", all(torch.equal(p, q) for p, q in zip(net.parameters(), net2.parameters())))
    # no_grad vs grad
    with torch.no_grad():
        n, _ = fwd(x)
    print(dtype, "no_grad diff", float((a.float() - n.float()).abs().max()))
