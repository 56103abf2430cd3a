"""Kernel-time breakdown of one eager training step (torch.profiler/CUPTI) — for SHARES only, never
for reported numbers.  usage: python tools/profile_step.py [--model xl] [--impl ours|reference]"""
import argparse, collections, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="xl"); ap.add_argument("--out", default="gpurun_out/step_profile.json")
ap.add_argument("--overlap", action="store_true"); ap.add_argument("--batch", type=int, default=4); ap.add_argument("--ctx", type=int, default=512)
args = ap.parse_args()
from cs336_a2_systems_b200.models import build_lm
from cs336_a2_systems_b200.optim import AdamW
from cs336_a2_systems_b200.train import Trainer
dev = torch.device("cuda")
model = build_lm(args.model, device=dev, context_length=args.ctx)
opt = AdamW(model.parameters(), lr=1e-4, betas=(0.9, 0.95), weight_decay=0.1)
tr = Trainer(model, opt, cuda_graph=False, overlap_optimizer=args.overlap)
x = torch.randint(0, 10000, (args.batch, args.ctx), device=dev); y = torch.roll(x, -1, 1)
for _ in range(3):
    tr.step(x, y)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    tr.step(x, y)
    torch.cuda.synchronize()
agg = collections.defaultdict(lambda: [0.0, 0])
for e in prof.events():
    if e.device_type == torch.autograd.DeviceType.CUDA:
        name = e.name.split("<")[0].split("(")[0][:70]
        if "gemm_bf16_kernel" in e.name:
            name = e.name[:110]
        agg[name][0] += e.device_time
        agg[name][1] += 1
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
t0 = min(e.time_range.start for e in evs); t1 = max(e.time_range.end for e in evs)
print(f"span first-kernel-start .. last-kernel-end: {(t1 - t0)/1e3:.2f} ms")
ad = sorted([e for e in evs if "adamw_multi" in e.name], key=lambda e: e.time_range.start)
if len(ad) > 1:
    busy_other = 0.0
    for a in ad[:6] + ad[-3:]:
        ov = [e for e in evs if e is not a and "adamw" not in e.name and e.time_range.start < a.time_range.end and e.time_range.end > a.time_range.start]
        print(f"  adamw launch at +{(a.time_range.start - t0)/1e3:7.2f} ms dur {a.device_time/1e3:6.3f} ms; {len(ov)} other kernels overlap it in time")
total = sum(v[0] for v in agg.values())
rows = sorted(agg.items(), key=lambda kv: -kv[1][0])
print(f"total kernel time {total/1e3:.2f} ms over {sum(v[1] for v in agg.values())} launches")
for n, (t, c) in rows[:40]:
    print(f"{t/1e3:9.3f} ms {100*t/total:5.1f}% {c:5d}x  {n}")
json.dump({"total_ms": total / 1e3, "kernels": [{"name": n, "ms": t / 1e3, "count": c} for n, (t, c) in rows]},
          open(args.out, "w"), indent=1)
