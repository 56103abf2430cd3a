"""Loss-curve parity: our native training step vs. the unmodified reference (same init, same batches).

Both train `--model` for `--steps` steps under bf16 autocast with AdamW(lr, betas (0.9,0.95), wd 0.1) on the
same synthetic token batches, starting from the reference model's random init (state dicts are compatible).
Prints the two loss curves and their max relative difference.
    python tools/compare_training.py --model small --steps 12
"""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "baseline"))
import torch
import ref_loader
from cs336_a2_systems_b200.models import TransformerLM, get_model_config
from cs336_a2_systems_b200.optim import AdamW
from cs336_a2_systems_b200.train import Trainer

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="small"); ap.add_argument("--steps", type=int, default=12)
ap.add_argument("--batch", type=int, default=4); ap.add_argument("--ctx", type=int, default=512)
ap.add_argument("--lr", type=float, default=3e-4); ap.add_argument("--graph", action="store_true")
ap.add_argument("--json", default=None)
args = ap.parse_args()
dev = torch.device("cuda")
ref = ref_loader.load()
cfg = get_model_config(args.model)
torch.manual_seed(0)
with torch.device(dev):
    rm = ref.BasicsTransformerLM(vocab_size=10000, context_length=args.ctx, rope_theta=10000.0, **cfg)
om = TransformerLM(vocab_size=10000, context_length=args.ctx, rope_theta=10000.0, **cfg).to(dev)
om.load_state_dict(rm.state_dict())
kw = dict(lr=args.lr, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1)
ropt = ref.AdamW(rm.parameters(), **kw)
tr = Trainer(om, AdamW(om.parameters(), **kw), cuda_graph=args.graph, overlap_optimizer=True)
g = torch.Generator().manual_seed(1)
pool = [torch.randint(0, 10000, (args.batch, args.ctx), generator=g).to(dev) for _ in range(4)]
lr_, lo_ = [], []
for s in range(args.steps):
    x = pool[s % len(pool)]; y = torch.roll(x, -1, 1)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        loss = ref.cross_entropy(rm(x), y)
    ropt.zero_grad(); loss.backward(); ropt.step()
    lr_.append(float(loss)); lo_.append(float(tr.step(x, y)))
rel = max(abs(a - b) / abs(a) for a, b in zip(lr_, lo_))
print("reference:", " ".join(f"{v:.4f}" for v in lr_))
print("ours     :", " ".join(f"{v:.4f}" for v in lo_))
print(f"max relative loss difference over {args.steps} steps: {rel:.4%}")
if args.json:
    json.dump({"model": args.model, "reference": lr_, "ours": lo_, "max_rel_diff": rel}, open(args.json, "w"))
