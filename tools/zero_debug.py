"""Two-rank debugging aid: trains the same tiny LM with (a) NVLink DDP + replicated AdamW and (b) the bucket-fused
ZeRO-1 step, and prints the per-parameter max difference after every step.  Run on a 2-GPU box:
``python tools/zero_debug.py``."""
import copy, os, sys
import torch, torch.distributed as dist, torch.multiprocessing as mp
sys.path.insert(0, os.getcwd())

def worker(rank, ws, port):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=ws, device_id=torch.device("cuda", rank))
    dev = torch.device("cuda", rank)
    from cs336_a2_systems_b200.models import TransformerLM
    from cs336_a2_systems_b200.ops import cross_entropy
    from cs336_a2_systems_b200.optim import AdamW
    from cs336_a2_systems_b200.parallel import DDPBucketed, ShardedOptimizer
    torch.manual_seed(0)
    base = TransformerLM(vocab_size=512, context_length=128, d_model=128, num_layers=2, num_heads=2, d_ff=256).to(dev)
    a = DDPBucketed(copy.deepcopy(base), 0.25, backend="symm")
    b = DDPBucketed(copy.deepcopy(base), 0.25, backend="symm")
    oa = AdamW(a.parameters(), lr=1e-3, weight_decay=0.1)
    ob = ShardedOptimizer(b.parameters(), AdamW, lr=1e-3, weight_decay=0.1)
    assert ob.enable_symmetric(b)
    g = torch.Generator().manual_seed(7 + rank)
    for step in range(3):
        x = torch.randint(0, 512, (2, 128), generator=g).to(dev)
        y = torch.roll(x, -1, 1)
        ls = []
        for m, o in ((a, oa), (b, ob)):
            m.on_train_batch_start()
            o.zero_grad()
            with torch.autocast("cuda", dtype=torch.bfloat16):
                loss = cross_entropy(m(x), y)
            loss.backward()
            m.finish_gradient_synchronization()
            o.step()
            ls.append(loss.item())
        torch.cuda.synchronize()
        if rank == 0:
            print("step", step, "loss", ls)
            pa = dict(a.module.named_parameters())
            for n, p in b.module.named_parameters():
                d = (p - pa[n]).abs().max().item()
                # which part of the tensor differs
                flat = (p - pa[n]).abs().flatten()
                bad = (flat > 2e-4).nonzero().flatten()
                print(f"  {n:45s} maxdiff {d:.3e} nbad {bad.numel()}/{flat.numel()} first {bad[:1].tolist()} last {bad[-1:].tolist()}")
    dist.barrier(); dist.destroy_process_group()

if __name__ == "__main__":
    mp.spawn(worker, args=(2, 29533), nprocs=2, join=True)
