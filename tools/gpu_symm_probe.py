"""Debug aid: all-gather of the packed bf16 row shards by pulling from peer memory (torch symmetric memory) vs NCCL.
torchrun --nproc-per-node N tools/gpu_symm_probe.py"""
import os, sys, time
import torch, torch.distributed as dist
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
Ma, D = 65536 // world, 128
full = torch.empty(2, world * Ma, D, dtype=torch.bfloat16, device=dev)
loc = torch.full((2, Ma, D), float(rank + 1), dtype=torch.bfloat16, device=dev)

def timeit(f, n=50):
    for _ in range(10): f()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): f()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3

def nccl_gather():
    with dist._coalescing_manager(group=dist.group.WORLD, device=dev, async_ops=False):
        dist.all_gather_into_tensor(full[0], loc[0]); dist.all_gather_into_tensor(full[1], loc[1])
t = timeit(nccl_gather)
if rank == 0: print(f"nccl coalesced all-gather 2 x {Ma*D*2/1e6:.1f} MB per rank: {t:.1f} us", flush=True)
try:
    import torch.distributed._symmetric_memory as symm
    bufs = [symm.empty((2, Ma, D), dtype=torch.bfloat16, device=dev) for _ in range(2)]
    hdls = [symm.rendezvous(b, dist.group.WORLD) for b in bufs]
    peers = [[h.get_buffer(p, (2, Ma, D), torch.bfloat16) for p in range(world)] for h in hdls]
    step = [0]
    def pull():
        i = step[0] & 1; step[0] += 1
        bufs[i].copy_(loc)
        hdls[i].barrier(channel=0)
        for s in range(world):
            p = (rank + s) % world
            full[:, p * Ma:(p + 1) * Ma].copy_(peers[i][p])
    t = timeit(pull)
    torch.cuda.synchronize()
    ok = all(bool((full[:, p * Ma:(p + 1) * Ma] == p + 1).all()) for p in range(world))
    if rank == 0: print(f"symmetric-memory pull ({world} copies + barrier): {t:.1f} us, correct={ok}", flush=True)
except Exception as e:
    if rank == 0: print("symmetric memory unavailable:", repr(e)[:300], flush=True)
dist.destroy_process_group()
