"""Debug aid: GPU timeline of one row-sharded CLIP step on rank 0 (torchrun --nproc-per-node N tools/gpu_sharded_timeline.py)."""
import os, sys
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "protein-clip_b200"))
import torch
from torch.profiler import profile, ProfilerActivity
import bench
d = bench.Dist(int(os.environ.get("WORLD_SIZE", "1")))
w = bench.ClipLarge(d)
w.record = False
if os.environ.get("TIMELINE_GRAPH", "1") == "1":
    print("graph:", w.enable_graph(), flush=True) if d.rank == 0 else w.enable_graph()
for _ in range(5):
    w.step()
torch.cuda.synchronize(); d.barrier()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3):
        w.step()
    torch.cuda.synchronize()
if d.rank == 0:
    evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    evs.sort(key=lambda e: e.time_range.start)
    t0 = evs[0].time_range.start
    n = len(evs) // 3
    prev_end = None
    for e in evs[n:2 * n]:
        gap = "" if prev_end is None else f"gap {e.time_range.start - prev_end:7.1f}"
        print(f"{e.time_range.start - t0:9.1f} us  dur {e.time_range.end - e.time_range.start:8.1f} us  {gap:14s} {e.name[:70]}")
        prev_end = e.time_range.end
d.close()
