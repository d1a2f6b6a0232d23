"""Debug aid: where the time of one GradCache-sized CLIP step (configs[1], B = 256 in 32 leaves) goes."""
import sys, os, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "protein-clip_b200"))
import torch
from pclip import ops, synth
xs, ys = synth.gradcache_leaves()
xs = [t.cuda().requires_grad_() for t in xs]
ys = [t.cuda().requires_grad_() for t in ys]

def step():
    for t in xs + ys:
        t.grad = None
    loss = ops.clip_loss(xs, ys, compute=ops.BF16)
    (loss * 65536.0).backward()

def wall(f, n=300):
    for _ in range(30): f()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): f()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e6

print("step us", wall(step))
dxs = [t.detach() for t in xs]; dys = [t.detach() for t in ys]
print("raw call (no autograd) us", wall(lambda: ops.clip_loss_raw(dxs, dys, ops.BF16, True)))
print("forward only us", wall(lambda: ops.clip_loss(xs, ys, compute=ops.BF16)))
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as p:
    for _ in range(20): step()
    torch.cuda.synchronize()
print(p.key_averages().table(sort_by="self_cpu_time_total", row_limit=28, max_name_column_width=60))
