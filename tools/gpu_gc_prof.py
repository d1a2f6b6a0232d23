"""Debug aid: where the time of one GradCache-sized CLIP step (configs[1], B = 256 in 32 leaves) goes.
python tools/gpu_gc_prof.py [profile]"""
import sys, os, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "protein-clip_b200"))
import torch
from pclip import ops, synth
xs, ys = synth.gradcache_leaves()
xs = [t.cuda().requires_grad_() for t in xs]
ys = [t.cuda().requires_grad_() for t in ys]
scale = torch.tensor(65536.0, device="cuda")


def direct():
    for t in xs + ys:
        t.grad = None
    return ops.clip_loss_gc_direct(xs, ys, scale)


def autograd():
    for t in xs + ys:
        t.grad = None
    loss = ops.clip_loss(xs, ys, compute=ops.BF16)
    (loss * 65536.0).backward()


def wall(f, n=300):
    for _ in range(30): f()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): f()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e6


print("direct path (train_gc) us", wall(direct))
print("autograd path us", wall(autograd))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); direct(); e1.record(); torch.cuda.synchronize()
print("one direct step, device time us", e0.elapsed_time(e1) * 1e3)
if len(sys.argv) > 1:
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as p:
        for _ in range(20): direct()
        torch.cuda.synchronize()
    print(p.key_averages().table(sort_by="self_cpu_time_total", row_limit=20, max_name_column_width=60))
