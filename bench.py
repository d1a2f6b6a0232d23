#!/usr/bin/env python
"""Benchmark of protein-clip's contrastive hot path on B200 (contract: see the task statement / DESIGN.md section 6).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

One JSON line on stdout (rank 0).  Workloads (BASELINE.json configs):
    clip_b65536  configs[3]: CLIP InfoNCE fwd+bwd, B = 65 536, D = 128, bf16 tensor-core path; rows sharded over the
                 N ranks with an embedding all-gather (strong scaling).  DEFAULT: it is the configuration the
                 metric's roofline target is quoted on and it fits one GPU (the B x B logits never reach HBM).
    clip_gc256   configs[1]: 16 + 16 GradCache leaves of [16, 128], grad_output = 65536 (launch-latency bound)
    filip_b256   configs[2]: FILIP token-wise late interaction, B = 256, T = 512/512
    topk_4m      configs[4]: top-10 of 2^22 bank rows for a stream of queries
The non-default workloads are also run briefly by the default invocation and reported under "workloads".
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
for _p in (REPO, os.path.join(REPO, "protein-clip_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402


def load_peaks():
    p = os.path.join(REPO, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        with open(p) as f:
            d = json.load(f)
        return dict(hbm=d["hbm_gbs"], tc_burst=d["bf16_tflops"], tc=d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                    src="measured")
    return dict(hbm=6650.0, tc_burst=1590.0, tc=1400.0, src="fallback")


# ---------------------------------------------------------------------------------------------------------
# clocks: sampled with NVML DURING the timed region
# ---------------------------------------------------------------------------------------------------------
class ClockSampler:
    REASONS = {0x4: "sw_power_cap", 0x8: "hw_slowdown", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
               0x80: "hw_power_brake_slowdown"}

    def __init__(self, index: int, period: float = 0.02):
        self.samples, self.mask, self.max_mhz, self.ok = [], 0, None, False
        self.period, self._stop = period, threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False
        self.t = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    self.mask |= int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
                except Exception:
                    self.mask |= int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
            except Exception:
                pass
            self._stop.wait(self.period)

    def __enter__(self):
        if self.ok:
            self.t.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self.ok:
            self.t.join(timeout=2)

    def summary(self):
        if not self.ok or not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": ["nvml unavailable"]}
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2], "sm_max_mhz": self.max_mhz, "samples": len(s),
                "reasons": [n for b, n in self.REASONS.items() if self.mask & b]}


# ---------------------------------------------------------------------------------------------------------
# distributed plumbing
# ---------------------------------------------------------------------------------------------------------
class Dist:
    def __init__(self, gpus: int):
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        self.on = self.world > 1
        if self.on:
            import torch.distributed as dist
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            torch.cuda.set_device(self.local)
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
            self.dist = dist
        elif gpus > 1:
            raise SystemExit("--gpus N > 1 must be launched with torch.distributed.run (one rank per GPU)")
        self.dev = torch.device("cuda", self.local)

    def barrier(self):
        if self.on:
            self.dist.barrier(device_ids=[self.local])

    def max_ms(self, ms: float) -> float:
        if not self.on:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t)

    def sum_int(self, v: int) -> int:
        if not self.on:
            return v
        t = torch.tensor([v], dtype=torch.int64, device=self.dev)
        self.dist.all_reduce(t)
        return int(t)

    def close(self):
        if self.on:
            self.dist.destroy_process_group()


def timed(dist: Dist, fn, steps: int, warmup: int, begin=None):
    """W untimed steps, then exactly K steps between barrier + synchronize; device time via CUDA events on the
    launching stream; returns max-over-ranks milliseconds for the K steps.  `begin` runs after the start event."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    if begin is not None:
        begin()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    return dist.max_ms(e0.elapsed_time(e1))


# ---------------------------------------------------------------------------------------------------------
# workloads
# ---------------------------------------------------------------------------------------------------------
class ClipLarge:
    """configs[3]: CLIP InfoNCE fwd+bwd at B = 65 536, D = 128 (ref: modules/training_utils.py:74-89), bf16."""
    name, metric, unit, dtype = "clip_b65536", "clip_infonce_fwd_bwd_pairs_per_s", "pairs/s", "bf16"
    scaling = "strong"

    def __init__(self, dist: Dist, B: int = 65536, D: int = 128):
        from pclip import ops, sharded, synth
        self.ops, self.dist, self.B, self.D = ops, dist, B, D
        P, r = dist.world, dist.rank
        self.Ma = B // P
        x, y = synth.clip_scaled_embeddings(B, D, cfg_id=4)          # seeded on the CPU generator
        sl = slice(r * self.Ma, (r + 1) * self.Ma)
        self.hx, self.hy = x[sl].contiguous().pin_memory(), y[sl].contiguous().pin_memory()
        self.x, self.y = self.hx.to(dist.dev), self.hy.to(dist.dev)
        self.sx, self.sy = torch.empty_like(self.x), torch.empty_like(self.y)   # e2e staging (H2D target)
        self.host_loss = torch.empty(1, dtype=torch.float32).pin_memory()
        self.job = sharded.ShardedClip(B, D, ops.BF16, dist.dev) if dist.on else None
        self.plan = self.job.s if self.job else ops.ClipPlan(B, B, D, ops.BF16, dist.dev)
        self.events, self.record = {"rowsumexp": [], "grad": []}, False
        self.units = B
        self.h2d = 2 * self.Ma * D * 4
        self.d2h = 4

    def _run(self, x, y):
        ev = None
        if self.record:      # one (start, end) CUDA-event pair per stage per step, on the launching stream
            ev = {k: (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for k in self.events}
            for k in ev:
                self.events[k].append(ev[k])
        if self.job:
            return self.job.step([x], [y], events=ev)
        return self.plan.step([x], [y], events=ev)

    def step(self):
        self._run(self.x, self.y)

    # End-to-end leg: every step copies its fp32 inputs from pinned host memory and reads the loss back.  The copy of
    # step i+1's inputs is issued on a second stream while step i computes (two staging buffers), as a training loop
    # with a prefetching loader does; e2e_begin() drops any copy issued before the timed region starts, so the K
    # timed steps contain K input copies of their own (the prefetch issued by the last step is the unused one).
    def e2e_begin(self):
        self._ready = [False, False]
        self._cur = 0

    def _issue_copy(self, b):
        with torch.cuda.stream(self.copy_stream):
            self.sx2[b].copy_(self.hx, non_blocking=True)
            self.sy2[b].copy_(self.hy, non_blocking=True)
            self.copied[b].record(self.copy_stream)
        self._ready[b] = True

    def e2e_step(self):
        if not hasattr(self, "copy_stream"):
            self.copy_stream = torch.cuda.Stream(self.dist.dev)
            self.sx2, self.sy2 = [self.sx, torch.empty_like(self.sx)], [self.sy, torch.empty_like(self.sy)]
            self.copied = [torch.cuda.Event(), torch.cuda.Event()]
            self.e2e_begin()
        b = self._cur
        main = torch.cuda.current_stream()
        if not self._ready[b]:
            self.copy_stream.wait_stream(main)
            self._issue_copy(b)
        main.wait_event(self.copied[b])
        self._ready[b] = False
        self._issue_copy(1 - b)       # the other buffer's last reader finished before the previous step's loss read
        loss, _, _ = self._run(self.sx2[b], self.sy2[b])
        self.host_loss.copy_(loss, non_blocking=True)
        main.synchronize()            # the caller reads the loss (ref: loss.item(), :72)
        self._cur = 1 - b

    def config(self):
        return {"workload": self.name, "B": self.B, "D": self.D, "compute": "bf16 tcgen05, fp32 accumulate",
                "sharding": f"rows/{self.dist.world} + all-gather" if self.dist.on else "single GPU",
                "l2": "no flush: per-step working set ~300 MB (fp32 in 67, bf16 packed 34, fp32 grads 67, column-sum "
                      "partials 134) > 126 MB L2"}

    def roofline(self, peaks, steps):
        ev = self.events
        out = {}
        for k in ("rowsumexp", "grad"):
            out[k] = sum(a.elapsed_time(b) for a, b in ev[k]) / len(ev[k])
        Ma, B, D = self.Ma, self.B, self.D
        flops = 4.0 * Ma * B * D                         # dX and dY GEMMs of this rank's rows (algorithmic)
        ach = flops / (out["grad"] * 1e-3) / 1e12
        return {"bound": "tensor", "kernel": "tc3_grad_kernel", "achieved": round(ach, 1), "peak": peaks["tc"],
                "unit": "TFLOP/s", "frac": round(ach / peaks["tc"], 4),
                # dram__bytes_read.sum + dram__bytes_write.sum of one launch at B = 65 536 on one GPU
                # (profiles/r01g_tc3_grad_kernel_ncu_full.txt: 49.8 + 82.4 MB); algorithmic bytes: 2 B D (2 + 4) = 100.7 MB
                "traffic": 132.1e6 if (Ma == B and B == 65536) else None,
                "peak_source": f"{peaks['src']} bf16_tflops_sustained",
                "algorithmic_flops_per_launch": flops, "kernel_ms": round(out["grad"], 4),
                "other_kernels_ms": {"tc2_sumexp_kernel (+2 small reduces)": round(out["rowsumexp"], 4)}}

    def cpu_sample(self, budget_s=12.0):
        """Oracle port on the host cores: rows [0, n) of the same B = 65 536 workload in fp32 (n / B of one
        fwd+bwd), n chosen so that the sample takes about `budget_s` seconds."""
        from oracle import restate as R
        from pclip import synth
        torch.set_num_threads(os.cpu_count() or 1)
        x, y = synth.clip_scaled_embeddings(self.B, self.D, cfg_id=4)
        t0 = time.perf_counter(); R.clip_rowblock_fwd_bwd(x, y, 0, 256); t1 = time.perf_counter() - t0
        t0 = time.perf_counter(); R.clip_rowblock_fwd_bwd(x, y, 256, 256); t1 = min(t1, time.perf_counter() - t0)
        n = int(max(256, min(self.B, budget_s / t1 * 256)) // 256 * 256)
        t0 = time.perf_counter()
        for i0 in range(0, n, 1024):
            R.clip_rowblock_fwd_bwd(x, y, i0, min(1024, n - i0))
        dt = time.perf_counter() - t0
        return {"value": round(n / dt, 1), "unit": self.unit, "cores": torch.get_num_threads(), "kind": "port",
                "sample": f"oracle/restate.py clip_rowblock_fwd_bwd, fp32, rows [0,{n}) of B={self.B} against all "
                          f"{self.B} columns in 1024-row blocks ({n}/{self.B} of one fwd+bwd) in {dt:.2f} s"}


class ClipGradCache:
    """configs[1]: _contrastive_loss_gc on 16 + 16 GradCache leaves of [16, 128] with grad_output = 65536
    (ref: modules/training_utils.py:137-139, :247-265), through the autograd entry point a user calls."""
    name = "clip_gc256"

    def __init__(self, dist: Dist):
        from pclip import ops, synth
        self.ops = ops
        xs, ys = synth.gradcache_leaves()
        self.xs = [t.to(dist.dev).requires_grad_() for t in xs]
        self.ys = [t.to(dist.dev).requires_grad_() for t in ys]

    def step(self):
        for t in self.xs + self.ys:
            t.grad = None
        loss = self.ops.clip_loss(self.xs, self.ys, compute=self.ops.BF16)
        (loss * 65536.0).backward()

    def report(self, dist, steps=200, warmup=20):
        self.ops.reset_launch_count()
        ms = timed(dist, self.step, steps, warmup)
        return {"us_per_step": round(ms / steps * 1e3, 2), "pairs_per_s": round(256 * steps / (ms * 1e-3), 1),
                "pclip_launches_per_step": self.ops.launch_count() / (steps + warmup),
                "note": "launch-latency bound (50 MFLOP): roofline fraction not meaningful; includes autograd overhead"}


class FilipB256:
    """configs[2]: FILIP token-wise late interaction fwd+bwd (normalise -> similarity -> row-wise InfoNCE -> backward),
    B = 256, T = 512/512 (ref: modules/models.py:238-239, :292-356; training_utils_filip.py:26-39), bf16 tensor cores."""
    name = "filip_b256"

    def __init__(self, dist: Dist, B=256, T=512, ragged=False):
        from pclip import ops, synth
        self.ops, self.B, self.T, self.ragged = ops, B, T, ragged
        hA, hB, mA, mB, tau = synth.filip_tokens(B, T, T, cfg_id=3, ragged=ragged)
        self.hA, self.hB = hA.to(dist.dev).requires_grad_(), hB.to(dist.dev).requires_grad_()
        self.lenA = mA.sum(1).to(torch.int32).to(dist.dev)
        self.lenB = mB.sum(1).to(torch.int32).to(dist.dev)
        self.tau = tau.to(dist.dev).requires_grad_()
        self.pair_tokens = float((mA.sum(1).double()[:, None] * mB.sum(1).double()[None, :]).sum())

    def step(self):
        ops = self.ops
        self.hA.grad = self.hB.grad = self.tau.grad = None
        nA = ops.normalize_scale(self.hA, None, eps=0.0)
        nB = ops.normalize_scale(self.hB, None, eps=0.0)
        simA, simB = ops.filip_similarity(nA, nB, self.lenA, self.lenB, self.tau, compute=ops.BF16)
        ops.filip_loss(simA, simB).backward()

    def report(self, dist, peaks, steps=5, warmup=2):
        self.ops.reset_launch_count()
        ms = timed(dist, self.step, steps, warmup) / steps
        flops = 2.0 * self.pair_tokens * 128              # the token-similarity contraction (algorithmic, once)
        return {"ms_per_step": round(ms, 3), "pairs_per_s": round(self.B / (ms * 1e-3), 1),
                "lengths": "PDB-like ragged" if self.ragged else "all 512",
                "algorithmic_tflops": round(flops / (ms * 1e-3) / 1e12, 1),
                "frac_of_tensor_peak": round(flops / (ms * 1e-3) / 1e12 / peaks["tc"], 4),
                "pclip_launches_per_step": self.ops.launch_count() / (steps + warmup)}


class Topk4M:
    """configs[4]: top-10 by cosine over a 2^22-row bank (ref: rank / top-k of training_utils.py:219-222, topk.py:47-61),
    bf16 tensor-core candidate selection + exact fp32 re-scoring; queries stream through in batches of Q."""
    name = "topk_4m"

    def __init__(self, dist: Dist, N=1 << 22, Q=1 << 18, k=10, q_total=None):
        from pclip import ops, synth
        self.ops, self.N, self.Q, self.k = ops, N, Q, k
        self.world = dist.world if q_total else 1
        if q_total:      # query-sharded: the same seeded bank and query set on every rank, this rank's slice of the queries
            self.bank, q, pos = synth.retrieval_bank(N, q_total, cfg_id=5, device=dist.dev)
            self.q, self.pos = q[dist.rank * Q:(dist.rank + 1) * Q].contiguous(), pos[dist.rank * Q:(dist.rank + 1) * Q]
            del q
        else:
            self.bank, self.q, self.pos = synth.retrieval_bank(N, Q, cfg_id=5, device=dist.dev)
        self.plan = ops.TopkPlan(Q, N, 128, k, ops.BF16, dist.dev)

    def step(self):
        self.plan.run(self.q, self.bank)

    def report(self, dist, peaks, steps=3, warmup=1):
        self.ops.reset_launch_count()
        ms = timed(dist, self.step, steps, warmup) / steps
        Qall = self.Q * self.world                     # whole-job queries per step (all ranks)
        flops = 2.0 * Qall * self.N * 128
        recall = float((self.plan.idx[:, 0] == self.pos).float().mean())
        return {"ms_per_step": round(ms, 3), "queries_per_s": round(Qall / (ms * 1e-3), 1), "Q_per_step": Qall,
                "n_gpus": self.world,
                "N": self.N, "k": self.k, "algorithmic_tflops": round(flops / (ms * 1e-3) / 1e12, 1),
                "frac_of_tensor_peak": round(flops / (ms * 1e-3) / 1e12 / peaks["tc"] / self.world, 4),
                "unproven_queries": int(self.plan.n_unsafe), "top1_is_planted_positive": round(recall, 4),
                "pclip_launches_per_step": self.ops.launch_count() / (steps + warmup)}


def run_reference(args):
    """--impl reference: the reference's CPU arithmetic for the path (oracle port: the reference is Python and
    cannot travel to the GPU box) on all host cores, each step a bounded sample of the same workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import restate as R
    from pclip import synth
    torch.set_num_threads(os.cpu_count() or 1)
    B, D, n = 65536, 128, 2048
    x, y = synth.clip_scaled_embeddings(B, D, cfg_id=4)
    for _ in range(args.warmup):
        R.clip_rowblock_fwd_bwd(x, y, 0, 256)
    t0 = time.perf_counter()
    for k in range(args.steps):
        for i0 in range(0, n, 1024):
            R.clip_rowblock_fwd_bwd(x, y, (k * n + i0) % B, 1024)
    dt = time.perf_counter() - t0
    v = round(n * args.steps / dt, 1)
    sample = (f"oracle/restate.py clip_rowblock_fwd_bwd fp32: each step = {n} rows of B={B} against all {B} columns "
              f"({n}/{B} of one fwd+bwd)")
    print(json.dumps({
        "impl": "reference", "metric": ClipLarge.metric, "value": v, "unit": ClipLarge.unit, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(dt / args.steps * 1e3, 3),
        "higher_is_better": True, "scaling": ClipLarge.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": ClipLarge.name, "B": B, "D": D, "sample_rows_per_step": n},
        "cpu_baseline": {"value": v, "unit": ClipLarge.unit, "cores": torch.get_num_threads(), "kind": "port",
                         "sample": sample},
        "e2e": {"value": v, "unit": ClipLarge.unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="clip_b65536")
    ap.add_argument("--no-extras", action="store_true", help="skip the secondary workloads and the CPU baseline")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: protein-clip_b200 has no CPU path")
    dist = Dist(args.gpus)
    peaks = load_peaks()
    from pclip import ops
    w = ClipLarge(dist)
    # ---- device-resident leg: `value` ----
    for _ in range(args.warmup):
        w.step()
    ops.reset_launch_count()
    w.record = True
    with ClockSampler(dist.local) as clk:
        ms = timed(dist, w.step, args.steps, 0)
    w.record = False
    launches = dist.sum_int(ops.launch_count())
    value = w.units * args.steps / (ms * 1e-3)
    roof = w.roofline(peaks, args.steps)
    # ---- end-to-end leg: host buffers, H2D + D2H inside the timed region ----
    ms_e2e = timed(dist, w.e2e_step, args.steps, max(3, args.warmup // 2), begin=getattr(w, "e2e_begin", None))
    e2e = w.units * args.steps / (ms_e2e * 1e-3)
    out = {
        "metric": w.metric, "value": round(value, 1), "unit": w.unit, "n_gpus": dist.world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(ms / args.steps, 4), "higher_is_better": True,
        "scaling": w.scaling, "vs_baseline": None, "dtype": w.dtype, "data": "synthetic", "config": w.config(),
        "clocks": clk.summary(), "roofline": roof,
        "e2e": {"value": round(e2e, 1), "unit": w.unit, "h2d_bytes_per_step": w.h2d * dist.world,
                "d2h_bytes_per_step": w.d2h * dist.world, "ms_per_step": round(ms_e2e / args.steps, 4),
                "pipeline": "H2D of step i+1 on a copy stream under step i (2 staging buffers); K copies inside the K "
                            "timed steps" if hasattr(w, "e2e_begin") else "serial"},
        "gpu_launches": launches,
        "step_tflops_algorithmic": round(6.0 * w.B * w.B * w.D / (ms / args.steps * 1e-3) / 1e12, 1),
        "step_frac_of_tensor_peak": round(6.0 * w.B * w.B * w.D / (ms / args.steps * 1e-3) / 1e12 / peaks["tc"]
                                          / max(args.gpus, 1), 4),   # whole-job flops over N GPUs' peak
    }
    if dist.rank == 0 and dist.world == 1 and not args.no_extras:
        del w.plan, w.job
        extras = {"clip_gc256": ClipGradCache(dist).report(dist)}
        for name, make in (("filip_b256", lambda: FilipB256(dist)), ("filip_b256_ragged", lambda: FilipB256(dist, ragged=True)),
                           ("topk_4m", lambda: Topk4M(dist))):
            try:
                wl = make()
                extras[name] = wl.report(dist, peaks)
                del wl
            except Exception as e:  # a secondary workload must not take the headline line down
                extras[name] = {"error": f"{type(e).__name__}: {e}"[:300]}
            torch.cuda.empty_cache()
        out["workloads"] = extras
        out["cpu_baseline"] = w.cpu_sample()
    elif not args.no_extras:
        # N > 1: configs[4] with the queries sharded over a replicated bank (pclip.sharded.ShardedTopk "queries" mode:
        # independent units, no collective).  Every rank runs it; the time is the max over ranks.
        del w.plan, w.job
        torch.cuda.empty_cache()
        try:
            wl = Topk4M(dist, Q=(1 << 20) // dist.world, q_total=1 << 20)
            extras = {"topk_4m_query_sharded": wl.report(dist, peaks)}
            del wl
        except Exception as e:
            extras = {"topk_4m_query_sharded": {"error": f"{type(e).__name__}: {e}"[:300]}}
        out["workloads"] = extras
        out["cpu_baseline"] = None
    elif dist.rank == 0:
        out["cpu_baseline"] = None
    if dist.rank == 0:
        print(json.dumps(out), flush=True)
    dist.close()


if __name__ == "__main__":
    main()
