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
    clip_b16     configs[0]: encoder tail + CLIP loss at batch 16, fp32 (the reference's own CPU-runnable case)
The non-default workloads are also run briefly by the default invocation and reported under "workloads", each with its
own cpu_baseline (the oracle port of the reference arithmetic on the host cores), an end-to-end leg with host buffers
and a roofline block for its dominant kernel.  Every line carries `check`: the loss and a 64-row gradient checksum of
the seeded input, so that the N = 1, 2, 4, 8 lines of a scaling run can be compared with each other; N > 1 runs add
`multirank_parity` (sharded CLIP / FILIP / top-k against the single-GPU result on rank 0's GPU).
    python bench.py --selftest        (under torch.distributed.run) runs only that multi-rank parity leg.
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


def profile_traffic(kernel: str):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of `kernel` from the newest committed ncu summary
    under profiles/ (files are named per round, so the lexicographically last one is the newest).  (bytes, file)."""
    import glob
    import re
    best = None
    for f in sorted(glob.glob(os.path.join(REPO, "profiles", f"*{kernel}*ncu_full.txt"))):
        txt = open(f).read()
        m = {k: float(v) for k, u, v in re.findall(r"(dram__bytes_(?:read|write)\.sum) \[(\w+)\] = ([0-9.]+)", txt)
             if u == "Mbyte"}
        if len(m) == 2:
            best = (sum(m.values()) * 1e6, os.path.relpath(f, REPO))
    return best or (None, None)


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
            if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":   # keeps NCCL's banner off stdout (one JSON line)
                os.environ["NCCL_DEBUG"] = "WARN"
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
        self.graph, self.graph_error = None, None
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
        if self.graph is not None and not self.record:
            self.graph.replay()
        else:
            self._run(self.x, self.y)

    def enable_graph(self):
        """N > 1: replay the whole sharded step (pack, 2 NCCL collectives, peer copies, sweep, finalize, gradient) as ONE
        CUDA graph launch: at 8 GPUs the host-side launches of a 0.7 ms step otherwise leave the GPU idle between kernels."""
        if self.job is None or os.environ.get("PCLIP_BENCH_GRAPH", "1") == "0":
            return False
        try:
            self.graph = self.job.capture([self.x], [self.y])
            self.graph_e2e = None
            return True
        except Exception as e:      # capture is an optimisation: fall back to eager launches, say so in the JSON
            self.graph = None
            self.graph_error = f"{type(e).__name__}: {e}"[:200]
            return False

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
        if self.graph is not None:    # N > 1: one captured graph per staging buffer (the graph reads the buffer in place)
            if getattr(self, "graph_e2e", None) is None:
                self.graph_e2e = [self.job.capture_graph([self.sx2[i]], [self.sy2[i]]) for i in (0, 1)]
            self.graph_e2e[b][0].replay()
            loss = self.graph_e2e[b][1][0]
        else:
            rec, self.record = self.record, False
            loss, _, _ = self._run(self.sx2[b], self.sy2[b])
            self.record = rec
        self.host_loss.copy_(loss, non_blocking=True)
        main.synchronize()            # the caller reads the loss (ref: loss.item(), :72)
        self._cur = 1 - b

    def config(self):
        return {"workload": self.name, "B": self.B, "D": self.D, "compute": "bf16 tcgen05, fp32 accumulate",
                "sharding": (f"rows/{self.dist.world}; Y all-gather (NCCL) in front of the forward sweep, X "
                             + ("pulled from peer memory by the copy engines under it" if self.job.pull is not None
                                else "in the same NCCL launch")) if self.dist.on else "single GPU",
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
                # dram__bytes_read.sum + dram__bytes_write.sum of one launch at B = 65 536 on one GPU, parsed from the
                # committed ncu capture; algorithmic bytes: 2 B D (2 + 4) = 100.7 MB
                "traffic": profile_traffic("tc3_grad_kernel")[0] if (Ma == B and B == 65536) else None,
                "traffic_source": profile_traffic("tc3_grad_kernel")[1] if (Ma == B and B == 65536) else None,
                "peak_source": f"{peaks['src']} bf16_tflops_sustained",
                "algorithmic_flops_per_launch": flops, "kernel_ms": round(out["grad"], 4),
                "other_kernels_ms": {"tc2_sumexp_kernel (+2 small reduces)": round(out["rowsumexp"], 4)}}

    def check(self):
        """Loss and gradient checksums of 64 fixed global rows (every B/64-th): the seeded input makes them constants
        that every N must reproduce (bf16 tensor-core arithmetic: compare to ~1e-5 relative)."""
        torch.cuda.synchronize()
        B, Ma, r = self.B, self.Ma, self.dist.rank
        rows = torch.arange(0, B, B // 64)
        mine = rows[(rows >= r * Ma) & (rows < (r + 1) * Ma)] - r * Ma
        v = torch.zeros(4, dtype=torch.float64, device=self.dist.dev)
        if len(mine):
            dx, dy = self.plan.dx[mine.to(self.dist.dev)].double(), self.plan.dy[mine.to(self.dist.dev)].double()
            v[0], v[1], v[2], v[3] = dx.sum(), dx.abs().sum(), dy.sum(), dy.abs().sum()
        if self.dist.on:
            self.dist.dist.all_reduce(v)
        return {"loss": float(self.plan.loss[0]), "dx64_sum": float(v[0]), "dx64_abs_sum": float(v[1]),
                "dy64_sum": float(v[2]), "dy64_abs_sum": float(v[3]), "rows": "every 1024th of 65 536"}

    def oracle_loss(self):
        """float64-accumulated CPU value of the loss on the bf16-rounded seeded input (two chunked passes over the
        65 536^2 logits, ref: modules/training_utils.py:85-89): the constant `check.loss` must match."""
        from oracle import restate as R
        from pclip import synth
        torch.set_num_threads(os.cpu_count() or 1)
        x, y = synth.clip_scaled_embeddings(self.B, self.D, cfg_id=4)
        x, y = R.bf16_round(x), R.bf16_round(y)
        Rs = torch.empty(self.B, dtype=torch.float64)
        Cs = torch.zeros(self.B, dtype=torch.float64)
        yt = y.t().contiguous()
        for i0 in range(0, self.B, 2048):
            e = (x[i0:i0 + 2048] @ yt).exp_()
            Rs[i0:i0 + 2048] = e.sum(1, dtype=torch.float64)
            Cs += e.sum(0, dtype=torch.float64)
        diag = (x.double() * y.double()).sum(-1)
        return float((Rs.log() + Cs.log() - 2 * diag).sum() / (2 * self.B))

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


def cpu_time(fn, reps=3):
    """best-of-`reps` wall seconds of fn() on the host cores after one warm-up call"""
    torch.set_num_threads(os.cpu_count() or 1)
    fn()
    best = float("inf")
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        best = min(best, time.perf_counter() - t0)
    return best


def e2e_block(value, unit, h2d, d2h, ms):
    return {"value": round(value, 1), "unit": unit, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
            "ms_per_step": round(ms, 4)}


class ClipSmall:
    """configs[0]: the reference's own default step on the CPU -- encoder tail (F.normalize * exp(tau/2), ref:
    modules/models.py:50-51) + _contrastive_loss(pep, rec.t()) (ref: modules/training_utils.py:69, :74-89) at
    B = 16, D = 128, fp32, forward + backward through the autograd entry points."""
    name = "clip_b16"

    def __init__(self, dist: Dist):
        from pclip import ops, synth
        self.ops, self.dev = ops, dist.dev
        e_pep, e_rec, tau = synth.clip_raw_embeddings(16, 128, cfg_id=1)
        self.h = [e_pep.pin_memory(), e_rec.pin_memory()]
        self.host_loss = torch.empty(1).pin_memory()
        self.e_pep, self.e_rec = e_pep.to(dist.dev).requires_grad_(), e_rec.to(dist.dev).requires_grad_()
        self.tau = tau.to(dist.dev).requires_grad_()
        self.cpu_in = (e_pep, e_rec, tau)

    def _fwd_bwd(self, a, b):
        ops = self.ops
        a.grad = b.grad = self.tau.grad = None
        pep, rec = ops.normalize_scale(a, self.tau), ops.normalize_scale(b, self.tau)
        loss = ops.clip_loss(pep, rec.t().t(), compute=ops.F32)   # rec.t() is what the reference passes; .t() again = the view
        loss.backward()
        return loss

    def step(self):
        self._fwd_bwd(self.e_pep, self.e_rec)

    def e2e_step(self):
        a = self.h[0].to(self.dev, non_blocking=True).requires_grad_()
        b = self.h[1].to(self.dev, non_blocking=True).requires_grad_()
        self.host_loss.copy_(self._fwd_bwd(a, b).detach().reshape(1), non_blocking=True)
        torch.cuda.current_stream().synchronize()

    def cpu_step(self):
        from oracle import restate as R
        e_pep, e_rec, tau = (t.clone().requires_grad_() for t in self.cpu_in)
        loss = R.clip_loss(R.normalize_scale(e_pep, tau), R.normalize_scale(e_rec, tau).t())
        loss.backward()

    def report(self, dist, steps=200, warmup=20):
        self.ops.reset_launch_count()
        ms = timed(dist, self.step, steps, warmup) / steps
        launches = self.ops.launch_count() / (steps + warmup)
        ms_e = timed(dist, self.e2e_step, steps, warmup) / steps
        t = cpu_time(self.cpu_step, reps=20)
        return {"us_per_step": round(ms * 1e3, 2), "pairs_per_s": round(16 / (ms * 1e-3), 1), "dtype": "f32",
                "pclip_launches_per_step": launches,
                "e2e": e2e_block(16 / (ms_e * 1e-3), "pairs/s", 2 * 16 * 128 * 4, 4, ms_e),
                "cpu_baseline": {"value": round(16 / t, 1), "unit": "pairs/s", "cores": torch.get_num_threads(),
                                 "kind": "port", "sample": "oracle/restate.py normalize_scale + clip_loss under torch "
                                 f"autograd, full size (B = 16), best of 20: {t * 1e6:.0f} us"},
                "roofline": {"bound": "launch latency", "note": "0.2 MFLOP and 16 KB per step: neither roofline applies; "
                             "the step is the launch + autograd overhead of 7 kernels"}}


class EncoderHead:
    """the encoder head behind a (stand-in) ESM trunk in evaluation mode -- projection 640 -> 128,
    per-residue FFN, masked mean, pooled FFN, normalise, temperature (ref: modules/models.py:46-51) -- at the scripts'
    sub-batch (16 sequences, depth 2 / 2) padded to 512 residues with PDB-like lengths: the fused two-launch path
    (pclip_encoder_head_fwd) beside the stock nn.Sequential path on the same GPU and the oracle on the host cores."""
    name = "encoder_head_b16"

    def __init__(self, dist: Dist, Bs=16, T=512, Din=640):
        import types
        from modules import models
        from pclip import ops, synth
        self.ops, self.dev, self.Bs, self.T, self.Din = ops, dist.dev, Bs, T, Din
        g = torch.Generator().manual_seed(synth.SEED + 8)
        hidden = torch.randn(Bs, T, Din, generator=g)
        lens = torch.clamp(synth.receptor_lengths(Bs, g, cap=2000) + 2, max=T)
        self.cpu = (hidden, synth.prefix_mask(lens, T), torch.tensor(1.0))
        self.h_hidden = hidden.pin_memory()
        self.hidden, self.mask, self.tau = hidden.to(dist.dev), self.cpu[1].to(dist.dev), self.cpu[2].to(dist.dev)

        class Trunk(torch.nn.Module):
            def forward(s, input_ids=None, attention_mask=None):
                return types.SimpleNamespace(last_hidden_state=self.cur)
        torch.manual_seed(synth.SEED + 8)
        self.enc = models.Encoder(Din, 128, 2, 2, 0.1, Trunk()).to(dist.dev).eval()
        self.sd = {k: v.detach().cpu() for k, v in self.enc.state_dict().items()}
        self.cur = self.hidden
        self.host_out = torch.empty(Bs, 128).pin_memory()

    def _seq(self):
        return {"input_ids": self.mask, "attention_mask": self.mask, "temperature": self.tau}

    def step(self):
        with torch.no_grad():
            return self.enc(self._seq())

    def e2e_step(self):
        self.cur = self.h_hidden.to(self.dev, non_blocking=True)
        self.host_out.copy_(self.step(), non_blocking=True)
        torch.cuda.current_stream().synchronize()
        self.cur = self.hidden

    def report(self, dist, steps=100, warmup=10):
        from oracle import restate as R
        self.ops.reset_launch_count()
        ms = timed(dist, self.step, steps, warmup) / steps
        launches = self.ops.launch_count() / (steps + warmup)
        os.environ["PCLIP_FUSED_HEAD"] = "0"
        try:
            ms_stock = timed(dist, self.step, steps, warmup) / steps
            stock = self.step()
        finally:
            del os.environ["PCLIP_FUSED_HEAD"]
        fused = self.step()
        ms_e = timed(dist, self.e2e_step, steps, warmup) / steps
        want = R.encoder_head(self.cpu[0].double(), self.cpu[1], {k: v.double() for k, v in self.sd.items()},
                              self.cpu[2].double())
        t = cpu_time(lambda: R.encoder_head(self.cpu[0], self.cpu[1], self.sd, self.cpu[2]), reps=5)
        err = lambda o: float((o.double().cpu() - want).abs().max() / want.abs().max())   # noqa: E731
        n_res = int(self.cpu[1].sum())
        flop = 2.0 * self.Bs * self.T * 128 * (self.Din + 2 * 128)
        return {"us_per_step": round(ms * 1e3, 2), "sequences_per_s": round(self.Bs / (ms * 1e-3), 1), "dtype": "f32",
                "shape": {"sequences": self.Bs, "padded_residues": self.T, "valid_residues": n_res, "Din": self.Din},
                "pclip_launches_per_step": launches,
                "stock_torch_path_same_gpu_us": round(ms_stock * 1e3, 2),
                "check": {"fused_vs_float64_oracle": err(fused), "stock_vs_float64_oracle": err(stock)},
                "e2e": e2e_block(self.Bs / (ms_e * 1e-3), "sequences/s", self.h_hidden.numel() * 4, self.Bs * 128 * 4, ms_e),
                "cpu_baseline": {"value": round(self.Bs / t, 1), "unit": "sequences/s", "cores": torch.get_num_threads(),
                                 "kind": "port", "sample": f"oracle/restate.py encoder_head, fp32, full size, best of 5: "
                                 f"{t * 1e3:.2f} ms"},
                "roofline": {"bound": "hbm", "achieved": round(self.hidden.numel() * 4 / (ms * 1e-3) / 1e9, 1),
                             "unit": "GB/s", "note": f"reads the {self.hidden.numel() * 4 / 1e6:.0f} MB of hidden states once; "
                             f"{flop / 1e9:.2f} GFLOP fp32 on the CUDA cores ({flop / (ms * 1e-3) / 1e12:.1f} TFLOP/s); at 16 "
                             "sequences the grid is 256 CTAs of which the all-pad ones exit at once"}}


class ClipGradCache:
    """configs[1]: _contrastive_loss_gc on 16 + 16 GradCache leaves of [16, 128] with grad_output = 65536
    (ref: modules/training_utils.py:137-139, :247-265), through the autograd entry point a user calls."""
    name = "clip_gc256"

    def __init__(self, dist: Dist):
        from pclip import ops, synth
        self.ops, self.dev = ops, dist.dev
        xs, ys = synth.gradcache_leaves()
        self.cpu_leaves = (xs, ys)
        self.hx, self.hy = torch.cat(xs).pin_memory(), torch.cat(ys).pin_memory()
        self.host_loss = torch.empty(1).pin_memory()
        self.xs = [t.to(dist.dev).requires_grad_() for t in xs]
        self.ys = [t.to(dist.dev).requires_grad_() for t in ys]
        self.scale = torch.tensor(65536.0, device=dist.dev)

    def _fwd_bwd(self, xs, ys):
        """What modules.training_utils.train_gc runs per big batch: loss + every leaf's .grad (times GradScaler's
        device-side scale) from ONE launch (pclip.ops.clip_loss_gc_direct)."""
        for t in xs + ys:
            t.grad = None
        return self.ops.clip_loss_gc_direct(xs, ys, self.scale)

    def _autograd(self):
        """the same step through torch.autograd (ops.clip_loss + backward over the 32 leaves)"""
        for t in self.xs + self.ys:
            t.grad = None
        loss = self.ops.clip_loss(self.xs, self.ys, compute=self.ops.BF16)
        (loss * 65536.0).backward()

    def step(self):
        self._fwd_bwd(self.xs, self.ys)

    def e2e_step(self):
        x = self.hx.to(self.dev, non_blocking=True)
        y = self.hy.to(self.dev, non_blocking=True)
        xs = [t.requires_grad_() for t in x.split(16)]       # the 16 + 16 leaves GradCache hands to the loss
        ys = [t.requires_grad_() for t in y.split(16)]
        self.host_loss.copy_(self._fwd_bwd(xs, ys).detach().reshape(1), non_blocking=True)
        torch.cuda.current_stream().synchronize()

    def cpu_step(self):
        from oracle import restate as R
        xs = [t.clone().requires_grad_() for t in self.cpu_leaves[0]]
        ys = [t.clone().requires_grad_() for t in self.cpu_leaves[1]]
        (R.clip_loss_gc(xs, ys) * 65536.0).backward()

    def report(self, dist, steps=200, warmup=20):
        self.ops.reset_launch_count()
        ms = timed(dist, self.step, steps, warmup) / steps
        launches = self.ops.launch_count() / (steps + warmup)
        ms_e = timed(dist, self.e2e_step, steps, warmup) / steps
        ms_a = timed(dist, self._autograd, steps, warmup) / steps
        t = cpu_time(self.cpu_step, reps=20)
        return {"us_per_step": round(ms * 1e3, 2), "pairs_per_s": round(256 / (ms * 1e-3), 1),
                "dtype": "f32 arithmetic (one-launch small-batch kernel)",
                "pclip_launches_per_step": launches,
                "through_torch_autograd_us_per_step": round(ms_a * 1e3, 2),
                "e2e": e2e_block(256 / (ms_e * 1e-3), "pairs/s", 2 * 256 * 128 * 4, 4, ms_e),
                "cpu_baseline": {"value": round(256 / t, 1), "unit": "pairs/s", "cores": torch.get_num_threads(),
                                 "kind": "port", "sample": "oracle/restate.py clip_loss_gc (cat of 16 + 16 leaves, fp32) "
                                 f"under torch autograd with grad_output 65536, full size, best of 20: {t * 1e6:.0f} us"},
                "roofline": {"bound": "launch latency", "note": "50 MFLOP per step: roofline fraction not meaningful; "
                             "the figure of merit is us per step through the autograd API"}}


class FilipB256:
    """configs[2]: FILIP token-wise late interaction fwd+bwd (normalise -> similarity -> row-wise InfoNCE -> backward),
    B = 256, T = 512/512 (ref: modules/models.py:238-239, :292-356; training_utils_filip.py:26-39), bf16 tensor cores."""
    name = "filip_b256"

    def __init__(self, dist: Dist, B=256, T=512, ragged=False, sharded=False):
        from pclip import ops, synth
        self.ops, self.B, self.T, self.ragged, self.dist = ops, B, T, ragged, dist
        hA, hB, mA, mB, tau = synth.filip_tokens(B, T, T, cfg_id=3, ragged=ragged)
        self.pair_tokens = float((mA.sum(1).double()[:, None] * mB.sum(1).double()[None, :]).sum())
        self.lens_cpu = (mA.sum(1), mB.sum(1))
        self.job = None
        if sharded and dist.on:      # A-row sharded over the ranks (pclip.sharded.ShardedFilip)
            from pclip import sharded as sh
            Bl = B // dist.world
            sl = slice(dist.rank * Bl, (dist.rank + 1) * Bl)
            hA, hB, mA, mB = hA[sl], hB[sl], mA[sl], mB[sl]
            self.job = sh.ShardedFilip(compute=ops.BF16)
        self.hostA, self.hostB = hA.contiguous().pin_memory(), hB.contiguous().pin_memory()
        self.hA, self.hB = self.hostA.to(dist.dev).requires_grad_(), self.hostB.to(dist.dev).requires_grad_()
        self.mA, self.mB = mA.to(dist.dev), mB.to(dist.dev)
        self.lenA = mA.sum(1).to(torch.int32).to(dist.dev)
        self.lenB = mB.sum(1).to(torch.int32).to(dist.dev)
        self.tau = tau.to(dist.dev).requires_grad_()
        self.host_loss = torch.empty(1).pin_memory()
        self.ev = None

    def _fwd_bwd(self, hA, hB):
        ops = self.ops
        hA.grad = hB.grad = self.tau.grad = None
        if self.job is not None:
            loss, _, _, _ = self.job.step(hA, hB, self.mA, self.mB, self.tau, normalize=True)
            return loss
        ev = self.ev
        nA = ops.normalize_scale(hA, None, eps=0.0)
        nB = ops.normalize_scale(hB, None, eps=0.0)
        if ev: ev[0].record()
        simA, simB = ops.filip_similarity(nA, nB, self.lenA, self.lenB, self.tau, compute=ops.BF16)
        if ev: ev[1].record()
        loss = ops.filip_loss(simA, simB)
        if ev: ev[2].record()
        loss.backward()
        if ev: ev[3].record()
        return loss

    def step(self):
        self._fwd_bwd(self.hA, self.hB)

    def e2e_step(self):
        a = self.hostA.to(self.dist.dev, non_blocking=True).requires_grad_()
        b = self.hostB.to(self.dist.dev, non_blocking=True).requires_grad_()
        self.host_loss.copy_(self._fwd_bwd(a, b).detach().reshape(1), non_blocking=True)
        torch.cuda.current_stream().synchronize()

    def cpu_baselines(self):
        """The reference arithmetic (oracle port, fp32, torch autograd like the reference) on the host cores:
        at the reference's own batch sizes (main_2protein_filip.py:33 uses 10) and, for B = 256, on a sample of A rows
        against all 256 B samples (both losses are row-wise, so rows are independent units of work)."""
        from oracle import restate as R
        from pclip import synth
        torch.set_num_threads(os.cpu_count() or 1)
        out = {}

        def run(hA, hB, mA, mB, tau, rows=None):
            a, b = hA.clone().requires_grad_(), hB.clone().requires_grad_()
            t = tau.clone().requires_grad_()
            nA, nB = R.filip_normalize(a), R.filip_normalize(b)
            sA, sB = R.filip_similarity(nA if rows is None else nA[rows], nB, mA if rows is None else mA[rows], mB, t)
            if rows is None:
                loss = R.filip_compute_loss(sA, sB)
            else:   # the sampled rows' terms of the two row-wise losses
                idx = torch.as_tensor(rows)
                loss = sum((-s[torch.arange(len(rows)), idx] + torch.logsumexp(s, 1)).sum() for s in (sA, sB)) / (2 * len(mB))
            loss.backward()

        for b in (10, 16, 32):
            hA, hB, mA, mB, tau = synth.filip_tokens(b, self.T, self.T, cfg_id=3, ragged=self.ragged)
            t0 = time.perf_counter(); run(hA, hB, mA, mB, tau); dt = time.perf_counter() - t0
            out[f"B{b}"] = {"pairs_per_s": round(b / dt, 2), "seconds": round(dt, 3)}
        hA, hB, mA, mB, tau = synth.filip_tokens(self.B, self.T, self.T, cfg_id=3, ragged=self.ragged)
        rows = [0, 64, 128, 192]
        t0 = time.perf_counter(); run(hA, hB, mA, mB, tau, rows); dt = time.perf_counter() - t0
        v = len(rows) / dt
        return {"value": round(v, 2), "unit": "pairs/s", "cores": torch.get_num_threads(), "kind": "port",
                "sample": f"oracle/restate.py filip_normalize + filip_similarity + row-wise InfoNCE under torch autograd, "
                          f"fp32: {len(rows)} of the {self.B} A samples against all {self.B} B samples "
                          f"({len(rows)}/{self.B} of one fwd+bwd; normalise of all tokens included) in {dt:.2f} s",
                "reference_batch_sizes": out}

    def report(self, dist, peaks, steps=5, warmup=3, cpu=True):
        self.ops.reset_launch_count()
        ms = timed(dist, self.step, steps, warmup) / steps
        launches = self.ops.launch_count() / (steps + warmup)
        flops = 2.0 * self.pair_tokens * 128              # the token-similarity contraction (algorithmic, once)
        out = {"ms_per_step": round(ms, 3), "pairs_per_s": round(self.B / (ms * 1e-3), 1),
               "lengths": "PDB-like ragged" if self.ragged else "all 512", "n_gpus": dist.world if self.job else 1,
               "algorithmic_tflops": round(flops / (ms * 1e-3) / 1e12, 1),
               "frac_of_tensor_peak": round(flops / (ms * 1e-3) / 1e12 / peaks["tc"] / (dist.world if self.job else 1), 4),
               "pclip_launches_per_step": launches}
        if self.job is None:
            # forward / backward split and the dominant kernel's stage time, CUDA events on the launching stream
            acc = [0.0, 0.0, 0.0]
            for _ in range(steps):
                self.ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
                self.step()
                torch.cuda.synchronize()
                for i in range(3):
                    acc[i] += self.ev[i].elapsed_time(self.ev[i + 1]) / steps
            self.ev = None
            out["split_ms"] = {"similarity_forward (tc_filip_both_kernel + prep + finalize)": round(acc[0], 3),
                               "loss_tail_forward": round(acc[1], 3),
                               "backward (loss tail, 2 x invert + gather, normalise)": round(acc[2], 3)}
            ach = flops / (acc[0] * 1e-3) / 1e12
            tr, src = profile_traffic("tc_filip_both_kernel")
            out["roofline"] = {"bound": "tensor", "kernel": "tc_filip_both_kernel (one sweep: row and column maxima of every tile)",
                               "achieved": round(ach, 1), "peak": peaks["tc"], "unit": "TFLOP/s",
                               "frac": round(ach / peaks["tc"], 4), "traffic": tr, "traffic_source": src,
                               "algorithmic_flops_per_step": flops, "kernel_ms": round(acc[0], 4),
                               "note": "algorithmic = 2 sum(lenA lenB) D, executed once (the round-1/2 kernels computed S "
                                       "and S^T); the MMA pipeline alone runs this sweep in 3.18 ms (power-limited "
                                       "tensor rate), the epilogue's ALU work is what is left"}
        ms_e = timed(dist, self.e2e_step, steps, warmup) / steps
        bytes_in = (self.hostA.numel() + self.hostB.numel()) * 4 * (dist.world if self.job else 1)
        out["e2e"] = e2e_block(self.B / (ms_e * 1e-3), "pairs/s", bytes_in, 4 * (dist.world if self.job else 1), ms_e)
        if cpu:
            out["cpu_baseline"] = self.cpu_baselines()
        return out


class Topk4M:
    """configs[4]: top-10 by cosine over a 2^22-row bank (ref: rank / top-k of training_utils.py:219-222, topk.py:47-61),
    bf16 tensor-core candidate selection + exact fp32 re-scoring; queries stream through in batches of Q against a
    bank whose bf16 image was prepared once (pclip_topk_prepare_bank)."""
    name = "topk_4m"

    def __init__(self, dist: Dist, N=1 << 22, Q=1 << 18, k=10, q_total=None):
        from pclip import ops, synth
        self.ops, self.N, self.Q, self.k, self.dist = ops, N, Q, k, dist
        self.world = dist.world if q_total else 1
        if q_total:      # query-sharded: the same seeded bank and query set on every rank, this rank's slice of the queries
            self.bank, q, pos = synth.retrieval_bank(N, q_total, cfg_id=5, device=dist.dev)
            self.q, self.pos = q[dist.rank * Q:(dist.rank + 1) * Q].contiguous(), pos[dist.rank * Q:(dist.rank + 1) * Q]
            del q
        else:
            self.bank, self.q, self.pos = synth.retrieval_bank(N, Q, cfg_id=5, device=dist.dev)
        self.plan = ops.TopkPlan(Q, N, 128, k, ops.BF16, dist.dev).prepare_bank(self.bank)

    def step(self):
        self.plan.run(self.q, self.bank)

    def check_sample(self, n=64):
        """`n` queries of the last result against float64 scores over the WHOLE bank (tests/test_gpu_retrieval.py
        check_topk's criteria): scores match, descending, every returned row is a top-k row or within the fp32 rounding
        band of the k-th score, and no row above the band is missing."""
        band = 2e-6
        sel = torch.arange(0, self.Q, self.Q // n, device=self.dist.dev)[:n]
        idx, sc = self.plan.idx[sel], self.plan.score[sel]
        s64 = self.q[sel].double() @ self.bank.double().t()                       # [n, N] float64 on the device
        kth = torch.topk(s64, self.k, dim=1).values[:, -1]
        got = s64.gather(1, idx)
        ok_scores = bool(((got - sc.double()).abs() <= band).all()) and bool((sc[:, :-1] >= sc[:, 1:]).all())
        ok_members = bool((got >= kth[:, None] - 2 * band).all())
        missing = int(((s64 > kth[:, None] + 2 * band).sum(1) >
                       (got > kth[:, None] + 2 * band).sum(1)).sum())
        ties = sc[:, :-1] == sc[:, 1:]
        ok_ties = bool((idx[:, :-1][ties] < idx[:, 1:][ties]).all())
        return {"queries_checked": int(n), "ok": ok_scores and ok_members and missing == 0 and ok_ties,
                "scores_match_f64": ok_scores, "members_in_band": ok_members, "rows_missing": missing,
                "ties_by_lower_index": ok_ties}

    def cpu_baseline(self, nq=1024, chunk=128):
        """torch.topk(q_chunk @ bank.T, k) over the full bank for `nq` queries on the host cores, extrapolated linearly
        in the number of queries (the reference itself only has rank-of-positive)."""
        from oracle import restate as R
        torch.set_num_threads(os.cpu_count() or 1)
        bank, q = self.bank.cpu(), self.q[:nq].cpu()
        R.topk_block(q[:chunk], bank, self.k)
        t0 = time.perf_counter()
        for i0 in range(0, nq, chunk):
            R.topk_block(q[i0:i0 + chunk], bank, self.k)
        dt = time.perf_counter() - t0
        return {"value": round(nq / dt, 1), "unit": "queries/s", "cores": torch.get_num_threads(), "kind": "port",
                "sample": f"oracle/restate.py topk_block (fp32 matmul + torch.topk) on {nq} queries x {self.N} bank rows in "
                          f"{chunk}-query chunks: {dt:.2f} s; queries/s is independent of the query count (linear)"}

    def report(self, dist, peaks, steps=3, warmup=3, cpu=True):
        self.ops.reset_launch_count()
        ms = timed(dist, self.step, steps, warmup) / steps
        launches = self.ops.launch_count() / (steps + warmup)
        Qall = self.Q * self.world                     # whole-job queries per step (all ranks)
        flops = 2.0 * Qall * self.N * 128
        recall = float((self.plan.idx[:, 0] == self.pos).float().mean())
        tr, src = profile_traffic("tc_topk_kernel")
        out = {"ms_per_step": round(ms, 3), "queries_per_s": round(Qall / (ms * 1e-3), 1), "Q_per_step": Qall,
               "n_gpus": self.world, "N": self.N, "k": self.k,
               "algorithmic_tflops": round(flops / (ms * 1e-3) / 1e12, 1),
               "frac_of_tensor_peak": round(flops / (ms * 1e-3) / 1e12 / peaks["tc"] / self.world, 4),
               "unproven_queries": int(self.plan.n_unsafe), "top1_is_planted_positive": round(recall, 4),
               "pclip_launches_per_step": launches,
               "roofline": {"bound": "tensor", "kernel": "tc_topk_kernel", "achieved": round(flops / (ms * 1e-3) / 1e12 / self.world, 1),
                            "peak": peaks["tc"], "unit": "TFLOP/s",
                            "frac": round(flops / (ms * 1e-3) / 1e12 / peaks["tc"] / self.world, 4),
                            "traffic": tr, "traffic_source": src,
                            "note": "whole call (query conversion, candidate kernel, fp32 re-scoring, exact fallback) "
                                    "over the algorithmic 2 Q N D flop"},
               "check": self.check_sample()}
        if self.world == 1:
            # end to end: fp32 queries from pinned host memory, (index, score) lists back to the host
            hq = self.q.cpu().pin_memory()
            hi = torch.empty(self.Q, self.k, dtype=torch.int64).pin_memory()
            hs = torch.empty(self.Q, self.k, dtype=torch.float32).pin_memory()
            dq = torch.empty_like(self.q)

            def e2e():
                dq.copy_(hq, non_blocking=True)
                idx, sc = self.plan.run(dq, self.bank)
                hi.copy_(idx, non_blocking=True)
                hs.copy_(sc, non_blocking=True)
                torch.cuda.current_stream().synchronize()
            ms_e = timed(dist, e2e, steps, 1) / steps
            out["e2e"] = e2e_block(self.Q / (ms_e * 1e-3), "queries/s", hq.numel() * 4, hi.numel() * 8 + hs.numel() * 4, ms_e)
            # small-batch regime: Q = 128 queries per call stream the bf16 bank once -> HBM-bound
            small = self.ops.TopkPlan(128, self.N, 128, self.k, self.ops.BF16, dist.dev).prepare_bank(self.bank)
            qs = self.q[:128].contiguous()
            ms_s = timed(dist, lambda: small.run(qs, self.bank), 20, 3) / 20
            gbs = self.N * 128 * 2 / (ms_s * 1e-3) / 1e9
            out["small_batch_q128"] = {"ms_per_call": round(ms_s, 4), "queries_per_s": round(128 / (ms_s * 1e-3), 1),
                                       "roofline": {"bound": "hbm", "achieved": round(gbs, 1), "peak": peaks["hbm"],
                                                    "unit": "GB/s", "frac": round(gbs / peaks["hbm"], 4),
                                                    "note": "algorithmic bytes = one pass over the prepared bf16 bank "
                                                            "(N D 2 = 1.07 GB); the 128 queries fill half a CTA's query "
                                                            "tile and the bank is split over the SMs"}}
            del small
            if cpu:
                out["cpu_baseline"] = self.cpu_baseline()
        return out


def multirank_parity(dist: Dist):
    """N > 1 ranks over NCCL against the single-GPU result, computed on every rank's own GPU (the content of
    tests/test_gpu_multirank.py, which needs 2 GPUs and is skipped on the driver's 1-GPU test box): sharded CLIP
    (fused one-collective exchange), A-row-sharded FILIP (all-gather + reduce-scatter) and both top-k shardings."""
    from pclip import ops, sharded, synth
    dev, world, rank = dist.dev, dist.world, dist.rank
    out = {}
    # ---- CLIP, bf16 ----
    B = 4096
    x, y = synth.clip_scaled_embeddings(B, 128, cfg_id=31, trained_like=True)
    Ma = B // world
    sl = slice(rank * Ma, (rank + 1) * Ma)
    job = sharded.ShardedClip(B, 128, ops.BF16, dev)
    loss, dx, dy = job.step([x[sl].to(dev)], [y[sl].to(dev)])
    l1, dx1, dy1, _, _ = ops.clip_loss_raw([x.to(dev)], [y.to(dev)], ops.BF16, True)
    out["clip_loss_rel_err"] = abs(float(loss) - float(l1)) / abs(float(l1))
    out["clip_grad_max_rel_err"] = max(float((dx - dx1[sl]).abs().max() / dx1.abs().max()),
                                       float((dy - dy1[sl]).abs().max() / dy1.abs().max()))
    # ---- FILIP, A rows sharded, normalise inside ----
    g = torch.Generator().manual_seed(9)
    Bf, T, D = 8 * world, 70, 128
    hA, hB = torch.randn(Bf, T, D, generator=g), torch.randn(Bf, T, D, generator=g)
    lenA, lenB = torch.randint(3, T + 1, (Bf,), generator=g), torch.randint(3, T + 1, (Bf,), generator=g)
    mA = (torch.arange(T)[None, :] < lenA[:, None]).long()
    mB = (torch.arange(T)[None, :] < lenB[:, None]).long()
    tau = torch.tensor(0.9)
    Bl = Bf // world
    fs = slice(rank * Bl, (rank + 1) * Bl)
    lf, dhA, dhB, dtau = sharded.ShardedFilip().step(hA[fs].to(dev), hB[fs].to(dev), mA[fs].to(dev), mB[fs].to(dev),
                                                     tau.to(dev), normalize=True)
    a1, b1, t1 = hA.to(dev).requires_grad_(), hB.to(dev).requires_grad_(), tau.to(dev).requires_grad_()
    sA, sB = ops.filip_similarity(ops.normalize_scale(a1, None, eps=0.0), ops.normalize_scale(b1, None, eps=0.0),
                                  mA.to(dev), mB.to(dev), t1)
    l1 = ops.filip_loss(sA, sB)
    l1.backward()
    out["filip_loss_rel_err"] = abs(float(lf) - float(l1.detach())) / abs(float(l1.detach()))
    out["filip_grad_max_rel_err"] = max(float((dhA - a1.grad[fs]).abs().max() / a1.grad.abs().max()),
                                        float((dhB - b1.grad[fs]).abs().max() / b1.grad.abs().max()))
    out["filip_dtau_abs_err"] = abs(float(dtau) - float(t1.grad))
    # ---- top-k, both shardings ----
    N, Q, k = 8192 * world, 512 * world, 10
    bank, q, _ = synth.retrieval_bank(N, Q, cfg_id=6, device="cpu")
    i1, s1, _ = ops.topk(q.to(dev), bank.to(dev), k)
    Ql, Nl = Q // world, N // world
    iq, sq = sharded.ShardedTopk(k, "queries").run(q[rank * Ql:(rank + 1) * Ql].to(dev), bank.to(dev))
    ib, sb = sharded.ShardedTopk(k, "bank").run(q.to(dev), bank[rank * Nl:(rank + 1) * Nl].contiguous().to(dev))
    out["topk_query_sharded_identical"] = bool(torch.equal(iq, i1[rank * Ql:(rank + 1) * Ql])
                                               and torch.equal(sq, s1[rank * Ql:(rank + 1) * Ql]))
    out["topk_bank_sharded_identical"] = bool(torch.equal(ib, i1) and torch.equal(sb, s1))
    # worst value over the ranks
    keys = sorted(out)
    v = torch.tensor([float(out[k_]) if not isinstance(out[k_], bool) else (0.0 if out[k_] else 1.0) for k_ in keys],
                     dtype=torch.float64, device=dev)
    dist.dist.all_reduce(v, op=dist.dist.ReduceOp.MAX)
    res = {}
    for k_, val in zip(keys, v.tolist()):
        res[k_] = (val == 0.0) if isinstance(out[k_], bool) else float(f"{val:.3e}")
    res["ok"] = bool(res["clip_loss_rel_err"] <= 1e-5 and res["clip_grad_max_rel_err"] <= 1e-4
                     and res["filip_loss_rel_err"] <= 1e-5 and res["filip_grad_max_rel_err"] <= 1e-4
                     and res["topk_query_sharded_identical"] and res["topk_bank_sharded_identical"])
    res["ranks"] = world
    return res


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


def extra(name, make, run):
    """A secondary workload must not take the headline line down."""
    try:
        wl = make()
        out = run(wl)
        del wl
    except Exception as e:
        out = {"error": f"{type(e).__name__}: {e}"[:300]}
    torch.cuda.empty_cache()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="clip_b65536")
    ap.add_argument("--no-extras", action="store_true", help="skip the secondary workloads and the CPU baselines")
    ap.add_argument("--selftest", action="store_true", help="multi-rank parity leg only (under torch.distributed.run)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: protein-clip_b200 has no CPU path")
    dist = Dist(args.gpus)
    if args.selftest:
        if not dist.on:
            raise SystemExit("--selftest needs >= 2 ranks: python -m torch.distributed.run --nproc-per-node N bench.py --selftest")
        res = multirank_parity(dist)
        if dist.rank == 0:
            print(json.dumps({"selftest": "multirank_parity", **res}), flush=True)
        dist.close()
        raise SystemExit(0 if res["ok"] else 1)
    peaks = load_peaks()
    from pclip import ops
    w = ClipLarge(dist)
    # ---- device-resident leg: `value` ----
    for _ in range(args.warmup):
        w.step()
    ops.reset_launch_count()
    w.record = True                       # eager launches with one CUDA-event pair per stage and step
    with ClockSampler(dist.local) as clk:
        ms = timed(dist, w.step, args.steps, 0)
    w.record = False
    launches = dist.sum_int(ops.launch_count())
    roof = w.roofline(peaks, args.steps)
    graphed = w.enable_graph()            # N > 1: the same K steps again as CUDA-graph replays; `value` is that run
    if graphed:
        roof["kernel_ms_measured_in"] = "an eager pass of the same K steps (events cannot be recorded inside a graph replay)"
        with ClockSampler(dist.local) as clk:
            ms_eager, ms = ms, timed(dist, w.step, args.steps, args.warmup)
    value = w.units * args.steps / (ms * 1e-3)
    check = w.check()
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
                            "timed steps" + ("; one captured graph per staging buffer" if graphed else "")
                            if hasattr(w, "e2e_begin") else "serial"},
        "gpu_launches": launches, "check": check,
        "launch_mode": ("one CUDA graph replay per step (capture of pack, Y all-gather, X peer copies on a side stream, "
                        "forward sweep, record all-gather, finalize, gradient); "
                        f"eager launches: {ms_eager / args.steps:.4f} ms/step" if graphed else
                        "eager" + (f" (graph capture failed: {w.graph_error})" if w.graph_error else "")),
        "step_tflops_algorithmic": round(6.0 * w.B * w.B * w.D / (ms / args.steps * 1e-3) / 1e12, 1),
        "step_frac_of_tensor_peak": round(6.0 * w.B * w.B * w.D / (ms / args.steps * 1e-3) / 1e12 / peaks["tc"]
                                          / max(args.gpus, 1), 4),   # whole-job flops over N GPUs' peak
    }
    extras = {}
    if dist.world == 1 and not args.no_extras:
        del w.plan, w.job
        torch.cuda.empty_cache()
        extras["clip_b16"] = extra("clip_b16", lambda: ClipSmall(dist), lambda wl: wl.report(dist))
        extras["encoder_head_b16"] = extra("encoder_head_b16", lambda: EncoderHead(dist), lambda wl: wl.report(dist))
        extras["clip_gc256"] = extra("clip_gc256", lambda: ClipGradCache(dist), lambda wl: wl.report(dist))
        extras["filip_b256"] = extra("filip_b256", lambda: FilipB256(dist), lambda wl: wl.report(dist, peaks))
        extras["filip_b256_ragged"] = extra("filip_b256_ragged", lambda: FilipB256(dist, ragged=True),
                                            lambda wl: wl.report(dist, peaks))
        extras["topk_4m"] = extra("topk_4m", lambda: Topk4M(dist), lambda wl: wl.report(dist, peaks))
        out["workloads"] = extras
        out["cpu_baseline"] = w.cpu_sample()
        out["check"]["loss_oracle_cpu"] = w.oracle_loss()
        out["check"]["loss_rel_err_vs_oracle"] = float(f"{abs(check['loss'] - out['check']['loss_oracle_cpu']) / abs(out['check']['loss_oracle_cpu']):.3e}")
    elif dist.world > 1 and not args.no_extras:
        # N > 1: parity of the sharded paths against the single-GPU results, then configs[2] with the A rows sharded
        # (all-gather of the B tokens, reduce-scatter of their gradient) and configs[4] with the queries sharded over a
        # replicated bank (independent units, no collective).  Every rank runs them; times are the max over ranks.
        del w.plan, w.job
        torch.cuda.empty_cache()
        try:
            out["multirank_parity"] = multirank_parity(dist)
        except Exception as e:
            out["multirank_parity"] = {"ok": False, "error": f"{type(e).__name__}: {e}"[:300]}
        torch.cuda.empty_cache()
        extras["filip_b256_a_row_sharded"] = extra("filip", lambda: FilipB256(dist, sharded=True),
                                                   lambda wl: wl.report(dist, peaks, cpu=False))
        extras["topk_4m_query_sharded"] = extra("topk", lambda: Topk4M(dist, Q=(1 << 20) // dist.world, q_total=1 << 20),
                                                lambda wl: wl.report(dist, peaks, cpu=False))
        out["workloads"] = extras
        out["cpu_baseline"] = None
    else:
        out["cpu_baseline"] = None
    if dist.rank == 0:
        print(json.dumps(out), flush=True)
    # tear-down: drop the captured graph (it holds NCCL work) before the process group goes; a watchdog ends the process
    # if NCCL's destructor still blocks -- the result line is already out
    w.graph = None
    if getattr(w, "job", None) is not None:
        w.job._graph = None
    torch.cuda.synchronize()
    threading.Timer(20.0, lambda: os._exit(0)).start()
    dist.close()
    os._exit(0)


if __name__ == "__main__":
    main()
