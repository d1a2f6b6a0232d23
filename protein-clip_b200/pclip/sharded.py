"""Row-sharded CLIP InfoNCE over the GPUs of one box (BASELINE.json configs[3]).

Rank p owns rows [p*B/P, (p+1)*B/P) of BOTH modalities.  One step:

  1. pack local rows to the compute dtype                                   (kernel, local)
  2. all-gather X and Y over NCCL / NVLink, both in ONE coalesced launch      (2 x B*D*2 bytes in bf16)
  3. bf16 (single pass): one sweep over X_p Y^T gives the row sums R_i and the diagonal of the local rows
     (complete) and the column sums C_j of ALL columns over the local rows (partial).  The three land in one
     contiguous per-rank record [C partial (B) | R (B/P) | S_ii (B/P)] and ONE all-gather of that record moves
     everything the ranks owe each other; each rank then reduces the column partials and finalises the loss and
     1/R, 1/C of all rows itself (64 K-element kernels) -- two NCCL launches per step instead of six: at 8 GPUs the
     step is 1.1 ms and each small collective costs 20-40 us of launch + latency.
     fp32 / first generation (two-sided): R_i from (X_p, Y) and C_i from (Y_p, X), both complete locally, then
     loss all-reduce and all-gathers of 1/R and 1/C.
  4. dX_local = sum_j P_ij y_j  and  dY_local = sum_j P_ji x_j               (kernel, local, complete)

Because every rank evaluates both the (X_p, Y) and the (Y_p, X) side, each gradient row block is complete locally:
there is NO gradient exchange (the price is that the score tiles of the two sides are both computed, which the
single-GPU path does as well).  The collective calls go through torch.distributed (plumbing); the arithmetic is
all in libpclip.so via the `stages` object (pclip.ops.ClipPlan on CUDA).  The stages object is injectable only so
that tests/test_sharded_gloo.py can drive this host logic on CPU ranks over gloo with the oracle as the stage
backend; the product constructs ClipPlan and nothing else.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


class _PeerRows:
    """This rank's packed X shard in symmetric memory (torch.distributed._symmetric_memory: every rank maps every peer's
    buffer into its own address space), and the copy-engine all-gather built on it: P plain device-to-device copies on a
    side stream that read the peers' shards over NVLink while the SMs run the forward sweep.  An NCCL all-gather in that
    position was measured and lost (its kernel takes SMs away from a whole-wave grid); copy engines take none."""

    def __init__(self, like: torch.Tensor, group):
        import torch.distributed._symmetric_memory as symm
        grp = group if group is not None else dist.group.WORLD
        self.P, self.rank = dist.get_world_size(grp), dist.get_rank(grp)
        self.local = symm.empty(tuple(like.shape), dtype=like.dtype, device=like.device)
        self.hdl = symm.rendezvous(self.local, grp)
        self.peers = [self.local if p == self.rank else self.hdl.get_buffer(p, tuple(like.shape), like.dtype)
                      for p in range(self.P)]
        self.stream = torch.cuda.Stream(like.device)

    @staticmethod
    def create(like, group):
        """None unless symmetric memory comes up on EVERY rank (the decision is all-reduced: a rank pulling from a peer
        that kept its shard in private memory would read garbage)."""
        try:
            pr = _PeerRows(like, group)
            ok = 1
        except Exception:
            pr, ok = None, 0
        flag = torch.tensor([ok], dtype=torch.int32, device=like.device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        return pr if int(flag.item()) == 1 else None

    def start(self, full: torch.Tensor):
        """Enqueue the gather of all shards into `full` [P * Ma, D] on the side stream, ordered after everything the
        current stream has enqueued so far.  The caller guarantees that every peer's shard is complete by then (here:
        the NCCL all-gather of Y that precedes it was entered by every peer after its pack)."""
        main = torch.cuda.current_stream(full.device)
        self.stream.wait_stream(main)
        Ma = self.local.shape[0]
        with torch.cuda.stream(self.stream):
            for k in range(1, self.P + 1):                      # ring order: no two ranks start on the same peer
                p = (self.rank + k) % self.P
                full[p * Ma:(p + 1) * Ma].copy_(self.peers[p], non_blocking=True)

    def join(self, dev):
        torch.cuda.current_stream(dev).wait_stream(self.stream)


class ShardedClip:
    def __init__(self, B: int, D: int, compute: int, device, group=None, stages=None):
        self.group = group
        self.P = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        if B % self.P:
            raise ValueError(f"B = {B} must divide evenly over {self.P} ranks")
        self.B, self.D, self.Ma = B, D, B // self.P
        if stages is None:
            from .ops import ClipPlan
            stages = ClipPlan(self.Ma, B, D, compute, device, diag_offset=self.rank * self.Ma)
        self.s = stages
        # fused exchange (step 3): needs the stage object's contiguous per-rank record
        self.fused = self.P > 1 and getattr(stages, "single_pass", False) and hasattr(stages, "stats_l")
        if self.fused:
            self.stats_all = torch.empty(self.P, stages.stats_l.numel(), dtype=stages.stats_l.dtype,
                                         device=stages.stats_l.device)
        self.coalesce = self.P > 1 and dist.get_backend(group) == "nccl"
        # X is not needed before the gradient: pull it over the copy engines under the forward sweep (PCLIP_PEER_PULL=0
        # keeps both gathers on NCCL in front of the sweep)
        self.pull = None
        if self.fused and self.coalesce and os.environ.get("PCLIP_PEER_PULL", "1") != "0":
            self.pull = _PeerRows.create(stages.xl, group)
            if self.pull is not None:
                stages.xl = self.pull.local

    def step(self, xs, ys, gscale=None, events=None):
        """xs / ys: this rank's row chunks.  Returns (global loss [1], dX_local [Ma, D], dY_local [Ma, D])."""
        s, g = self.s, self.group

        def rec(name, i):
            if events is not None and name in events:
                events[name][i].record(torch.cuda.current_stream(s.dev))

        s.pack(xs, ys)
        if self.pull is not None:
            # Y by NCCL: the sweep needs it, and its completion here means every peer has packed (stream order on the
            # peer: pack X, pack Y, all-gather).  X then travels on the copy engines under the sweep.  Re-use of the shard
            # buffer by the next step is safe: a peer cannot leave this step's record all-gather before this rank has
            # entered it, which is after the join below.
            dist.all_gather_into_tensor(s.yf, s.yl, group=g)
            self.pull.start(s.xf)
        elif self.P > 1:
            # (Tried: gather Y first and let X travel under the forward sweep AS A SECOND NCCL GATHER.  The NCCL kernel
            # then holds SMs the whole-wave grid of the sweep counts on: sweep 222 -> 262 us, two gathers 2 x 52 us
            # instead of 75 us.)
            if self.coalesce:
                with dist._coalescing_manager(group=g):      # one ncclGroup: a single launch for both gathers
                    dist.all_gather_into_tensor(s.xf, s.xl, group=g)
                    dist.all_gather_into_tensor(s.yf, s.yl, group=g)
            else:
                dist.all_gather_into_tensor(s.xf, s.xl, group=g)
                dist.all_gather_into_tensor(s.yf, s.yl, group=g)
        s.transpose_full()
        if self.fused:
            rec("rowsumexp", 0); s.sumexp(); rec("rowsumexp", 1)
            if self.pull is not None:
                self.pull.join(s.dev)
            dist.all_gather_into_tensor(self.stats_all.view(-1), s.stats_l, group=g)
            s.finalize_full(self.stats_all, self.B)
        else:
            if getattr(s, "single_pass", False):
                # one sweep over X_p Y^T: row sums complete, column sums partial over this rank's rows
                rec("rowsumexp", 0); s.sumexp(); rec("rowsumexp", 1)
                if self.P > 1:
                    dist.all_reduce(s.colsum_f, group=g)
                s.take_local_colsum()
            else:
                rec("rowsumexp", 0); s.rowsumexp(); rec("rowsumexp", 1)
            s.finalize(self.B)
            if self.P > 1:
                dist.all_reduce(s.loss, group=g)
                dist.all_gather_into_tensor(s.rinv_f, s.rinv, group=g)
                dist.all_gather_into_tensor(s.cinv_f, s.cinv, group=g)
        rec("grad", 0); s.grad(self.B, gscale); rec("grad", 1)
        return s.loss, s.dx, s.dy

    # -----------------------------------------------------------------------------------------------------
    # CUDA-graph replay.  At 8 GPUs one step is ~0.8 ms of which the two tensor-core kernels are 0.64: the rest is two
    # NCCL launches and ~20 small kernels whose HOST launch cost (about 5 us each through Python + ctypes) leaves the GPU
    # idle between them.  All buffers of a step are owned by the plan, so the whole step -- pack, both collectives,
    # forward sweep, finalize, gradient -- is captured once and replayed with one launch.
    # -----------------------------------------------------------------------------------------------------
    def capture_graph(self, xs, ys, gscale=None):
        """Capture step(xs, ys) into a CUDA graph and return (graph, outputs).  Several graphs over different input
        tensors (e.g. the two staging buffers of a prefetching loader) may be captured; they share the plan's buffers,
        so their replays must be stream-ordered."""
        dev = self.s.dev
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):            # warm-up outside the capture: NCCL channel set-up, lazy workspaces
            for _ in range(3):
                self.step(xs, ys, gscale)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            out = self.step(xs, ys, gscale)
        return graph, out

    def capture(self, xs, ys, gscale=None):
        """Capture step(xs, ys) into a CUDA graph.  xs / ys (and gscale) must be the tensors every later replay reads
        (fill them in place before replay()).  Returns self."""
        self._graph, self._out = self.capture_graph(xs, ys, gscale)
        return self

    def replay(self):
        """One launch for the whole step captured by capture(); returns (loss [1], dX_local, dY_local)."""
        self._graph.replay()
        return self._out


class ShardedFilip:
    """FILIP token-wise loss with the A-batch rows sharded over the ranks (configs[2]).

    Both FILIP losses are row-wise (ref: modules/training_utils_filip.py:26-39 -- simB is NOT transposed), so row a of
    the loss needs hA[a] and ALL of hB: rank p keeps its B/P samples of A, all-gathers hB (+ masks), and evaluates the
    [B/P, B] blocks of simA / simB, both log-sum-exps and dhA completely locally.  What every rank holds only a part of
    is dhB (each rank's rows contribute to every B sample): one reduce-scatter (sum) returns each rank its own B/P
    samples of it; the loss and d tau are all-reduced scalars.  Token length TB must be padded to the same value on all
    ranks.  `backend` (default pclip.ops) supplies `filip_similarity` and `filip_loss`; tests inject the oracle."""

    def __init__(self, group=None, backend=None, compute=None):
        self.group = group
        self.P = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        if backend is None:
            from . import ops as backend
        self.be = backend
        self.kw = {} if compute is None else {"compute": compute}     # pclip.ops.F32 / BF16 (default: PCLIP_COMPUTE)
        self.rs_native = dist.get_backend(group) == "nccl"       # gloo has no reduce-scatter: all-reduce + slice

    def step(self, hA_l, hB_l, maskA_l, maskB_l, tau, normalize: bool = False):
        """Local samples hA_l [Bl, TA, D], hB_l [Bl, TB, D], right-padded masks [Bl, T*]; tau 0-d.
        normalize=True applies ExtendedFILIP.forward's h / ||h|| (ref: modules/models.py:238-239) to the local shards
        first -- what a training step does -- so that only NORMALISED B tokens travel and the gradients returned are
        those of the raw tokens.  Returns (global loss, dhA_l, dhB_l, dtau): gradients of the GLOBAL mean loss."""
        g, P, r = self.group, self.P, self.rank
        Bl = hA_l.shape[0]
        B = Bl * P
        hA = hA_l.detach().clone().requires_grad_(True)
        hB_raw = hB_l.detach().clone().requires_grad_(normalize)
        hA_in = self.be.normalize_scale(hA, None, eps=0.0) if normalize else hA
        hB_n = self.be.normalize_scale(hB_raw, None, eps=0.0) if normalize else hB_raw
        hB_src, mB_src = hB_n.detach().contiguous(), maskB_l.contiguous()
        if P > 1:
            hB_f = torch.empty((B,) + tuple(hB_src.shape[1:]), dtype=hB_src.dtype, device=hB_src.device)
            mB_f = torch.empty((B,) + tuple(mB_src.shape[1:]), dtype=mB_src.dtype, device=mB_src.device)
            dist.all_gather_into_tensor(hB_f, hB_src, group=g)
            dist.all_gather_into_tensor(mB_f, mB_src, group=g)
        else:
            hB_f, mB_f = hB_src.clone(), mB_src
        hB_f.requires_grad_(True)
        tau_l = torch.as_tensor(tau).detach().clone().to(hA.device).requires_grad_(True)
        simA, simB = self.be.filip_similarity(hA_in, hB_f, maskA_l, mB_f, tau_l, **self.kw)
        # filip_loss is the mean over the LOCAL rows; Bl / B of it is this rank's share of the global mean
        loss = self.be.filip_loss(simA, simB, diag_offset=r * Bl) * (Bl / B)
        loss.backward()
        loss_g, dtau = loss.detach().clone().reshape(1), tau_l.grad.detach().clone().reshape(1)
        dhB_f = hB_f.grad
        if P > 1:
            dist.all_reduce(loss_g, group=g)
            dist.all_reduce(dtau, group=g)
            if self.rs_native:
                dhB_l = torch.empty_like(hB_src)
                dist.reduce_scatter_tensor(dhB_l, dhB_f.contiguous(), group=g)
            else:
                dist.all_reduce(dhB_f, group=g)
                dhB_l = dhB_f[r * Bl:(r + 1) * Bl].clone()
        else:
            dhB_l = dhB_f
        if normalize:      # back through the local normalise of the B shard
            hB_n.backward(dhB_l)
            dhB_l = hB_raw.grad
        return loss_g, hA.grad, dhB_l, dtau


class ShardedTopk:
    """Cosine / dot-product top-k over the GPUs of one box (configs[4]).

    mode "queries" (default): the bank is replicated (1.07 GB in bf16 at N = 2^22 fits every GPU), each rank answers
        its own slice of the queries -- independent units, NO collective; the caller keeps the per-rank results.
    mode "bank": for banks that do not fit one GPU.  Every rank holds all queries and N / P bank rows, finds its local
        top-k with GLOBAL indices (`index_base`), and one all-gather of the [Q, k] (index, score) lists is followed by
        the k-way merge (`pclip_topk_merge`: score descending, ties to the lower global index) on every rank."""

    def __init__(self, k: int, mode: str = "queries", group=None, backend=None):
        if mode not in ("queries", "bank"):
            raise ValueError("mode must be 'queries' or 'bank'")
        self.k, self.mode, self.group = k, mode, group
        self.P = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        if backend is None:
            from . import ops as backend
        self.be = backend

    def run(self, q: torch.Tensor, bank: torch.Tensor, rows_before: int | None = None):
        """mode "queries": q = this rank's queries, bank = the full bank -> (idx, score) of those queries.
        mode "bank": q = ALL queries, bank = this rank's rows, `rows_before` = number of bank rows on lower ranks
        (default rank * len(bank): equal shards) -> the global (idx, score) of all queries, identical on every rank."""
        if self.mode == "queries":
            idx, score = self.be.topk(q, bank, self.k)[:2]
            return idx, score
        base = self.rank * bank.shape[0] if rows_before is None else rows_before
        idx, score = self.be.topk(q, bank, self.k, index_base=base)[:2]
        if self.P == 1:
            return idx, score
        idx_all = torch.empty((self.P,) + tuple(idx.shape), dtype=idx.dtype, device=idx.device)
        score_all = torch.empty((self.P,) + tuple(score.shape), dtype=score.dtype, device=score.device)
        # (concatenation along dim 0 of the flattened view: the layout every backend accepts)
        dist.all_gather_into_tensor(idx_all.view(-1, idx.shape[-1]), idx.contiguous(), group=self.group)
        dist.all_gather_into_tensor(score_all.view(-1, score.shape[-1]), score.contiguous(), group=self.group)
        return self.be.topk_merge(idx_all, score_all)
