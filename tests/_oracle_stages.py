"""Oracle-backed stand-in for pclip.ops.ClipPlan, used ONLY by the CPU / gloo tests of the row-sharding host logic
(pclip/sharded.py).  Same attribute names and stage methods, arithmetic from oracle/restate.py in float64."""
import torch


class OracleStages:
    def __init__(self, Ma, Nb, D, diag_offset, single_pass=False):
        self.single_pass = single_pass
        z = lambda *s: torch.zeros(*s, dtype=torch.float64)  # noqa: E731
        self.Ma, self.Nb, self.D, self.off, self.dev = Ma, Nb, D, diag_offset, torch.device("cpu")
        self.xl, self.yl, self.xf, self.yf = z(Ma, D), z(Ma, D), z(Nb, D), z(Nb, D)
        self.rinv, self.cinv, self.rinv_f, self.cinv_f = z(Ma), z(Ma), z(Nb), z(Nb)
        self.loss, self.dx, self.dy = z(1), z(Ma, D), z(Ma, D)
        if Ma == Nb:
            self.xf, self.yf, self.rinv_f, self.cinv_f = self.xl, self.yl, self.rinv, self.cinv
        elif single_pass:
            self.stats_l = z(Nb + 2 * Ma)        # [C partial | R | S_ii], as pclip.ops.ClipPlan lays it out

    def pack(self, xs, ys):
        self.xl.copy_(torch.cat(list(xs)).double())
        self.yl.copy_(torch.cat(list(ys)).double())

    def transpose_full(self):
        pass

    def rowsumexp(self):
        self.R = (self.xl @ self.yf.t()).exp().sum(1)      # ref: training_utils.py:85-87
        self.C = (self.yl @ self.xf.t()).exp().sum(1)      # :88 (column sums == row sums of the transposed side)
        self.diag = (self.xl * self.yf[self.off:self.off + self.Ma]).sum(-1)

    def sumexp(self):
        e = (self.xl @ self.yf.t()).exp()
        self.R = e.sum(1)
        self.colsum_f = e.sum(0)                             # partial over this rank's rows
        self.diag = (self.xl * self.yf[self.off:self.off + self.Ma]).sum(-1)
        if hasattr(self, "stats_l"):
            self.stats_l.copy_(torch.cat([self.colsum_f, self.R, self.diag]))

    def finalize_full(self, stats_all, B):
        Nb, Ma = self.Nb, self.Ma
        C = stats_all[:, :Nb].sum(0)
        R = stats_all[:, Nb:Nb + Ma].reshape(-1)
        d = stats_all[:, Nb + Ma:].reshape(-1)
        self.loss[0] = (R.log() + C.log() - 2 * d).sum() / (2 * B)      # global on every rank
        self.rinv_f.copy_(1 / R)
        self.cinv_f.copy_(1 / C)
        pos = slice(self.off, self.off + Ma)
        self.rinv.copy_(self.rinv_f[pos])
        self.cinv.copy_(self.cinv_f[pos])

    def take_local_colsum(self):
        self.C = self.colsum_f[self.off:self.off + self.Ma].clone()

    def finalize(self, B):
        self.loss[0] = (self.R.log() + self.C.log() - 2 * self.diag).sum() / (2 * B)
        self.rinv.copy_(1 / self.R)
        self.cinv.copy_(1 / self.C)

    def grad(self, B, gscale=None):
        g = 1.0 if gscale is None else float(gscale)
        k = g / (2 * B)
        pos = slice(self.off, self.off + self.Ma)
        e = (self.xl @ self.yf.t()).exp()
        self.dx.copy_(k * ((e * (self.rinv[:, None] + self.cinv_f[None, :])) @ self.yf - 2 * self.yf[pos]))
        et = (self.yl @ self.xf.t()).exp()
        self.dy.copy_(k * ((et * (self.cinv[:, None] + self.rinv_f[None, :])) @ self.xf - 2 * self.xf[pos]))
