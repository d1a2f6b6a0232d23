"""Prints kernel-vs-torch numbers for the CLIP kernels (debug aid for gpurun; not a test)."""
import sys, os, traceback
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "protein-clip_b200"))
import torch
from pclip import ops, synth

torch.set_printoptions(precision=5, linewidth=200)


def ref(x, y):
    x = x.double().requires_grad_(); y = y.double().requires_grad_()
    s = x @ y.t(); e = s.exp()
    loss = 0.5 * ((e.sum(1).log() - s.diagonal()).mean() + (e.sum(0).log() - s.diagonal()).mean())
    loss.backward()
    return loss.detach(), x.grad, y.grad, e.sum(1).log().detach(), e.sum(0).log().detach()


for mode, name in ((ops.F32, "f32"), (ops.BF16, "bf16")):
    for B in (16, 128, 256, 1000):
        try:
            x, y = synth.clip_scaled_embeddings(B, 128, cfg_id=B)
            xr, yr = (x, y) if mode == ops.F32 else (x.bfloat16().float(), y.bfloat16().float())
            lo, dxo, dyo, rlo, clo = ref(xr, yr)
            loss, dx, dy, rl, cl = ops.clip_loss_raw([x.cuda()], [y.cuda()], mode, True, want_lse=True)
            torch.cuda.synchronize()
            print(f"[{name} B={B}] loss {float(loss):.7f} ref {float(lo):.7f} | row_lse err {float((rl.cpu()-rlo).abs().max()):.2e}"
                  f" col_lse err {float((cl.cpu()-clo).abs().max()):.2e} | dx rel {float((dx.cpu()-dxo).abs().max()/dxo.abs().max()):.2e}"
                  f" dy rel {float((dy.cpu()-dyo).abs().max()/dyo.abs().max()):.2e}")
            if float((rl.cpu() - rlo).abs().max()) > 1e-3:
                print("  row_lse kernel", rl[:8].cpu().tolist()); print("  row_lse ref   ", rlo[:8].tolist())
            if float((dx.cpu() - dxo).abs().max() / dxo.abs().max()) > 2e-2:
                print("  dx kernel", dx[0, :8].cpu().tolist()); print("  dx ref   ", dxo[0, :8].tolist())
                print("  dx kernel r127", dx[min(B,128) - 1, :8].cpu().tolist()); print("  dx ref r127   ", dxo[min(B,128) - 1, :8].tolist())
        except Exception:
            traceback.print_exc()
            sys.exit(1)
print("debug done")
