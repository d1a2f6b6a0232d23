import sys, os, time, ctypes, faulthandler
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "protein-clip_b200"))
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import numpy as np, torch
from pclip import ops, _lib
faulthandler.dump_traceback_later(60, exit=True)
def load(name):
    z = np.load(os.path.join(os.path.dirname(__file__), "..", "tests", "golden", name + ".npz"))
    return {k: torch.from_numpy(z[k]) for k in z.files}
dbg = torch.zeros(16, dtype=torch.int32).pin_memory()
h = _lib.lib(); h.pclip_debug_set.argtypes = [ctypes.c_void_p]; h.pclip_debug_set(dbg.data_ptr())
g = load("filip_b4_full")
mode = ops.BF16
hA = g["hA"].cuda().requires_grad_(); hB = g["hB"].cuda().requires_grad_(); tau = g["tau"].cuda().requires_grad_()
nA = ops.normalize_scale(hA, None, eps=0.0); nB = ops.normalize_scale(hB, None, eps=0.0)
simA, simB = ops.filip_similarity(nA, nB, g["maskA"].cuda(), g["maskB"].cuda(), tau, compute=mode)
loss = ops.filip_loss(simA, simB)
torch.cuda.synchronize(); print("loss ok", float(loss), flush=True)
loss.backward()
for k in range(6):
    time.sleep(1.0); print("dbg", dbg.tolist(), flush=True)
torch.cuda.synchronize()
print("bwd ok", float((hA.grad.cpu()-g["dhA"]).abs().max() / g["dhA"].abs().max()), flush=True)
