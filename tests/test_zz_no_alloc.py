"""pclip_rowwise_infonce_noalloc (static __device__ scratch, no allocation) against pclip_rowwise_infonce (caller-owned
workspace), both through the C ABI.  The entry point has no counterpart in the reference;
the check is agreement with the existing entry point, within the 1e-5 the row-wise loss is held to
elsewhere (tests/test_gpu_filip.py) -- in fact bit-for-bit, since kernel and summation order are the same.  Collected
last on purpose (file name), and self-contained: passes alone in a fresh process and after the other tests."""
import pytest
import torch

pytestmark = pytest.mark.gpu
TOL = 1e-5


def _call(L, fn, lg, diag_offset, scale, loss, dl, *extra):
    with torch.cuda.device(lg.device):
        L.check(fn(L.ptr(lg), lg.shape[0], lg.shape[1], diag_offset, scale, L.ptr(loss), L.ptr(dl), *extra,
                   L.stream_ptr(lg.device)))


@pytest.mark.parametrize("B,Bc,off", [(256, 256, 0), (37, 101, 50), (4099, 4200, 64)])
def test_noalloc_equals_workspace_entry_two_calls_in_a_row(B, Bc, off):
    from pclip import _lib as L
    lib = L.lib()
    dev = torch.device("cuda", 0)
    g = torch.Generator().manual_seed(B)
    results = []
    for call in range(2):        # second call: different logits, the static scratch still holds the first call's partials
        lg = (3.0 * torch.randn(B, Bc, generator=g)).to(dev)
        scale = 0.5 / B
        with torch.cuda.device(dev):
            n = int(lib.pclip_rowsum_workspace_bytes(B))
        owner, ws = L.workspace(n, dev)
        loss_ref, dl_ref = torch.zeros(1, device=dev), torch.empty(B, Bc, device=dev)
        _call(L, lib.pclip_rowwise_infonce, lg, off, scale, loss_ref, dl_ref, ws, n)
        loss_new, dl_new = torch.zeros(1, device=dev), torch.full((B, Bc), float("nan"), device=dev)
        _call(L, lib.pclip_rowwise_infonce_noalloc, lg, off, scale, loss_new, dl_new)
        torch.cuda.synchronize(dev)
        assert abs(float(loss_new) - float(loss_ref)) <= TOL * abs(float(loss_ref)), (call, float(loss_new), float(loss_ref))
        assert float((dl_new - dl_ref).abs().max()) <= TOL * float(dl_ref.abs().max()), call
        assert torch.equal(loss_new, loss_ref) and torch.equal(dl_new, dl_ref)      # same kernel, same order
        # loss only (dlogits NULL), accumulated onto a non-zero value like the existing entry point does
        acc_ref, acc_new = torch.full((1,), 2.5, device=dev), torch.full((1,), 2.5, device=dev)
        _call(L, lib.pclip_rowwise_infonce, lg, off, scale, acc_ref, None, ws, n)
        _call(L, lib.pclip_rowwise_infonce_noalloc, lg, off, scale, acc_new, None)
        assert torch.equal(acc_new, acc_ref)
        # independent value: torch's own log-softmax
        rows = torch.arange(B, device=dev)
        want = float((torch.logsumexp(lg.double(), 1) - lg.double()[rows, rows + off]).sum() * scale)
        assert abs(float(loss_new) - want) <= TOL * abs(want)
        results.append(float(loss_new))
    assert results[0] != results[1]


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    from pclip import _lib as L
    lib = L.lib()
    dev = torch.device("cuda", 0)
    cap = int(lib.pclip_rowwise_infonce_noalloc_max_rows())
    assert cap == 32768
    B = cap + 1
    lg = torch.zeros(1, device=dev)                                                  # never dereferenced: rejected first
    loss = torch.zeros(1, device=dev)
    with torch.cuda.device(dev):
        rc = lib.pclip_rowwise_infonce_noalloc(L.ptr(lg), B, B, 0, 1.0, L.ptr(loss), None, L.stream_ptr(dev))
    assert rc == -2                                                                  # PCLIP_ERR_UNSUPPORTED
    msg = lib.pclip_last_error().decode()
    assert "32768" in msg and "exceeds" in msg
    torch.cuda.synchronize(dev)
    assert float(loss) == 0.0
