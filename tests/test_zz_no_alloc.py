"""tadnn_head_forward_noalloc — the classifier-head forward for callers that cannot allocate.

A program that drives libtadnn.so through its C ABI may be unable to provide the scratch buffer
that tadnn_head_forward asks for: it may sit in a loop where a cudaMalloc (a device-wide
synchronisation of unbounded latency) is off limits, its device memory may already be fully owned
by another pool, or it may be a small C client with no allocator at all.  The no-alloc entry
point keeps its 2 KB of scratch inside the library instead.

This test goes through the C ABI with ctypes, exactly as such a client would, and checks the
new entry point against the existing one on two calls in a row (the second call sees the
scratch contents the first one left behind).  The file name makes pytest collect it last, so a
failure here cannot hide the result of any other test in a run that stops at the first failure.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _args(K, B, seed, dropout_p):
    g = torch.Generator(device="cuda").manual_seed(seed)
    t = {
        "p2": torch.randn(B, 64, device="cuda", generator=g).bfloat16(),
        "w1": torch.randn(100, 64, device="cuda", generator=g) * 0.2,
        "b1": torch.randn(100, device="cuda", generator=g) * 0.1,
        "w2": torch.randn(10, 100, device="cuda", generator=g) * 0.2,
        "b2": torch.randn(10, device="cuda", generator=g) * 0.1,
        "target": torch.randint(0, 10, (B,), device="cuda", generator=g),
    }
    outs = []
    for _ in range(2):
        o = {"hidden": torch.zeros(B, 100, device="cuda"), "logp": torch.zeros(B, 10, device="cuda"),
             "loss": torch.zeros((), device="cuda")}
        h = K.HeadArgs()
        h.p2, h.w1, h.b1, h.w2, h.b2 = (t[k].data_ptr() for k in ("p2", "w1", "b1", "w2", "b2"))
        h.target = t["target"].data_ptr()
        h.B, h.D, h.H, h.NC = B, 64, 100, 10
        h.dropout_p, h.seed = dropout_p, 1234 + seed
        h.hidden, h.logp, h.loss = o["hidden"].data_ptr(), o["logp"].data_ptr(), o["loss"].data_ptr()
        outs.append((h, o))
    return t, outs


def test_noalloc_matches_workspace_entry_point_twice_in_a_row():
    from tadnn import _native
    from tadnn.ops import cnn as K
    lib = _native.lib()
    lib.tadnn_head_forward_noalloc.argtypes = [C.POINTER(K.HeadArgs), C.c_void_p]
    stream = C.c_void_p(_native.stream_ptr())
    ws = K.head_workspace(torch.device("cuda", torch.cuda.current_device()))
    for call, (B, p) in enumerate([(4096, 0.5), (777, 0.0)]):  # second call: different size, stale scratch
        keep, ((h_ref, o_ref), (h_new, o_new)) = _args(K, B, seed=call, dropout_p=p)
        _native.check(lib.tadnn_head_forward(C.byref(h_ref), C.c_void_p(ws.data_ptr()), stream), "tadnn_head_forward")
        _native.check(lib.tadnn_head_forward_noalloc(C.byref(h_new), stream), "tadnn_head_forward_noalloc")
        torch.cuda.synchronize()
        assert torch.equal(o_new["hidden"], o_ref["hidden"])
        assert torch.equal(o_new["logp"], o_ref["logp"])
        ref, new = o_ref["loss"].item(), o_new["loss"].item()
        assert ref == ref and abs(new - ref) <= 1e-5 * max(1.0, abs(ref)), (call, new, ref)
        # and against plain PyTorch fp32 when dropout is off
        if p == 0.0:
            x = keep["p2"].float()
            hid = torch.relu(x @ keep["w1"].t() + keep["b1"])
            logp = torch.log_softmax(hid @ keep["w2"].t() + keep["b2"], dim=-1)
            want = torch.nn.functional.nll_loss(logp, keep["target"]).item()
            assert abs(new - want) <= 2e-4 * max(1.0, abs(want))
        del keep


def test_noalloc_rejects_batches_beyond_its_limit():
    from tadnn import _native
    from tadnn.ops import cnn as K
    lib = _native.lib()
    lib.tadnn_head_forward_noalloc.argtypes = [C.POINTER(K.HeadArgs), C.c_void_p]
    h = K.HeadArgs()
    h.B, h.D, h.H, h.NC = 65537, 64, 100, 10
    rc = lib.tadnn_head_forward_noalloc(C.byref(h), C.c_void_p(0))
    assert rc != 0
    msg = lib.tadnn_last_error().decode()
    assert "65536" in msg and "exceeds" in msg
