"""crnn_clip_adam_noalloc: the optimiser step for callers that cannot give the library a workspace.

Some users of this library run where device memory is not theirs to take: every free byte already
belongs to another allocator, or a cudaMalloc on their path would stall the GPU or fail.  For them
the library carries a small fixed scratch array inside its own image and this entry point uses it
instead of a caller workspace.  The test drives it through the C-ABI exactly as such a caller would
(no workspace argument), twice in a row - the array is shared and keeps whatever the previous call
left in it - and compares parameters, both Adam slots and the reported global norm with what
crnn_clip_adam produces from the same inputs.  Tolerance: the 1e-4 relative bound used by the
parity tests (the two calls group the norm's partial sums differently, so bit equality is not
promised).  The file sorts last so that a run stopping at the first failure has already executed
every other test; it also passes on its own in a fresh process.
"""
import argparse

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

RTOL = 1e-4


def _model(model="lstm", H=48, L=2, S=6, B=5, V=13):
    from distributed_char_rnn_b200.model import Model
    ns = argparse.Namespace(model=model, rnn_size=H, num_layers=L, seq_length=S, batch_size=B, vocab_size=V,
                            grad_clip=0.05, output_keep_prob=1.0, input_keep_prob=1.0)
    return Model(ns, device="cuda:0", seed=4)


def _close(a, b, what):
    a, b = a.double(), b.double()
    scale = max(float(b.abs().max()), 1e-6)
    assert float((a - b).abs().max()) <= RTOL * scale, what


@pytest.mark.parametrize("model", ["lstm", "gru"])
def test_noalloc_matches_workspace_version_twice(lib, model):
    from distributed_char_rnn_b200._lib import check, ptr
    m = _model(model)
    rng = np.random.default_rng(0)
    ref = {k: getattr(m, k).clone() for k in ("params", "adam_m", "adam_v")}     # updated through the workspace version
    new = {k: getattr(m, k).clone() for k in ("params", "adam_m", "adam_v")}     # updated through the no-alloc version
    gn_ref = torch.zeros(1, device="cuda")
    gn_new = torch.zeros(1, device="cuda")
    for step in (1, 2):                                      # second call finds the first call's partial sums in the array
        x = rng.integers(0, 13, (5, 6)).astype(np.int32)
        y = rng.integers(0, 13, (5, 6)).astype(np.int32)
        m.params.copy_(ref["params"])
        m.gradients(x, y, m.zero_state())                    # fills m.grads (P + 2 floats)
        check(lib.crnn_clip_adam(m._h, ptr(ref["params"]), ptr(m.grads), ptr(ref["adam_m"]), ptr(ref["adam_v"]),
                                 0.002, 0.05, step, 1.0, ptr(gn_ref), ptr(m.workspace), None))
        check(lib.crnn_clip_adam_noalloc(m._h, ptr(new["params"]), ptr(m.grads), ptr(new["adam_m"]), ptr(new["adam_v"]),
                                         0.002, 0.05, step, 1.0, ptr(gn_new), None))
        torch.cuda.synchronize()
        assert float(gn_ref) > 0.05                          # clipping is active, so the norm matters
        _close(gn_new, gn_ref, "global norm, step %d" % step)
        for k in ref:
            _close(new[k], ref[k], "%s, step %d" % (k, step))
        assert not torch.equal(ref["params"], m.params)       # the update did something


def test_noalloc_does_not_allocate_device_memory(lib):
    """Free device memory as the driver reports it is the same before and after the calls."""
    from distributed_char_rnn_b200._lib import check, ptr
    m = _model("rnn")
    m.gradients(np.zeros((5, 6), np.int32), np.ones((5, 6), np.int32), m.zero_state())
    check(lib.crnn_clip_adam_noalloc(m._h, ptr(m.params), ptr(m.grads), ptr(m.adam_m), ptr(m.adam_v),
                                     0.002, 5.0, 1, 1.0, None, None))              # set-up call for this GPU
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for step in (2, 3, 4):
        check(lib.crnn_clip_adam_noalloc(m._h, ptr(m.params), ptr(m.grads), ptr(m.adam_m), ptr(m.adam_v),
                                         0.002, 5.0, step, 1.0, None, None))
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0


def test_noalloc_rejects_models_beyond_its_scratch(lib):
    """134,217,728 floats is the limit; a model above it is refused with a message naming the limit."""
    import ctypes
    from distributed_char_rnn_b200 import _lib
    cfg = _lib.CrnnConfig(_lib.MODEL_IDS["rnn"], 64, 1, 1, 1, 1100000, 0, 0)    # 2 x 64 x 1.1 M vocabulary floats > limit
    h = ctypes.c_void_p()
    _lib.check(lib.crnn_create(ctypes.byref(cfg), ctypes.byref(h)))
    try:
        assert lib.crnn_flat_size(h) > 8192 * 16384
        one = torch.zeros(4, device="cuda")                  # never dereferenced: the size check comes first
        rc = lib.crnn_clip_adam_noalloc(h, _lib.ptr(one), _lib.ptr(one), _lib.ptr(one), _lib.ptr(one),
                                        0.002, 5.0, 1, 1.0, None, None)
        assert rc != 0
        msg = lib.crnn_last_error().decode()
        assert "134217728" in msg and "crnn_clip_adam" in msg
    finally:
        lib.crnn_destroy(h)
