"""The oracle checked against itself: two derivations, finite differences, invariants.

Parity of the floating-point half is UNPINNED (no TensorFlow 1.8 here, no numeric golden in
the reference) — see oracle/char_rnn_oracle.py.  What can be pinned is that the hand-derived
BPTT is the true gradient of the restated forward graph, by two independent routes.
"""
import json
import math
import os

import numpy as np
import pytest
import torch

from conftest import ROOT
from oracle import char_rnn_oracle as O
from oracle.torch_restatement import TorchStep

MODELS = ("lstm", "gru", "rnn")


def _setup(model, H=6, L=2, S=4, B=3, V=5, seed=0, dtype=np.float64):
    cfg = O.Config(model, H, L, S, B, V, 5.0)
    p = O.init_params(cfg, seed, dtype)
    rng = np.random.default_rng(seed + 1)
    for k in p:
        if k.endswith("bias"):
            p[k] = p[k] + rng.normal(size=p[k].shape) * 0.1
    return cfg, p, rng


@pytest.mark.parametrize("model", MODELS)
def test_numpy_vs_torch_autograd_fp64(model):
    cfg, p, rng = _setup(model)
    tr, ts = O.Trainer(cfg, p, np.float64), TorchStep(cfg, p, torch.float64)
    tr.begin_epoch(0)
    ts.begin_epoch(0)
    for it in range(4):                                   # steps 1.. use the carried state
        x, y = rng.integers(0, 5, (3, 4)), rng.integers(0, 5, (3, 4))
        l1, l2 = tr.step(x, y), ts.step(x, y)
        assert abs(l1 - l2) < 1e-12
        g1, g2 = tr.last["fb"]["grads"], ts.grads_numpy()
        for k in g1:
            assert np.abs(g1[k] - g2[k]).max() < 1e-12, k
        assert abs(tr.last["gnorm"] - ts.last_gnorm) < 1e-12
        p2 = ts.params_numpy()
        for k in tr.params:
            assert np.abs(tr.params[k] - p2[k]).max() < 1e-12, k


@pytest.mark.parametrize("model", MODELS)
def test_finite_differences(model):
    cfg, p, rng = _setup(model, H=4, L=2, S=3, B=2, V=5)
    x, y = rng.integers(0, 5, (2, 3)), rng.integers(0, 5, (2, 3))
    st = [tuple(rng.normal(size=s.shape) * 0.3 for s in layer) for layer in O.zero_state(cfg, np.float64)]
    fb = O.forward_backward(cfg, p, x, y, st, np.float64)
    eps = 1e-6
    for name in p:
        flat = p[name].reshape(-1)
        for idx in rng.choice(flat.size, size=min(6, flat.size), replace=False):
            old = flat[idx]
            flat[idx] = old + eps
            lp, _ = O.loss_from_logits(O.forward(cfg, p, x, st, np.float64)[0], y)
            flat[idx] = old - eps
            lm, _ = O.loss_from_logits(O.forward(cfg, p, x, st, np.float64)[0], y)
            flat[idx] = old
            num = (lp - lm) / (2 * eps)
            ana = fb["grads"][name].reshape(-1)[idx]
            assert abs(num - ana) < 1e-7 + 1e-5 * abs(ana), (name, idx, num, ana)


@pytest.mark.parametrize("model", MODELS)
def test_invariants(model):
    cfg = O.Config(model, 16, 2, 5, 4, 65, 5.0)
    p = O.init_params(cfg, 0)
    rng = np.random.default_rng(0)
    x, y = rng.integers(0, 65, (4, 5)), rng.integers(0, 65, (4, 5))
    fb = O.forward_backward(cfg, p, x, y, O.zero_state(cfg), np.float32)
    assert abs(float(fb["loss"]) - math.log(65)) < 0.2          # loss at init ~ ln V
    assert abs(fb["grads"]["rnnlm/softmax_b"].sum()) < 1e-6     # rows of dlogits sum to 0
    # un-deduplicated slices norm >= dense scatter-added norm is NOT guaranteed, but both are finite
    assert np.isfinite(fb["emb_slices_sqnorm"]) and fb["emb_slices_sqnorm"] > 0
    # softmax_b is glorot-initialised, not zero; GRU gate bias is one
    assert np.abs(p["rnnlm/softmax_b"]).max() > 0
    if model == "gru":
        assert (p["rnnlm/multi_rnn_cell/cell_0/gru_cell/gates/bias"] == 1.0).all()


def test_tvars_order_and_param_count():
    cfg = O.Config("lstm", 128, 2, 50, 50, 65)
    names = [n for n, _ in O.tvar_specs(cfg)]
    assert names[:3] == ["rnnlm/softmax_w", "rnnlm/softmax_b", "embedding"]
    assert names[3].endswith("cell_0/lstm_cell/kernel") and names[6].endswith("cell_1/lstm_cell/bias")
    assert sum(int(np.prod(s)) for _, s in O.tvar_specs(cfg)) == 279873
    assert sum(int(np.prod(s)) for _, s in O.tvar_specs(O.Config("gru", 512, 2, 100, 256, 96))) == 3247200
    assert sum(int(np.prod(s)) for _, s in O.tvar_specs(O.Config("rnn", 2048, 4, 256, 512, 256))) == 34611456


def test_clip_scale_tf_semantics():
    assert O.clip_scale(10.0, 5.0) == np.float32(0.5)
    assert O.clip_scale(2.0, 5.0) == np.float32(1.0)           # always multiplied, by exactly 1
    assert O.clip_scale(0.0, 5.0) == np.float32(1.0)


def test_embedding_norm_uses_slices_not_dense():
    """tf.global_norm sees IndexedSlices.values: repeated ids are NOT summed first."""
    cfg = O.Config("rnn", 4, 1, 3, 2, 3, 5.0)
    p = O.init_params(cfg, 1)
    x = np.zeros((2, 3), dtype=np.int64)                       # every token is id 0
    y = np.ones((2, 3), dtype=np.int64)
    fb = O.forward_backward(cfg, p, x, y, O.zero_state(cfg), np.float64)
    dense = (fb["grads"]["embedding"] ** 2).sum()
    assert abs(fb["emb_slices_sqnorm"] - (fb["dX"] ** 2).sum()) < 1e-15
    assert abs(dense - fb["emb_slices_sqnorm"]) > 1e-9         # the two norms differ


def test_golden_regression():
    """Loss curve of the fp32 oracle on a fixed seed (self-generated; guards refactors)."""
    path = os.path.join(ROOT, "tests", "golden", "oracle_losses.json")
    gold = json.load(open(path))
    for model, ref in gold.items():
        cfg = O.Config(model, 16, 2, 8, 4, 11, 5.0)
        tr = O.Trainer(cfg, O.init_params(cfg, 5), np.float32)
        tr.begin_epoch(0)
        rng = np.random.default_rng(9)
        got = []
        for _ in range(5):
            x, y = rng.integers(0, 11, (4, 8)), rng.integers(0, 11, (4, 8))
            got.append(float(tr.step(x, y)))
        assert np.allclose(got, ref, rtol=2e-5, atol=0), (model, got, ref)
