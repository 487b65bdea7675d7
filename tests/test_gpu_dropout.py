"""Dropout (``--input_keep_prob`` / ``--output_keep_prob``; reference ``model.py:31-34,58-59``) on the CUDA path.

TF draws its masks from its own Philox streams, which no other implementation reproduces; what IS
checkable is that, GIVEN the masks a training call used, loss / gradients / state equal the oracle's
with the same masks.  ``crnn_dropout_mask`` returns exactly those multipliers (they are a pure function
of seed, training-call index, site and element), so each call can be replayed on the CPU.
"""
import argparse

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from oracle import char_rnn_oracle as O  # noqa: E402


def _relerr(a, b, floor):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return np.abs(a - b).max() / max(np.abs(b).max(), floor)


def _model(model, H, L, S, B, V, ikp, okp, precision, seed=3):
    from distributed_char_rnn_b200.model import Model
    cfg = O.Config(model, H, L, S, B, V, 5.0)
    params = O.init_params(cfg, seed)
    args = argparse.Namespace(model=model, rnn_size=H, num_layers=L, seq_length=S, batch_size=B, vocab_size=V,
                              grad_clip=5.0, output_keep_prob=okp, input_keep_prob=ikp)
    m = Model(args, device="cuda:0", precision=precision, seed=seed)
    m.set_variables(params)
    return cfg, params, m


@pytest.mark.parametrize("ikp,okp", [(1.0, 0.5), (0.8, 1.0), (0.7, 0.6)])
def test_masks_have_the_stated_distribution(ikp, okp):
    L, S, B, H = 2, 20, 30, 64
    cfg, params, m = _model("lstm", H, L, S, B, 17, ikp, okp, "fp32")
    masks = m.dropout_masks(0)
    assert len(masks) == L + 1
    # site 0 and the inner sites carry output-mask x input-mask, the top site the output mask only
    for site, mk in enumerate(masks):
        keep = okp if site == L else ikp * okp
        scale = 1.0 / okp if site == L else 1.0 / (ikp * okp)
        vals = np.unique(mk)
        assert all(v == 0.0 or abs(v - scale) <= 1e-5 * scale for v in vals), (site, vals)
        frac = (mk != 0).mean()
        assert abs(frac - keep) < 4 * np.sqrt(keep * (1 - keep) / mk.size) + 1e-9, (site, frac, keep)
        if keep < 1:
            assert abs(mk.mean() - 1.0) < 0.05                     # inverted dropout keeps the expectation
    # a different training call draws different masks; the same call index is reproducible
    again = m.dropout_masks(0)
    nxt = m.dropout_masks(1)
    assert all((a == b).all() for a, b in zip(masks, again))
    assert any((a != b).any() for a, b in zip(masks, nxt))
    # sites are independent draws
    assert (masks[0] != masks[1]).any()


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("model,H,L,S,B,V,ikp,okp", [
    ("lstm", 32, 2, 6, 5, 11, 0.8, 0.7), ("gru", 64, 2, 8, 20, 17, 1.0, 0.5), ("rnn", 48, 3, 7, 12, 13, 0.6, 1.0),
    ("lstm", 128, 2, 10, 50, 65, 0.9, 0.9), ("gru", 32, 1, 5, 9, 7, 0.5, 0.5),
])
def test_gradients_with_dropout_match_oracle_given_the_masks(model, H, L, S, B, V, ikp, okp, precision):
    cfg, params, m = _model(model, H, L, S, B, V, ikp, okp, precision)
    bf16 = precision == "bf16"
    rtol, floor = (1e-2, 1e-2) if bf16 else (1e-4, 1e-3)          # same bars as test_gpu_tensor / test_gpu_parity
    rng = np.random.default_rng(S * B)
    for call in range(2):                                          # two calls: the mask index advances
        x, y = rng.integers(0, V, (B, S)), rng.integers(0, V, (B, S))
        st = [tuple((rng.normal(size=s.shape) * 0.5).astype(np.float32) for s in layer) for layer in O.zero_state(cfg)]
        assert m.training_calls == call
        masks = m.dropout_masks(call)
        fb = O.forward_backward(cfg, params, x, y, st, np.float64, bf16_gemm=bf16, masks=masks)
        sflat = torch.from_numpy(np.concatenate([np.stack(layer).reshape(-1) for layer in st]).astype(np.float32)).cuda()
        loss, sout = m.gradients(x.astype(np.int32), y.astype(np.int32), sflat)
        assert abs(loss - fb["loss"]) <= (2e-3 if bf16 else 1e-4) * abs(fb["loss"])
        got = m.gradients_numpy()
        gmax = max(np.abs(g).max() for g in fb["grads"].values())
        for k, g in fb["grads"].items():
            assert _relerr(got[k], g, floor * gmax) < rtol, (call, k)
        ref_state = np.concatenate([np.stack(layer).reshape(-1) for layer in fb["final_state"]])
        assert _relerr(sout.cpu().numpy(), ref_state, floor) < rtol
        # the masks did something: the unmasked graph gives another loss
        plain = O.forward_backward(cfg, params, x, y, st, np.float64, bf16_gemm=bf16)
        assert abs(plain["loss"] - fb["loss"]) > 1e-4


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_keep_prob_one_is_the_plain_graph_and_sampling_ignores_dropout(precision):
    from distributed_char_rnn_b200.model import Model
    model, H, L, S, B, V = "lstm", 32, 2, 6, 8, 11
    cfg, params, m1 = _model(model, H, L, S, B, V, 1.0, 1.0, precision)
    rng = np.random.default_rng(0)
    x, y = rng.integers(0, V, (B, S)).astype(np.int32), rng.integers(0, V, (B, S)).astype(np.int32)
    l1, _ = m1.gradients(x, y, m1.zero_state())
    fb = O.forward_backward(cfg, params, x, y, O.zero_state(cfg), np.float64, bf16_gemm=precision == "bf16")
    assert abs(l1 - fb["loss"]) <= 2e-3 * abs(fb["loss"])
    # an inference model built from args that ask for dropout runs without it (model.py:31 `training and ...`)
    args = argparse.Namespace(model=model, rnn_size=H, num_layers=L, seq_length=S, batch_size=B, vocab_size=V,
                              grad_clip=5.0, output_keep_prob=0.5, input_keep_prob=0.5)
    ms = Model(args, training=False, device="cuda:0", precision=precision)
    ms.set_variables(params)
    _, p1, _ = ms.forward(np.array([[3]], np.int32), ms.zero_state())
    _, p2, _ = ms.forward(np.array([[3]], np.int32), ms.zero_state())
    assert torch.equal(p1, p2)
    c1 = O.Config(model, H, L, 1, 1, V, 5.0)
    ref = O.softmax(O.forward(c1, params, np.array([[3]]), O.zero_state(c1), np.float64,
                                        bf16_gemm=precision == "bf16")[0])
    assert _relerr(p1.cpu().numpy(), ref, 1e-3) < (1e-2 if precision == "bf16" else 1e-4)


def test_training_with_dropout_follows_the_oracle_trainer_step_by_step():
    """Five optimiser steps with state carry, each replayed on the oracle with that step's masks."""
    model, H, L, S, B, V = "gru", 32, 2, 6, 10, 13
    cfg, params, m = _model(model, H, L, S, B, V, 0.8, 0.7, "fp32")
    tr = O.Trainer(cfg, params, np.float32)
    tr.begin_epoch(0)
    m.assign_lr(0.002)
    rng = np.random.default_rng(4)
    for k in range(5):
        x, y = rng.integers(0, V, (B, S)), rng.integers(0, V, (B, S))
        masks = m.dropout_masks(m.training_calls)
        loss, _ = m.step(x.astype(np.int32), y.astype(np.int32))
        ref = float(tr.step(x, y, masks=masks))
        assert abs(loss - ref) <= 2e-4 * abs(ref), (k, loss, ref)
    got = m.variables_numpy()
    for name, v in tr.params.items():
        assert _relerr(got[name], v, 1e-3) < 5e-4, name
