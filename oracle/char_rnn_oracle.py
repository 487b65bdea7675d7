"""CPU oracle for the char-RNN training step.  TEST INFRASTRUCTURE, NOT PRODUCT.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this file.  Nothing under
``distributed_char_rnn_b200/`` imports it.

What it restates
----------------
The static TF-1.8 graph the reference builds in ``model.py:8-98`` and runs once
per batch at ``train.py:199``: embedding lookup (``model.py:54-63``), the
``MultiRNNCell`` unrolled by ``legacy_seq2seq.rnn_decoder`` (``model.py:28-36,72``),
vocab projection + ``sequence_loss_by_example`` (``model.py:76-85``),
``tf.gradients`` + ``clip_by_global_norm`` (``model.py:88-92``) and
``AdamOptimizer.apply_gradients`` (``model.py:93-98``), plus the state carry of the
training loop (``train.py:185-199``).

PARITY UNPINNED for the floating-point half: the arithmetic lives in the
third-party wheel ``tensorflow==1.8`` (``requirements.txt:2``), which is absent from
``/root/reference`` and not installable here, and the reference holds no numeric
golden vector for this path (``.travis.yml`` only checks that files are
non-empty).  The restatement follows the published TF r1.8 semantics
(``rnn_cell_impl.py`` BasicRNNCell/GRUCell/LSTMCell, ``seq2seq.py``
sequence_loss_by_example, ``clip_ops.py`` clip_by_global_norm, ``training_ops.cc``
ApplyAdam, ``variable_scope.py`` default glorot_uniform initializer) as listed in
SURVEY.md §8(a5-a11).  It is cross-checked by (i) an independent PyTorch-autograd
restatement (``oracle/torch_restatement.py``) and (ii) central finite differences,
both in ``tests/test_oracle.py``.  The integer half (vocab, tensor, batches, shards)
IS pinned: the reference's own ``utils.py`` / ``data_splitter.py`` run here.

Hand-derived BPTT, NumPy, dtype selectable (float32 = the 1e-4 comparator,
float64 = the derivation check).
"""
import math

import numpy as np

GATES = {"lstm": 4, "gru": 3, "rnn": 1}


class Config(object):
    """The model-shaping subset of the reference's argparse Namespace (``train.py:37-55``)."""

    def __init__(self, model="lstm", rnn_size=128, num_layers=2, seq_length=50,
                 batch_size=50, vocab_size=65, grad_clip=5.0):
        if model not in GATES:
            # model.py:24-25
            raise Exception("model type not supported: {}".format(model))
        self.model = model
        self.rnn_size = rnn_size
        self.num_layers = num_layers
        self.seq_length = seq_length
        self.batch_size = batch_size
        self.vocab_size = vocab_size
        self.grad_clip = grad_clip


def tvar_specs(cfg):
    """``tf.trainable_variables()`` order and shapes (``model.py:48-54,72``)."""
    H, V = cfg.rnn_size, cfg.vocab_size
    specs = [("rnnlm/softmax_w", (H, V)), ("rnnlm/softmax_b", (V,)), ("embedding", (V, H))]
    for l in range(cfg.num_layers):
        p = "rnnlm/multi_rnn_cell/cell_%d/" % l
        if cfg.model == "lstm":
            specs += [(p + "lstm_cell/kernel", (2 * H, 4 * H)), (p + "lstm_cell/bias", (4 * H,))]
        elif cfg.model == "gru":
            specs += [(p + "gru_cell/gates/kernel", (2 * H, 2 * H)), (p + "gru_cell/gates/bias", (2 * H,)),
                      (p + "gru_cell/candidate/kernel", (2 * H, H)), (p + "gru_cell/candidate/bias", (H,))]
        else:
            specs += [(p + "basic_rnn_cell/kernel", (2 * H, H)), (p + "basic_rnn_cell/bias", (H,))]
    return specs


def init_params(cfg, seed=0, dtype=np.float32):
    """TF default initialisation, numpy ``PCG64(seed)``.

    ``tf.get_variable`` without an initializer → glorot_uniform ``U(-l, l)``,
    ``l = sqrt(6/(fan_in+fan_out))`` (1-D: fan_in = fan_out = n), so ``softmax_b`` is
    NOT zero.  Cell biases come from the cells' own initializers: 0, except the GRU
    gate bias which is 1.0.  The LSTM forget bias 1.0 is added at run time.
    """
    rng = np.random.Generator(np.random.PCG64(seed))
    params = {}
    for name, shape in tvar_specs(cfg):
        if name.endswith("/bias"):
            fill = 1.0 if name.endswith("gates/bias") else 0.0
            params[name] = np.full(shape, fill, dtype=dtype)
            continue
        fan_in, fan_out = (shape[0], shape[0]) if len(shape) == 1 else shape
        lim = math.sqrt(6.0 / (fan_in + fan_out))
        params[name] = rng.uniform(-lim, lim, size=shape).astype(np.float32).astype(dtype)
    return params


def zero_state(cfg, dtype=np.float32):
    """``cell.zero_state`` (``model.py:43``): per layer ``(c, h)`` for lstm, ``(h,)`` otherwise."""
    B, H = cfg.batch_size, cfg.rnn_size
    n = 2 if cfg.model == "lstm" else 1
    return [tuple(np.zeros((B, H), dtype) for _ in range(n)) for _ in range(cfg.num_layers)]


def _sigmoid(x):
    return 1.0 / (1.0 + np.exp(-x))


def _bf16_round(a):
    """Round-to-nearest-even to bfloat16, returned in the input dtype."""
    a32 = np.ascontiguousarray(a, dtype=np.float32)
    u = a32.view(np.uint32)
    r = ((u >> 16) & 1) + np.uint32(0x7FFF)
    out = ((u + r) & np.uint32(0xFFFF0000)).view(np.float32)
    return out.astype(a.dtype)


class _Mat(object):
    """matmul with optional bf16 rounding of both inputs (accumulation stays wide)."""

    def __init__(self, bf16):
        self.bf16 = bf16

    def __call__(self, a, b):
        if self.bf16:
            return _bf16_round(a) @ _bf16_round(b)
        return a @ b


def _cell_names(cfg, l):
    p = "rnnlm/multi_rnn_cell/cell_%d/" % l
    if cfg.model == "lstm":
        return (p + "lstm_cell/kernel", p + "lstm_cell/bias")
    if cfg.model == "gru":
        return (p + "gru_cell/gates/kernel", p + "gru_cell/gates/bias",
                p + "gru_cell/candidate/kernel", p + "gru_cell/candidate/bias")
    return (p + "basic_rnn_cell/kernel", p + "basic_rnn_cell/bias")


def forward(cfg, params, x, state, dtype=np.float32, bf16_gemm=False, keep=False, masks=None):
    """Embedding + unrolled cells + logits (``model.py:54-77``).

    ``x``: ``[B, S]`` ids.  Returns ``(logits [B*S, V] rows ordered b*S+t, final_state,
    tape)``; ``tape`` holds what BPTT needs when ``keep`` is set.
    """
    mm = _Mat(bf16_gemm)
    B, S = x.shape
    H, L = cfg.rnn_size, cfg.num_layers
    emb = params["embedding"].astype(dtype)
    inp = [emb[x[:, t]] for t in range(S)]                       # model.py:55,62-63
    state = [tuple(s.astype(dtype) for s in st) for st in state]
    tape = {"x": x, "layers": []}
    # masks (dropout, model.py:31-34,58-59): L arrays [S,B,H] multiplying what each cell reads (embedding or
    # previous cell's output mask x the cell's input mask) and one for the top output as the projection sees it
    for l in range(L):
        if masks is not None:
            inp = [inp[t] * masks[l][t].astype(dtype) for t in range(S)]
        names = _cell_names(cfg, l)
        P = [params[n].astype(dtype) for n in names]
        rec = {"in": inp, "hprev": [], "g": []}
        outs = []
        if cfg.model == "lstm":
            c, h = state[l]
            W, b = P
            for t in range(S):
                z = mm(np.concatenate([inp[t], h], 1), W) + b
                i, j, f, o = np.split(z, 4, axis=1)               # LSTMCell: i, j, f, o
                si, sf, so, tj = _sigmoid(i), _sigmoid(f + dtype(1.0)), _sigmoid(o), np.tanh(j)
                rec["hprev"].append(h)
                cn = sf * c + si * tj
                tc = np.tanh(cn)
                h = so * tc
                rec["g"].append((si, tj, sf, so, c, tc))
                c = cn
                outs.append(h)
            state[l] = (c, h)
        elif cfg.model == "gru":
            (h,) = state[l]
            Wg, bg, Wc, bc = P
            for t in range(S):
                v = _sigmoid(mm(np.concatenate([inp[t], h], 1), Wg) + bg)
                r, u = np.split(v, 2, axis=1)                     # GRUCell: r, u
                rs = r * h                                        # reset BEFORE the matmul
                cand = np.tanh(mm(np.concatenate([inp[t], rs], 1), Wc) + bc)
                rec["hprev"].append(h)
                rec["g"].append((r, u, cand, rs))
                h = u * h + (dtype(1.0) - u) * cand
                outs.append(h)
            state[l] = (h,)
        else:
            (h,) = state[l]
            W, b = P
            for t in range(S):
                rec["hprev"].append(h)
                h = np.tanh(mm(np.concatenate([inp[t], h], 1), W) + b)
                rec["g"].append((h,))
                outs.append(h)
            state[l] = (h,)
        tape["layers"].append(rec)
        inp = outs
    # model.py:73: concat(outputs, 1) -> [B, S*H] -> reshape [-1, H]: row = b*S + t
    if masks is not None:
        inp = [inp[t] * masks[L][t].astype(dtype) for t in range(S)]
    out = np.stack(inp, axis=1).reshape(B * S, H)
    logits = mm(out, params["rnnlm/softmax_w"].astype(dtype)) + params["rnnlm/softmax_b"].astype(dtype)
    tape["out"] = out
    return logits, state, (tape if keep else None)


def softmax(logits):
    m = logits.max(axis=1, keepdims=True)
    e = np.exp(logits - m)
    return e / e.sum(axis=1, keepdims=True)


def loss_from_logits(logits, y):
    """``sequence_loss_by_example`` with one step and unit weights, then ``model.py:85``.

    ``cost = sum(xent / (1 + 1e-12)) / B / S``; the 1e-12 vanishes in fp32.
    """
    n = logits.shape[0]
    m = logits.max(axis=1, keepdims=True)
    lse = np.log(np.exp(logits - m).sum(axis=1)) + m[:, 0]
    xent = lse - logits[np.arange(n), y.reshape(-1)]
    return xent.sum() / logits.dtype.type(n), xent


def forward_backward(cfg, params, x, y, state, dtype=np.float32, bf16_gemm=False, masks=None):
    """One ``tf.gradients(cost, tvars)`` evaluation (``model.py:80-91``).

    Returns a dict with ``loss``, ``grads`` (dense, keyed like ``params``; the
    embedding entry is the scatter-added dense gradient), ``emb_slices_sqnorm``
    (sum of squares of the UN-deduplicated ``IndexedSlices.values``, the term
    ``tf.global_norm`` actually sees — SURVEY §8 a10), ``final_state``, ``logits``.
    Gradients do not flow into the fed initial state (truncated BPTT).
    """
    mm = _Mat(bf16_gemm)
    B, S = x.shape
    H, L, V = cfg.rnn_size, cfg.num_layers, cfg.vocab_size
    one = dtype(1.0)
    logits, final_state, tape = forward(cfg, params, x, state, dtype, bf16_gemm, keep=True, masks=masks)
    loss, _ = loss_from_logits(logits, y)
    n = B * S
    dlogits = softmax(logits)
    dlogits[np.arange(n), y.reshape(-1)] -= one
    dlogits /= dtype(n)
    grads = {}
    Wsm = params["rnnlm/softmax_w"].astype(dtype)
    grads["rnnlm/softmax_w"] = mm(tape["out"].T, dlogits)
    grads["rnnlm/softmax_b"] = dlogits.sum(axis=0)
    dout = mm(dlogits, Wsm.T).reshape(B, S, H)
    dys = [dout[:, t, :] for t in range(S)]
    if masks is not None:
        dys = [dys[t] * masks[L][t].astype(dtype) for t in range(S)]
    for l in reversed(range(L)):
        rec = tape["layers"][l]
        names = _cell_names(cfg, l)
        P = [params[nm].astype(dtype) for nm in names]
        dxs = [None] * S
        dh = np.zeros((B, H), dtype)
        if cfg.model == "lstm":
            W, b = P
            dW, db = np.zeros_like(W), np.zeros_like(b)
            dc = np.zeros((B, H), dtype)
            for t in reversed(range(S)):
                si, tj, sf, so, cprev, tc = rec["g"][t]
                dht = dys[t] + dh
                do = dht * tc * so * (one - so)
                dct = dc + dht * so * (one - tc * tc)
                df = dct * cprev * sf * (one - sf)
                di = dct * tj * si * (one - si)
                dj = dct * si * (one - tj * tj)
                dc = dct * sf
                dz = np.concatenate([di, dj, df, do], 1)
                xh = np.concatenate([rec["in"][t], rec["hprev"][t]], 1)
                dW += mm(xh.T, dz)
                db += dz.sum(axis=0)
                dxh = mm(dz, W.T)
                dxs[t], dh = dxh[:, :H], dxh[:, H:]
            grads[names[0]], grads[names[1]] = dW, db
        elif cfg.model == "gru":
            Wg, bg, Wc, bc = P
            dWg, dbg, dWc, dbc = (np.zeros_like(a) for a in P)
            for t in reversed(range(S)):
                r, u, cand, rs = rec["g"][t]
                hp = rec["hprev"][t]
                dht = dys[t] + dh
                du = dht * (hp - cand)
                dcz = dht * (one - u) * (one - cand * cand)
                xrs = np.concatenate([rec["in"][t], rs], 1)
                dWc += mm(xrs.T, dcz)
                dbc += dcz.sum(axis=0)
                dxrs = mm(dcz, Wc.T)
                drs = dxrs[:, H:]
                dgz = np.concatenate([drs * hp * r * (one - r), du * u * (one - u)], 1)
                xh = np.concatenate([rec["in"][t], hp], 1)
                dWg += mm(xh.T, dgz)
                dbg += dgz.sum(axis=0)
                dxh = mm(dgz, Wg.T)
                dxs[t] = dxrs[:, :H] + dxh[:, :H]
                dh = dht * u + drs * r + dxh[:, H:]
            grads[names[0]], grads[names[1]], grads[names[2]], grads[names[3]] = dWg, dbg, dWc, dbc
        else:
            W, b = P
            dW, db = np.zeros_like(W), np.zeros_like(b)
            for t in reversed(range(S)):
                (hn,) = rec["g"][t]
                dz = (dys[t] + dh) * (one - hn * hn)
                xh = np.concatenate([rec["in"][t], rec["hprev"][t]], 1)
                dW += mm(xh.T, dz)
                db += dz.sum(axis=0)
                dxh = mm(dz, W.T)
                dxs[t], dh = dxh[:, :H], dxh[:, H:]
            grads[names[0]], grads[names[1]] = dW, db
        if masks is not None:
            dxs = [dxs[t] * masks[l][t].astype(dtype) for t in range(S)]
        dys = dxs
    dX = np.stack(dys, axis=1)                                   # [B, S, H] = IndexedSlices.values
    dE = np.zeros((V, H), dtype)
    np.add.at(dE, x.reshape(-1), dX.reshape(n, H))
    grads["embedding"] = dE
    return {"loss": loss, "grads": grads, "emb_slices_sqnorm": (dX.astype(np.float64) ** 2).sum(),
            "final_state": final_state, "logits": logits, "dX": dX}


def global_norm(cfg, fb):
    """``tf.global_norm``: dense grads by their own L2, the embedding by its slices."""
    ss = fb["emb_slices_sqnorm"]
    for name, g in fb["grads"].items():
        if name != "embedding":
            ss += (g.astype(np.float64) ** 2).sum()
    return math.sqrt(ss)


def clip_scale(gnorm, clip, dtype=np.float32):
    """``clip_by_global_norm``: ``scale = clip * min(1/norm, 1/clip)`` in the graph dtype."""
    c, g = dtype(clip), dtype(gnorm)
    with np.errstate(divide="ignore"):
        return c * min(dtype(1.0) / g, dtype(1.0) / c)


class AdamState(object):
    """TF ``AdamOptimizer`` slots: ``m``, ``v`` per variable and the two beta-power scalars."""

    def __init__(self, params, dtype=np.float32, beta1=0.9, beta2=0.999, eps=1e-8):
        self.m = {k: np.zeros_like(v, dtype=dtype) for k, v in params.items()}
        self.v = {k: np.zeros_like(v, dtype=dtype) for k, v in params.items()}
        self.b1, self.b2, self.eps = dtype(beta1), dtype(beta2), dtype(eps)
        self.b1p, self.b2p = dtype(beta1), dtype(beta2)
        self.dtype = dtype


def adam_apply(params, grads, st, lr):
    """``ApplyAdam`` (``training_ops.cc``): eps outside the bias correction.

    The sparse embedding path (``_apply_sparse_shared``) decays ``m``/``v`` of every
    row and adds the deduplicated slices, which equals the dense update on the
    scatter-added gradient used here.
    """
    d = st.dtype
    one = d(1.0)
    alpha = d(lr) * np.sqrt(one - st.b2p) / (one - st.b1p)
    for k in params:
        g = grads[k].astype(d)
        st.m[k] += (g - st.m[k]) * (one - st.b1)
        st.v[k] += (g * g - st.v[k]) * (one - st.b2)
        params[k] -= (alpha * st.m[k]) / (np.sqrt(st.v[k]) + st.eps)
    st.b1p *= st.b1
    st.b2p *= st.b2


class Trainer(object):
    """The reference's training loop body (``train.py:185-199``) as a stepping object."""

    def __init__(self, cfg, params, dtype=np.float32, learning_rate=0.002, decay_rate=0.97,
                 bf16_gemm=False):
        self.cfg, self.dtype, self.bf16 = cfg, dtype, bf16_gemm
        self.params = {k: v.astype(dtype).copy() for k, v in params.items()}
        self.adam = AdamState(self.params, dtype)
        self.learning_rate, self.decay_rate = learning_rate, decay_rate
        self.lr = dtype(0.0)                                     # model.py:87
        self.global_step = 0
        self.state = zero_state(cfg, dtype)

    def begin_epoch(self, e):
        """``train.py:187-190``: lr = learning_rate * decay_rate**e (fp32), state = zeros."""
        self.lr = np.float32(self.learning_rate) * np.float32(self.decay_rate) ** np.float32(e)
        self.state = zero_state(self.cfg, self.dtype)

    def step(self, x, y, world=None, masks=None):
        """One ``session.run([cost, final_state, train_op])``.  ``masks``: this step's dropout multipliers.

        ``world``: optional list of ``(x_r, y_r, state_r)`` for the N-rank contract
         — not used for single-process runs.
        """
        fb = forward_backward(self.cfg, self.params, x, y, self.state, self.dtype, self.bf16, masks=masks)
        gn = global_norm(self.cfg, fb)
        scale = clip_scale(gn, self.cfg.grad_clip, self.dtype)
        grads = {k: g * scale for k, g in fb["grads"].items()}
        adam_apply(self.params, grads, self.adam, self.lr)
        self.global_step += 1
        self.state = fb["final_state"]
        self.last = {"gnorm": gn, "scale": scale, "fb": fb}
        return fb["loss"]
