"""Second, independent CPU restatement of the reference graph: PyTorch-CPU autograd.

TEST INFRASTRUCTURE, NOT PRODUCT (same import rule as ``char_rnn_oracle.py``).

Two jobs:
  1. cross-check the hand-derived BPTT of ``char_rnn_oracle.py`` (no TensorFlow can be
     run here, so two derivations of the same graph have to agree — fp64, ~1e-10);
  2. be the TIMED CPU baseline of ``bench.py`` (``cpu_baseline`` / ``--impl reference``):
     it mirrors the op sequence TF 1.8 executes for ``model.py:8-98`` — per step per
     layer ``concat([x,h]) -> matmul -> bias_add -> split -> sigmoid/tanh -> cell update``,
     a static unroll over ``seq_length``, reverse-mode autodiff through the unrolled
     graph, TF-style global-norm clip (embedding counted by its un-deduplicated
     slices) and TF-style Adam — on all host threads torch can use.
"""
import math

import numpy as np
import torch


def _t(a, dtype):
    return torch.from_numpy(np.ascontiguousarray(a)).to(dtype)


class TorchStep(object):
    def __init__(self, cfg, params, dtype=torch.float32, learning_rate=0.002, decay_rate=0.97):
        self.cfg, self.dtype = cfg, dtype
        self.names = list(params.keys())
        self.p = {k: _t(v, dtype).requires_grad_(True) for k, v in params.items()}
        self.m = {k: torch.zeros_like(v) for k, v in self.p.items()}
        self.v = {k: torch.zeros_like(v) for k, v in self.p.items()}
        self.b1, self.b2, self.eps = 0.9, 0.999, 1e-8
        self.b1p, self.b2p = self.b1, self.b2
        self.learning_rate, self.decay_rate = learning_rate, decay_rate
        self.lr = 0.0
        self.global_step = 0
        self.state = self.zero_state()

    def zero_state(self):
        c = self.cfg
        n = 2 if c.model == "lstm" else 1
        return [tuple(torch.zeros(c.batch_size, c.rnn_size, dtype=self.dtype) for _ in range(n))
                for _ in range(c.num_layers)]

    def begin_epoch(self, e):
        self.lr = float(np.float32(self.learning_rate) * np.float32(self.decay_rate) ** np.float32(e))
        self.state = self.zero_state()

    def _cell(self, l, x, st):
        c = self.cfg
        pre = "rnnlm/multi_rnn_cell/cell_%d/" % l
        if c.model == "lstm":
            cs, h = st
            z = torch.matmul(torch.cat([x, h], 1), self.p[pre + "lstm_cell/kernel"]) + self.p[pre + "lstm_cell/bias"]
            i, j, f, o = torch.split(z, c.rnn_size, dim=1)
            cn = torch.sigmoid(f + 1.0) * cs + torch.sigmoid(i) * torch.tanh(j)
            hn = torch.sigmoid(o) * torch.tanh(cn)
            return hn, (cn, hn)
        if c.model == "gru":
            (h,) = st
            v = torch.sigmoid(torch.matmul(torch.cat([x, h], 1), self.p[pre + "gru_cell/gates/kernel"])
                              + self.p[pre + "gru_cell/gates/bias"])
            r, u = torch.split(v, c.rnn_size, dim=1)
            cand = torch.tanh(torch.matmul(torch.cat([x, r * h], 1), self.p[pre + "gru_cell/candidate/kernel"])
                              + self.p[pre + "gru_cell/candidate/bias"])
            hn = u * h + (1.0 - u) * cand
            return hn, (hn,)
        (h,) = st
        hn = torch.tanh(torch.matmul(torch.cat([x, h], 1), self.p[pre + "basic_rnn_cell/kernel"])
                        + self.p[pre + "basic_rnn_cell/bias"])
        return hn, (hn,)

    def forward_backward(self, x, y):
        c = self.cfg
        B, S = x.shape
        xt = torch.from_numpy(np.ascontiguousarray(x)).long()
        yt = torch.from_numpy(np.ascontiguousarray(y)).long().reshape(-1)
        for v in self.p.values():
            v.grad = None
        inputs = self.p["embedding"][xt]                         # [B, S, H]
        inputs.retain_grad()
        state = [tuple(s.detach() for s in st) for st in self.state]
        outs = []
        for t in range(S):
            inp = inputs[:, t, :]
            for l in range(c.num_layers):
                inp, state[l] = self._cell(l, inp, state[l])
            outs.append(inp)
        out = torch.cat(outs, 1).reshape(-1, c.rnn_size)
        logits = torch.matmul(out, self.p["rnnlm/softmax_w"]) + self.p["rnnlm/softmax_b"]
        xent = torch.logsumexp(logits, 1) - logits[torch.arange(B * S), yt]
        cost = xent.sum() / B / S
        cost.backward()
        final_state = [tuple(s.detach() for s in st) for st in state]
        return cost.detach(), final_state, inputs.grad, logits.detach()

    def step(self, x, y):
        cost, final_state, dX, _ = self.forward_backward(x, y)
        c = self.cfg
        with torch.no_grad():
            ss = (dX.double() ** 2).sum()
            for k, v in self.p.items():
                if k != "embedding":
                    ss = ss + (v.grad.double() ** 2).sum()
            gn = math.sqrt(float(ss))
            scale = c.grad_clip * min(1.0 / gn if gn > 0 else float("inf"), 1.0 / c.grad_clip)
            alpha = self.lr * math.sqrt(1.0 - self.b2p) / (1.0 - self.b1p)
            for k, v in self.p.items():
                g = v.grad * scale
                self.m[k] += (g - self.m[k]) * (1.0 - self.b1)
                self.v[k] += (g * g - self.v[k]) * (1.0 - self.b2)
                v -= (alpha * self.m[k]) / (torch.sqrt(self.v[k]) + self.eps)
            self.b1p *= self.b1
            self.b2p *= self.b2
        self.global_step += 1
        self.state = final_state
        self.last_gnorm = gn
        return float(cost)

    def grads_numpy(self):
        return {k: v.grad.detach().numpy().copy() for k, v in self.p.items()}

    def params_numpy(self):
        return {k: v.detach().numpy().copy() for k, v in self.p.items()}
