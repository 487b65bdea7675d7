"""Host-side mirror of the reference's ``Model`` (``model.py:8-140``) over ``libcrnn.so``.

The reference object is a bag of TF graph nodes that ``train.py`` feeds and fetches
(``input_data targets initial_state final_state cost train_op lr global_step logits
probs cell sample()``).  Here the same names exist, but the graph is a set of
hand-written sm_100a kernels behind the C-ABI and "running the graph" is
:meth:`Model.step` (= ``sess.run([cost, final_state, train_op], feed)``,
``train.py:199``) or :meth:`Model.forward` (= ``sess.run([probs, final_state], feed)``,
``model.py:124``).  PyTorch is used for device memory, streams and
``torch.distributed`` only.

Data-parallel mode replaces the reference's parameter server (``train.py:126-133``):
every rank owns a full replica, gradients live in ONE flat buffer and are summed with a
single ``all_reduce`` before the fused clip+Adam kernel runs identically on every rank.
"""
import ctypes
import math
import os

import numpy as np
import torch

from . import _lib, parallel
from ._lib import CrnnConfig, GRAD_TAIL, MODEL_IDS, PREC_IDS, check, ptr


class _Cell(object):
    """Stand-in for ``rnn.MultiRNNCell``: only ``zero_state`` is used by the drivers."""

    def __init__(self, model):
        self._m = model

    def zero_state(self, batch_size=None, dtype=None):
        return self._m.zero_state()


class Model(object):
    def __init__(self, args, training=True, device=None, seed=None, precision="fp32", process_group=None):
        self.args = args
        if not training:                       # model.py:11-13
            args.batch_size = 1
            args.seq_length = 1
        if args.model == "nas":
            raise Exception("model type not supported on this engine: nas (rnn.NASCell is out of scope)")
        if args.model not in MODEL_IDS:        # model.py:24-25
            raise Exception("model type not supported: {}".format(args.model))
        # DropoutWrapper per cell + tf.nn.dropout on the embedded inputs, training graphs only (model.py:31-34,58-59)
        ikp = float(getattr(args, "input_keep_prob", 1.0)) if training else 1.0
        okp = float(getattr(args, "output_keep_prob", 1.0)) if training else 1.0
        if not (0.0 < ikp <= 1.0 and 0.0 < okp <= 1.0):
            raise Exception("keep probabilities must be in (0, 1]")
        if not torch.cuda.is_available():
            raise _lib.CrnnError("no CUDA device: distributed_char_rnn_b200 has no CPU path")
        self.training = training
        self.device = torch.device(device if device is not None else "cuda:%d" % torch.cuda.current_device())
        torch.cuda.set_device(self.device)
        self._lib = _lib.load()
        self.pg = process_group
        self.world = torch.distributed.get_world_size(process_group) if process_group is not None else 1

        # every replica draws its own masks for its own shard (the reference's workers do: independent TF RNGs); the
        # variable initialiser below stays on `seed` alone, so all replicas start from the same weights
        rank = torch.distributed.get_rank(process_group) if process_group is not None else 0
        entropy = None if seed is None else [int(seed), int(rank)]
        self.dropout_seed = int(np.random.SeedSequence(entropy).generate_state(1)[0] & 0x7FFFFFFF) if (ikp < 1 or okp < 1) else 0
        # "auto": the fp32-grade tensor-core mode where the shape allows it (rnn_size % 16 == 0 and the hi + lo
        # recurrent slices fit in shared memory), else the fp32 FFMA kernels - both CUDA, both at the 1e-4 parity bar.
        candidates = ["bf16x3", "fp32"] if precision == "auto" else [precision]
        h = ctypes.c_void_p()
        for i, prec in enumerate(candidates):
            cfg = CrnnConfig(MODEL_IDS[args.model], args.rnn_size, args.num_layers, args.seq_length, args.batch_size,
                             args.vocab_size, PREC_IDS[prec], self.device.index or 0, ikp, okp, self.dropout_seed, 0)
            rc = self._lib.crnn_create(ctypes.byref(cfg), ctypes.byref(h))
            if rc == 0:
                break
            if rc != _lib.CRNN_ERR_UNSUP or i + 1 == len(candidates):
                check(rc)
        self.precision = prec
        self.training_calls = 0                 # crnn_forward_backward calls so far = index of the next dropout mask set
        self._h = h
        self.P = int(self._lib.crnn_flat_size(h))
        self.layout = []
        name = ctypes.create_string_buffer(256)
        off, rows, cols = ctypes.c_int64(), ctypes.c_int64(), ctypes.c_int64()
        for i in range(self._lib.crnn_num_tensors(h)):
            check(self._lib.crnn_param_layout(h, i, name, 256, ctypes.byref(off), ctypes.byref(rows), ctypes.byref(cols)))
            self.layout.append((name.value.decode(), off.value, rows.value, cols.value))

        dev, f32 = self.device, torch.float32
        self.params = torch.zeros(self.P, dtype=f32, device=dev)
        self.grads = torch.zeros(self.P + GRAD_TAIL, dtype=f32, device=dev)
        self.adam_m = torch.zeros(self.P, dtype=f32, device=dev) if training else None
        self.adam_v = torch.zeros(self.P, dtype=f32, device=dev) if training else None
        self.workspace = torch.empty(int(self._lib.crnn_workspace_bytes(h)), dtype=torch.uint8, device=dev)
        self._nstate = 2 if args.model == "lstm" else 1
        self._state = torch.zeros(int(self._lib.crnn_state_size(h)), dtype=f32, device=dev)
        self._gnorm = torch.zeros(1, dtype=f32, device=dev)
        B, S = args.batch_size, args.seq_length
        # pinned staging for ids and the loss: the feed/fetch of train.py:194-199
        self._xy_host = torch.empty((2, B, S), dtype=torch.int32).pin_memory()
        self._xy_dev = torch.empty((2, B, S), dtype=torch.int32, device=dev)
        self._loss_host = torch.empty(1, dtype=f32).pin_memory()
        self._xy_busy = None                    # event: the device copy of the pinned ids has been issued and not yet awaited
        # Gradient exchange of the replicas in two buckets (CRNN_AR_OVERLAP=1): the top cells' bucket goes out while BPTT is
        # still in the lower layers.  Measured on B200, config 3: it LOSES - 2 GPUs 4.70 -> 4.74 ms per step, 8 GPUs 4.75 ->
        # 4.80 ms (twice each) - the NCCL kernel takes SMs and L2 bandwidth from the latency-bound recurrence kernels it
        # runs beside, which costs more than the ~0.06 ms of exchange it hides.  Off by default: one all-reduce after BPTT.
        self._ar_offset = self._ar_stream = self._ar_event = None
        if self.world > 1 and training and int(os.environ.get("CRNN_AR_OVERLAP", "0")):
            first, off1 = parallel.bucket_split(self.layout, args.num_layers)
            if off1 is not None:
                self._ar_offset, self._ar_stream = off1, torch.cuda.Stream(device=dev)
                self._ar_event = torch.cuda.Event()
                self._ar_event.record()         # (creates the CUDA event whose handle the library records)
                check(self._lib.crnn_set_bucket_event(h, first, ctypes.c_void_p(self._ar_event.cuda_event)))
        self.input_data, self.targets = self._xy_host[0], self._xy_host[1]
        self.cell = _Cell(self)
        self.lr = 0.0                          # model.py:87
        self.global_step = 0
        self.cost = None
        self.final_state = self._state
        self.initial_state = self.zero_state()
        self._logits = self._probs = None
        self.init_variables(seed)

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                self._lib.crnn_destroy(self._h)
                self._h = None
        except Exception:
            pass

    # -- variables ---------------------------------------------------------------------
    def init_variables(self, seed=None):
        """TF default initialisers: glorot-uniform kernels, embedding and
        ``softmax_b``; zero cell biases; GRU gate bias 1.0.  ``global_variables_initializer``."""
        rng = np.random.Generator(np.random.PCG64(seed))
        flat = np.zeros(self.P, dtype=np.float32)
        for name, off, rows, cols in self.layout:
            n = rows * cols
            if name.endswith("/bias"):
                flat[off:off + n] = 1.0 if name.endswith("gates/bias") else 0.0
                continue
            fan_in, fan_out = (cols, cols) if name.endswith("softmax_b") else (rows, cols)
            lim = math.sqrt(6.0 / (fan_in + fan_out))
            flat[off:off + n] = rng.uniform(-lim, lim, size=n).astype(np.float32)
        self.params.copy_(torch.from_numpy(flat))
        if self.adam_m is not None:
            self.adam_m.zero_()
            self.adam_v.zero_()
        self.global_step = 0
        self.sync_replicas()

    def sync_replicas(self):
        """All ranks start from rank 0's variables (the chief initialises, ``train.py:157-165``)."""
        if self.pg is not None:
            torch.distributed.broadcast(self.params, src=0, group=self.pg)

    def variable_names(self):
        return [t[0] for t in self.layout]

    def _view(self, flat, name):
        for n, off, rows, cols in self.layout:
            if n == name:
                v = flat[off:off + rows * cols]
                return v if rows == 1 else v.view(rows, cols)
        raise KeyError(name)

    def get_variable(self, name):
        return self._view(self.params, name)

    def get_gradient(self, name):
        return self._view(self.grads, name)

    def set_variables(self, values):
        """Load ``{tf_variable_name: array}`` (e.g. the oracle's initial weights)."""
        for name, val in values.items():
            dst = self.get_variable(name)
            dst.copy_(torch.as_tensor(np.ascontiguousarray(val, dtype=np.float32)).reshape(dst.shape))

    def variables_numpy(self):
        return {n: self.get_variable(n).detach().cpu().numpy().copy() for n in self.variable_names()}

    def gradients_numpy(self):
        return {n: self.get_gradient(n).detach().cpu().numpy().copy() for n in self.variable_names()}

    # -- state -------------------------------------------------------------------------
    def zero_state(self):
        """``cell.zero_state`` (``model.py:43``): flat ``[L][n][B][H]`` zeros on the device."""
        return torch.zeros_like(self._state)

    def state_tuples(self, state=None):
        """The reference's view of a state: per layer ``(c, h)`` for lstm, ``(h,)`` otherwise."""
        a = self.args
        s = (self._state if state is None else state).view(a.num_layers, self._nstate, a.batch_size, a.rnn_size)
        return [tuple(s[l, i] for i in range(self._nstate)) for l in range(a.num_layers)]

    def assign_lr(self, value):
        """``tf.assign(model.lr, ...)`` (``train.py:147-148``): kept as an fp32 value."""
        self.lr = float(np.float32(value))

    # -- running the graph -------------------------------------------------------------
    def _stage_ids(self, x, y=None):
        """Feed: host ids -> pinned staging -> device (counted in ``h2d_bytes_per_step``)."""
        if isinstance(x, torch.Tensor) and x.is_cuda:
            xd = x if x.dtype == torch.int32 else x.to(torch.int32)
            yd = None if y is None else (y if y.dtype == torch.int32 else y.to(torch.int32))
            return xd, yd, 0
        if self._xy_busy is not None:           # the previous step's copy may still be reading the pinned buffer
            self._xy_busy.synchronize()
        else:
            self._xy_busy = torch.cuda.Event()
        self._xy_host[0].copy_(torch.as_tensor(np.asarray(x)))
        if y is not None:
            self._xy_host[1].copy_(torch.as_tensor(np.asarray(y)))
            self._xy_dev.copy_(self._xy_host, non_blocking=True)
            self._xy_busy.record()
            return self._xy_dev[0], self._xy_dev[1], self._xy_host.numel() * 4
        self._xy_dev[0].copy_(self._xy_host[0], non_blocking=True)
        self._xy_busy.record()
        return self._xy_dev[0], None, self._xy_host[0].numel() * 4

    @staticmethod
    def _ld(t):
        assert t.dim() == 2 and t.stride(1) == 1, "ids must be a [B,S] matrix with unit column stride"
        return t.stride(0)

    def step_async(self, x, y, state=None):
        """Enqueue one training step; returns the device scalar holding the cost.

        ``x``/``y``: ``[B,S]`` ids (numpy / CPU tensors are staged through pinned memory;
        CUDA int32 tensors, e.g. column slices of the loader matrix, are used in place).
        ``state``: flat device state to start from (default: the carried state, which
        is what ``train.py:194-199`` feeds back every batch).
        """
        assert self.training
        xd, yd, _ = self._stage_ids(x, y)
        assert self._ld(xd) == self._ld(yd)
        if state is not None and state is not self._state:
            self._state.copy_(state)
        st = ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        self.global_step += 1
        self.training_calls += 1
        a = self.args
        if self.world == 1:
            check(self._lib.crnn_train_step(self._h, ptr(xd), ptr(yd), self._ld(xd), ptr(self._state),
                                            ptr(self.params), ptr(self.grads), ptr(self.adam_m), ptr(self.adam_v),
                                            self.lr, a.grad_clip, self.global_step, ptr(self.workspace), st))
        else:
            check(self._lib.crnn_forward_backward(self._h, ptr(xd), ptr(yd), self._ld(xd), ptr(self._state),
                                                  ptr(self._state), ptr(self.params), ptr(self.grads),
                                                  ptr(self.workspace), st))
            if self._ar_offset is None:
                parallel.allreduce_flat(self.grads, self.pg)             # the ONE exchange per step
            else:                                                        # ... cut in two so its first half hides behind BPTT's tail
                parallel.exchange_buckets(self.grads, self._ar_offset, self.P, self.pg, self._ar_stream, self._ar_event)
            check(self._lib.crnn_clip_adam(self._h, ptr(self.params), ptr(self.grads), ptr(self.adam_m),
                                           ptr(self.adam_v), self.lr, a.grad_clip, self.global_step,
                                           1.0 / self.world, ptr(self._gnorm), ptr(self.workspace), st))
        self.final_state = self._state
        return self.grads[self.P + 1]

    def step(self, x, y, state=None):
        """``sess.run([cost, final_state, train_op], feed)`` (``train.py:199``) -> ``(cost, final_state)``."""
        loss_dev = self.step_async(x, y, state)
        self._loss_host.copy_(loss_dev.view(1), non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        self.cost = float(self._loss_host[0]) / self.world
        return self.cost, self._state

    def gradients(self, x, y, state=None):
        """``tf.gradients(cost, tvars)`` only (no update); leaves them in ``self.grads``."""
        xd, yd, _ = self._stage_ids(x, y)
        sin = self._state if state is None else state
        sout = torch.empty_like(self._state)
        st = ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        check(self._lib.crnn_forward_backward(self._h, ptr(xd), ptr(yd), self._ld(xd), ptr(sin), ptr(sout),
                                              ptr(self.params), ptr(self.grads), ptr(self.workspace), st))
        self.training_calls += 1
        torch.cuda.current_stream(self.device).synchronize()
        return float(self.grads[self.P + 1]), sout

    def dropout_masks(self, call_index):
        """The multipliers training call ``call_index`` applies: ``L`` cell-input sites + the top output,
        each ``[S, B, H]`` (time-major).  Lets a run with dropout be replayed on the CPU oracle."""
        a = self.args
        out = []
        for site in range(a.num_layers + 1):
            m = torch.empty((a.seq_length, a.batch_size, a.rnn_size), dtype=torch.float32, device=self.device)
            check(self._lib.crnn_dropout_mask(self._h, call_index, site, ptr(m), None))
            out.append(m.cpu().numpy())
        torch.cuda.synchronize()
        return out

    def forward(self, x, state=None, want_logits=True, want_probs=True):
        """``sess.run([logits, probs, final_state], feed)``: rows ordered ``b*S+t`` (``model.py:73``)."""
        a = self.args
        xd, _, _ = self._stage_ids(x)
        sin = self._state if state is None else state
        sout = torch.empty_like(self._state)
        n = a.batch_size * a.seq_length
        logits = torch.empty((n, a.vocab_size), dtype=torch.float32, device=self.device) if want_logits else None
        probs = torch.empty((n, a.vocab_size), dtype=torch.float32, device=self.device) if want_probs else None
        st = ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        check(self._lib.crnn_forward(self._h, ptr(xd), self._ld(xd), ptr(sin), ptr(sout), ptr(self.params),
                                     ptr(logits), ptr(probs), ptr(self.workspace), st))
        self._logits, self._probs = logits, probs
        return logits, probs, sout

    @property
    def logits(self):
        return self._logits

    @property
    def probs(self):
        return self._probs

    # -- sampling (model.py:105-140) ---------------------------------------------------
    def sample(self, sess, chars, vocab, num=200, prime='The ', sampling_type=1, rng=None):
        """Same contract as the reference; ``sess`` is ignored (there is no session)."""
        rng = np.random if rng is None else rng
        state = self.zero_state()
        x = np.zeros((1, 1), dtype=np.int32)
        for char in prime[:-1]:
            x[0, 0] = vocab[char]
            _, _, state = self.forward(x, state, want_logits=False, want_probs=False)

        def weighted_pick(weights):
            t = np.cumsum(weights)
            s = np.sum(weights)
            return int(np.searchsorted(t, rng.rand(1) * s))

        ret = prime
        char = prime[-1]
        for _ in range(num):
            x[0, 0] = vocab[char]
            _, probs, state = self.forward(x, state, want_logits=False)
            p = probs[0].cpu().numpy()
            if sampling_type == 0:
                pick = int(np.argmax(p))
            elif sampling_type == 2:
                pick = weighted_pick(p) if char == ' ' else int(np.argmax(p))
            else:
                pick = weighted_pick(p)
            pick = min(pick, len(chars) - 1)
            char = chars[pick]
            ret += char
        return ret

    # -- checkpoint payload (train.py:145: Saver(tf.global_variables())) ---------------
    def state_dict(self):
        d = {"variables": self.variables_numpy(), "global_step": self.global_step, "lr": self.lr}
        if self.adam_m is not None:
            d["adam_m"] = {n: self._view(self.adam_m, n).cpu().numpy().copy() for n in self.variable_names()}
            d["adam_v"] = {n: self._view(self.adam_v, n).cpu().numpy().copy() for n in self.variable_names()}
        return d

    def load_state_dict(self, d):
        self.set_variables(d["variables"])
        self.global_step = int(d.get("global_step", 0))
        self.lr = float(d.get("lr", 0.0))
        if self.adam_m is not None and "adam_m" in d:
            for n in self.variable_names():
                self._view(self.adam_m, n).copy_(torch.as_tensor(d["adam_m"][n]).reshape(self._view(self.adam_m, n).shape))
                self._view(self.adam_v, n).copy_(torch.as_tensor(d["adam_v"][n]).reshape(self._view(self.adam_v, n).shape))
