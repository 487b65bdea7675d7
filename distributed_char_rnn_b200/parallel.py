"""Host-side protocol of the data-parallel replicas that replace the reference's ps/worker split
(``train.py:117-133,157-163``; SURVEY §8 e).

* one replica per ``--job_name worker`` process, rank = ``--task_index``, world = number of
  ``--worker_hosts`` (or the launcher's ``RANK`` / ``WORLD_SIZE``);
* rank r trains on shard ``data-<r>.npy`` of ``data_splitter.py``;
* per step ONE sum all-reduce of the flat gradient buffer ``[P + 2]``: the P gradient floats in
  ``tf.trainable_variables()`` order plus ``[P] = sum |dX|^2`` of the local embedding slices and
  ``[P+1] =`` local cost; afterwards every rank applies the same clip + Adam with
  ``grad_mult = 1/world`` (global norm ``= mult * sqrt(sum dense^2 + [P])``), so replicas stay
  identical without any further exchange.

Nothing here touches the GPU: ``tests/test_distributed_cpu.py`` drives it over ``gloo``.
"""
import os

import numpy as np

ALIGN = 64          # floats; must match build_layout() in csrc/crnn_api.cu
GRAD_TAIL = 2


def cluster_from_flags(args, env=None):
    """``(rank, world, master_addr, master_port)`` from the reference's flags or torchrun's env."""
    env = os.environ if env is None else env
    workers = args.worker_hosts.split(",") if getattr(args, "worker_hosts", None) else []
    world = int(env.get("WORLD_SIZE", len(workers) or 1))
    rank = int(env.get("RANK", args.task_index if getattr(args, "task_index", None) is not None else 0))
    if "MASTER_ADDR" in env:
        addr, port = env["MASTER_ADDR"], int(env.get("MASTER_PORT", 29500))
    elif workers:
        host, _, p = workers[0].partition(":")
        addr, port = host or "127.0.0.1", int(p or 29500)
    else:
        addr, port = "127.0.0.1", 29500
    if not 0 <= rank < world:
        raise ValueError("task_index %d outside [0, %d)" % (rank, world))
    return rank, world, addr, port


def tvar_shapes(model, H, L, V):
    """Names and shapes in ``tf.trainable_variables()`` order (``model.py:48-54,72``)."""
    out = [("rnnlm/softmax_w", (H, V)), ("rnnlm/softmax_b", (V,)), ("embedding", (V, H))]
    for l in range(L):
        p = "rnnlm/multi_rnn_cell/cell_%d/" % l
        if model == "lstm":
            out += [(p + "lstm_cell/kernel", (2 * H, 4 * H)), (p + "lstm_cell/bias", (4 * H,))]
        elif model == "gru":
            out += [(p + "gru_cell/gates/kernel", (2 * H, 2 * H)), (p + "gru_cell/gates/bias", (2 * H,)),
                    (p + "gru_cell/candidate/kernel", (2 * H, H)), (p + "gru_cell/candidate/bias", (H,))]
        elif model == "rnn":
            out += [(p + "basic_rnn_cell/kernel", (2 * H, H)), (p + "basic_rnn_cell/bias", (H,))]
        else:
            raise Exception("model type not supported: {}".format(model))
    return out


def flat_layout(model, H, L, V):
    """``([(name, offset, shape)], P)``: every tensor starts on a 64-float boundary."""
    off, out = 0, []
    for name, shape in tvar_shapes(model, H, L, V):
        out.append((name, off, shape))
        off = (off + int(np.prod(shape)) + ALIGN - 1) // ALIGN * ALIGN
    return out, off


def pack(tensors, layout, P, tail=(0.0, 0.0)):
    """Dict of arrays -> flat fp32 buffer ``[P + GRAD_TAIL]`` (padding stays zero)."""
    flat = np.zeros(P + GRAD_TAIL, dtype=np.float32)
    for name, off, shape in layout:
        flat[off:off + int(np.prod(shape))] = np.asarray(tensors[name], dtype=np.float32).reshape(-1)
    flat[P], flat[P + 1] = tail
    return flat


def unpack(flat, layout):
    return {name: flat[off:off + int(np.prod(shape))].reshape(shape) for name, off, shape in layout}


def allreduce_flat(flat_tensor, group=None):
    """The one exchange per step (sum; scaling by 1/world happens inside clip+Adam)."""
    import torch.distributed as dist
    dist.all_reduce(flat_tensor, op=dist.ReduceOp.SUM, group=group)
    return flat_tensor


def bucket_split(layout, num_layers):
    """``(first_layer, offset)`` of the EARLY bucket of the gradient exchange: the cells ``first_layer .. L-1`` (the top
    half of the stack) are a contiguous tail ``[offset, P)`` of the flat buffer and are final well before BPTT has
    reached the bottom layers.  ``layout`` rows are ``(name, offset, ...)``.  One layer: no early bucket (``(0, None)``)."""
    if num_layers < 2:
        return 0, None
    first = num_layers - max(1, num_layers // 2)
    key = "cell_%d/" % first
    for row in layout:
        if key in row[0]:
            return first, int(row[1])
    raise KeyError(key)


def exchange_buckets(flat, offset, P, group=None, early_stream=None, early_event=None):
    """Sum all-reduce of the flat gradient buffer ``[P + 2]`` in two buckets: ``[offset, P)`` (the top cells, issued on
    ``early_stream`` as soon as ``early_event`` - recorded by the library inside BPTT, ``crnn_set_bucket_event`` - has
    fired) and the rest ``[0, offset) + [P, P+2)`` on the current stream.  The caller's stream then waits for the
    early bucket.  Without a stream / event (CPU, gloo) both buckets simply follow each other."""
    import torch
    import torch.distributed as dist
    early, rest, tail = flat[offset:P], flat[:offset], flat[P:]
    if early_stream is not None:
        early_stream.wait_event(early_event)
        with torch.cuda.stream(early_stream):
            dist.all_reduce(early, op=dist.ReduceOp.SUM, group=group)
    else:
        dist.all_reduce(early, op=dist.ReduceOp.SUM, group=group)
    if flat.is_cuda:                             # NCCL: the two pieces travel in one grouped launch
        with dist._coalescing_manager(group=group, device=flat.device):
            dist.all_reduce(rest, op=dist.ReduceOp.SUM, group=group)
            dist.all_reduce(tail, op=dist.ReduceOp.SUM, group=group)
    else:
        dist.all_reduce(rest, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(tail, op=dist.ReduceOp.SUM, group=group)
    if early_stream is not None:
        torch.cuda.current_stream(flat.device).wait_stream(early_stream)
    return flat


def global_norm_from_flat(flat, layout, P, mult):
    """What ``crnn_clip_adam`` computes: dense tensors by their own squares, the embedding by
    the slices' sum of squares carried in ``flat[P]`` (``tf.global_norm`` on IndexedSlices)."""
    ss = float(flat[P])
    for name, off, shape in layout:
        if name != "embedding":
            g = flat[off:off + int(np.prod(shape))].astype(np.float64)
            ss += float((g * g).sum())
    return mult * np.sqrt(ss)
