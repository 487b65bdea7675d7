"""The N-replica protocol on CPU (gloo, world_size 2): N ranks + one all-reduce of the flat
gradient buffer == one process on the stacked batch (SURVEY §8 e).  The per-rank gradients come
from the oracle here; on the GPU box the same check runs against the CUDA kernels."""
import os
import socket

import numpy as np
import pytest

torch = pytest.importorskip("torch")
import torch.distributed as dist                                   # noqa: E402
import torch.multiprocessing as mp                                 # noqa: E402

from distributed_char_rnn_b200 import parallel as PL               # noqa: E402
from oracle import char_rnn_oracle as O                            # noqa: E402


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, model, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    H, L, S, B, V = 8, 2, 4, 3, 7
    cfg = O.Config(model, H, L, S, B, V, 0.05)                    # clip small enough to be active
    params = O.init_params(cfg, 0, np.float64)
    rng = np.random.default_rng(100 + rank)                       # rank r's own shard
    x, y = rng.integers(0, V, (B, S)), rng.integers(0, V, (B, S))
    fb = O.forward_backward(cfg, params, x, y, O.zero_state(cfg, np.float64), np.float64)
    lay, P = PL.flat_layout(model, H, L, V)
    flat = torch.from_numpy(PL.pack(fb["grads"], lay, P, tail=(fb["emb_slices_sqnorm"], fb["loss"])).astype(np.float64))
    first, off1 = PL.bucket_split(lay, L)
    assert first == 1 and lay[[n for n, _, _ in lay].index("rnnlm/multi_rnn_cell/cell_1/" + lay[-2][0].split("cell_1/")[1])][1] >= off1
    two = flat.clone()
    PL.allreduce_flat(flat)
    PL.exchange_buckets(two, off1, P)                             # the same exchange cut in two buckets (+ the tail)
    assert torch.equal(two, flat)
    flat = flat.numpy()
    mult = 1.0 / world
    gn = PL.global_norm_from_flat(flat, lay, P, mult)
    np.savez(os.path.join(out_dir, "r%d.npz" % rank), flat=flat, gn=gn, x=x, y=y, loss=flat[P + 1] * mult)
    dist.destroy_process_group()


@pytest.mark.parametrize("model", ["lstm", "gru", "rnn"])
def test_two_replicas_equal_one_process_on_stacked_batch(tmp_path, model):
    world, port = 2, _free_port()
    mp.spawn(_worker, args=(world, port, model, str(tmp_path)), nprocs=world, join=True)
    r = [np.load(os.path.join(str(tmp_path), "r%d.npz" % i)) for i in range(world)]
    assert (r[0]["flat"] == r[1]["flat"]).all()                   # every rank holds the same reduced buffer
    H, L, S, B, V = 8, 2, 4, 3, 7
    cfg = O.Config(model, H, L, S, world * B, V, 0.05)
    params = O.init_params(O.Config(model, H, L, S, B, V), 0, np.float64)
    x, y = np.concatenate([r[0]["x"], r[1]["x"]]), np.concatenate([r[0]["y"], r[1]["y"]])
    fb = O.forward_backward(cfg, params, x, y, O.zero_state(cfg, np.float64), np.float64)
    lay, P = PL.flat_layout(model, H, L, V)
    got = PL.unpack(r[0]["flat"][:P] / world, lay)
    for k, g in fb["grads"].items():
        assert np.abs(got[k] - g).max() <= 1e-6 * max(np.abs(g).max(), 1e-6), k     # the wire format is fp32
    assert abs(float(r[0]["loss"]) - fb["loss"]) < 1e-6 * fb["loss"]
    assert abs(float(r[0]["gn"]) - O.global_norm(cfg, fb)) < 1e-6 * float(r[0]["gn"])   # incl. the un-deduplicated slices term
    assert float(r[0]["gn"]) > 0.05


def test_bucket_split_is_a_contiguous_tail():
    for model, L in (("lstm", 2), ("gru", 3), ("rnn", 4), ("lstm", 1)):
        lay, P = PL.flat_layout(model, 16, L, 9)
        first, off = PL.bucket_split(lay, L)
        if L == 1:
            assert off is None
            continue
        assert first == L - max(1, L // 2)
        early = [n for n, o, _ in lay if o >= off]
        assert early and all("cell_%d/" % l in n for n in early for l in [int(n.split("cell_")[1].split("/")[0])] if l >= first)
        assert all(int(n.split("cell_")[1].split("/")[0]) >= first for n in early)
        assert all(("cell_" not in n) or int(n.split("cell_")[1].split("/")[0]) < first for n, o, _ in lay if o < off)


def test_reference_shards_give_each_rank_223_batches(tmp_path):
    """Config 2: tinyshakespeare split in two by data_splitter, B = S = 50 per rank."""
    import json
    from conftest import ROOT
    ts = json.load(open(os.path.join(ROOT, "tests", "golden", "data_golden.json")))["tinyshakespeare"]
    assert ts["num_batches"] == 446 and ts["shards_concat_equal_full"]
    assert 557697 // (50 * 50) == 223
