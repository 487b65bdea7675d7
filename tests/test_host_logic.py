"""Host-side mirror of the reference's drivers: CLI surface, on-disk files, checkpoint naming."""
import argparse
import json
import os
import pickle

import numpy as np
import pytest

from conftest import ROOT
from distributed_char_rnn_b200 import checkpoint as C
from distributed_char_rnn_b200 import data_splitter, parallel, sample, train

GOLD = json.load(open(os.path.join(ROOT, "tests", "golden", "cli_golden.json")))


def _actions(parser):
    return {a.option_strings[-1]: a for a in parser._actions if a.option_strings and a.dest != "help"}


@pytest.mark.parametrize("name,parser", [("train.py", train.build_parser()), ("sample.py", sample.build_parser()),
                                          ("data_splitter.py", data_splitter.build_parser())])
def test_cli_flags_and_defaults_match_reference(name, parser):
    acts = _actions(parser)
    for flag, spec in GOLD[name].items():
        assert flag in acts, "%s lacks %s" % (name, flag)
        a = acts[flag]
        assert list(a.option_strings) == spec["names"]
        if spec.get("action") == "store_true":
            assert a.default is False and a.nargs == 0
        else:
            assert a.default == spec.get("default"), flag
        if "choices" in spec:
            assert list(a.choices) == spec["choices"]
        if spec.get("required"):
            assert a.required
        if spec.get("type") in ("int", "float"):
            assert a.type.__name__ == spec["type"]


def test_train_defaults_are_config1():
    a = train.build_parser().parse_args([])
    assert (a.model, a.rnn_size, a.num_layers, a.seq_length, a.batch_size) == ("lstm", 128, 2, 50, 50)
    assert (a.grad_clip, a.learning_rate, a.decay_rate, a.num_epochs, a.save_every) == (5.0, 0.002, 0.97, 50, 1000)


def test_worker_without_tensor_file_raises_like_reference(tmp_path):
    a = train.build_parser().parse_args(["--distributed", "--job_name", "worker", "--task_index", "0",
                                         "--worker_hosts", "127.0.0.1:9000", "--data_dir", str(tmp_path)])
    with pytest.raises(Exception, match="Tensor file must be provided."):
        train.train(a)


def test_ps_job_is_a_noop(capsys):
    a = train.build_parser().parse_args(["--distributed", "--job_name", "ps", "--task_index", "0"])
    assert train.train(a) == 0
    assert "no parameter server" in capsys.readouterr().out


def test_cluster_from_flags():
    ns = argparse.Namespace(worker_hosts="10.0.0.1:9000,10.0.0.2:9001", task_index=1)
    assert parallel.cluster_from_flags(ns, env={}) == (1, 2, "10.0.0.1", 9000)
    assert parallel.cluster_from_flags(ns, env={"RANK": "0", "WORLD_SIZE": "8", "MASTER_ADDR": "127.0.0.1",
                                                "MASTER_PORT": "1234"}) == (0, 8, "127.0.0.1", 1234)
    with pytest.raises(ValueError):
        parallel.cluster_from_flags(argparse.Namespace(worker_hosts="a:1", task_index=3), env={})


def test_flat_layout_is_tvars_order_and_aligned():
    lay, P = parallel.flat_layout("lstm", 128, 2, 65)
    assert [n for n, _, _ in lay][:3] == ["rnnlm/softmax_w", "rnnlm/softmax_b", "embedding"]
    assert all(off % 64 == 0 for _, off, _ in lay) and P % 64 == 0
    assert sum(int(np.prod(s)) for _, _, s in lay) == 279873
    g = {n: np.random.default_rng(0).normal(size=s).astype(np.float32) for n, _, s in lay}
    flat = parallel.pack(g, lay, P, tail=(3.0, 4.0))
    back = parallel.unpack(flat, lay)
    assert all((back[n] == g[n]).all() for n in g) and flat[P] == 3.0 and flat[P + 1] == 4.0
    used = np.zeros(P, bool)
    for n, off, s in lay:
        used[off:off + int(np.prod(s))] = True
    assert (flat[:P][~used] == 0).all()
    with pytest.raises(Exception, match="model type not supported"):
        parallel.flat_layout("nas", 8, 1, 5)


def _state(seed=0):
    rng = np.random.default_rng(seed)
    names = ["embedding", "rnnlm/softmax_w"]
    return {"variables": {n: rng.normal(size=(3, 4)).astype(np.float32) for n in names},
            "adam_m": {n: rng.normal(size=(3, 4)).astype(np.float32) for n in names},
            "adam_v": {n: rng.random(size=(3, 4)).astype(np.float32) for n in names},
            "global_step": 15, "lr": 0.0015}


def test_checkpoint_files_and_roundtrip(tmp_path):
    d = str(tmp_path)
    assert C.latest_checkpoint(d) is None
    assert C.save(d, _state(0), 0) == os.path.join(d, "model.ckpt")          # what train.py prints
    C.save(d, _state(1), 14)
    for f in ("model.ckpt-0.npz", "model.ckpt-0.index", "model.ckpt-14.npz", "model.ckpt-14.index", "checkpoint"):
        assert os.path.getsize(os.path.join(d, f)) > 0                       # .travis.yml: test -s model.ckpt-14.index
    text = open(os.path.join(d, "checkpoint")).read()
    assert text.startswith('model_checkpoint_path: "model.ckpt-14"') and 'all_model_checkpoint_paths: "model.ckpt-0"' in text
    latest = C.latest_checkpoint(d)
    assert latest == os.path.join(d, "model.ckpt-14")
    got, want = C.load(latest), _state(1)
    assert got["global_step"] == 15 and abs(got["lr"] - 0.0015) < 1e-9
    for part in ("variables", "adam_m", "adam_v"):
        for n in want[part]:
            assert (got[part][n] == want[part][n]).all()


def test_init_from_asserts_match_reference_messages(tmp_path):
    d = str(tmp_path / "save")
    a = train.build_parser().parse_args(["--init_from", d])
    loader = argparse.Namespace(chars=("a", "b"), vocab={"a": 0, "b": 1})
    with pytest.raises(AssertionError, match="must be a a path"):
        train.check_init_from(a, loader)
    os.makedirs(d)
    with pytest.raises(AssertionError, match="config.pkl file does not exist"):
        train.check_init_from(a, loader)
    saved = train.build_parser().parse_args(["--rnn_size", "64"])
    pickle.dump(saved, open(os.path.join(d, "config.pkl"), "wb"))
    with pytest.raises(AssertionError, match="chars_vocab.pkl.pkl file does not exist"):
        train.check_init_from(a, loader)
    pickle.dump((("a", "b"), {"a": 0, "b": 1}), open(os.path.join(d, "chars_vocab.pkl"), "wb"))
    with pytest.raises(AssertionError, match="No checkpoint found"):
        train.check_init_from(a, loader)
    C.save(d, _state(), 3)
    with pytest.raises(AssertionError, match="disagree on 'rnn_size'"):
        train.check_init_from(a, loader)
    pickle.dump(train.build_parser().parse_args([]), open(os.path.join(d, "config.pkl"), "wb"))
    assert train.check_init_from(a, loader).endswith("model.ckpt-3")
    with pytest.raises(AssertionError, match="character set"):
        train.check_init_from(a, argparse.Namespace(chars=("a",), vocab={"a": 0}))


def test_nas_and_bad_keep_prob_are_rejected_explicitly():
    torch = pytest.importorskip("torch")
    from distributed_char_rnn_b200.model import Model
    ns = argparse.Namespace(model="nas", rnn_size=8, num_layers=1, seq_length=2, batch_size=2, vocab_size=5, grad_clip=5.0)
    with pytest.raises(Exception, match="nas"):
        Model(ns)
    ns = argparse.Namespace(model="bogus", rnn_size=8, num_layers=1, seq_length=2, batch_size=2, vocab_size=5, grad_clip=5.0)
    with pytest.raises(Exception, match="model type not supported: bogus"):       # model.py:25
        Model(ns)
    ns = argparse.Namespace(model="lstm", rnn_size=8, num_layers=1, seq_length=2, batch_size=2, vocab_size=5,
                            grad_clip=5.0, output_keep_prob=0.0, input_keep_prob=1.0)
    with pytest.raises(Exception, match="keep probabilities"):
        Model(ns)
