"""Bit-exact integer contract: vocab construction, id tensor, batches, shards.

Two anchors: (1) golden vectors produced by the REFERENCE's own utils.py / data_splitter.py
(tests/golden/make_data_golden.py) — run everywhere; (2) the reference modules themselves,
imported from /root/reference when present (build container only).
"""
import hashlib
import importlib.util
import json
import os
import pickle
import shutil
import subprocess
import sys

import numpy as np
import pytest

from conftest import REFERENCE, ROOT, have_reference
from distributed_char_rnn_b200 import utils as U
from distributed_char_rnn_b200 import data_splitter as DS

GOLD = json.load(open(os.path.join(ROOT, "tests", "golden", "data_golden.json")))


def _write(tmp_path, text):
    d = str(tmp_path)
    with open(os.path.join(d, "input.txt"), "wb") as f:
        f.write(text.encode("utf-8"))
    return d


@pytest.mark.parametrize("name", sorted(GOLD["corpora"]))
def test_golden_corpora(tmp_path, name):
    g = GOLD["corpora"][name]
    d = _write(tmp_path, g["text"])
    ld = U.TextLoader(d, g["batch_size"], g["seq_length"])
    assert list(ld.chars) == g["chars"]
    assert ld.vocab == {c: i for i, c in enumerate(g["chars"])}
    assert ld.vocab_size == g["vocab_size"] and ld.num_batches == g["num_batches"]
    saved = np.load(os.path.join(d, "data.npy"))
    assert str(saved.dtype) == g["tensor_dtype"] and saved.tolist() == g["tensor"]
    assert pickle.load(open(os.path.join(d, "vocab.pkl"), "rb")) == tuple(g["chars"])
    assert ld.x_batches[0].tolist() == g["x0"] and ld.y_batches[0].tolist() == g["y0"]
    assert ld.x_batches[-1].tolist() == g["x_last"] and ld.y_batches[-1].tolist() == g["y_last"]
    # wrap-around label and contiguity of row streams across batches (utils.py:75-82)
    assert ld.y_batches[-1][-1, -1] == ld.tensor[0]
    if ld.num_batches > 1:
        S = g["seq_length"]
        assert ld.x_batches[1][0, 0] == ld.tensor[S]
    # reload path gives the same object state
    ld2 = U.TextLoader(d, g["batch_size"], g["seq_length"])
    assert ld2.chars == ld.chars and (ld2.tensor == ld.tensor).all()
    X, Y = ld.batch_matrices()
    assert X.dtype == np.int32 and (X[:, :g["seq_length"]] == np.array(g["x0"])).all()
    assert (Y[:, -g["seq_length"]:] == np.array(g["y_last"])).all()
    # shards
    for p in g["split_ok"]:
        od = os.path.join(d, "sh%d" % p)
        DS.split(d, p, od)
        for i in range(p):
            a = np.load(os.path.join(od, "data-%d.npy" % i))
            assert hashlib.sha256(a.astype("<i8").tobytes()).hexdigest() == g["shards"][str(p)][i]
    for p in g["split_bad"]:
        with pytest.raises(ValueError):
            DS.split(d, p, os.path.join(d, "bad"))


def test_next_batch_pointer(tmp_path):
    g = GOLD["corpora"]["ascii"]
    ld = U.TextLoader(_write(tmp_path, g["text"]), 4, 7)
    seen = [ld.next_batch() for _ in range(ld.num_batches)]
    assert ld.pointer == ld.num_batches
    ld.reset_batch_pointer()
    x, y = ld.next_batch()
    assert (x == seen[0][0]).all() and (y == seen[0][1]).all() and ld.pointer == 1


def test_not_enough_data(tmp_path):
    d = _write(tmp_path, "tiny")
    with pytest.raises(AssertionError, match="Not enough data"):
        U.TextLoader(d, 50, 50)


def test_empty_corpus(tmp_path):
    d = _write(tmp_path, "")
    with pytest.raises(ValueError):
        U.create_vocab_file(os.path.join(d, "input.txt"), os.path.join(d, "vocab.pkl"))


def test_shard_never_overwritten(tmp_path):
    """a worker given a shard but no vocab.pkl must not clobber the shard."""
    d = _write(tmp_path, GOLD["corpora"]["ascii"]["text"])
    shard = os.path.join(d, "data-0.npy")
    np.save(shard, np.arange(100))
    before = open(shard, "rb").read()
    with pytest.raises(Exception, match="data_splitter"):
        U.TextLoader(d, 2, 5, tensor_file=shard)
    assert open(shard, "rb").read() == before


def test_int32_shards_accepted(tmp_path):
    """Shipped shards are <i4 (made on Windows); both dtypes must load."""
    d = _write(tmp_path, GOLD["corpora"]["ascii"]["text"])
    ld = U.TextLoader(d, 4, 7)
    shard = os.path.join(d, "s.npy")
    np.save(shard, ld.tensor.astype("<i4"))
    ld2 = U.TextLoader(d, 4, 7, tensor_file=shard)
    assert (ld2.tensor == ld.tensor).all() and ld2.vocab == ld.vocab


@pytest.mark.skipif(not have_reference(), reason="/root/reference only exists in the build container")
class TestAgainstReference:
    def _ref(self):
        spec = importlib.util.spec_from_file_location("ref_utils", os.path.join(REFERENCE, "utils.py"))
        m = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(m)
        return m

    def test_tinyshakespeare_known_answers(self, tmp_path):
        ts = GOLD["tinyshakespeare"]
        d = str(tmp_path)
        shutil.copy(os.path.join(REFERENCE, "data/tinyshakespeare/input.txt"), d)
        ld = U.TextLoader(d, 50, 50)
        assert "".join(ld.chars) == ts["chars"] and ld.vocab_size == 65
        full = np.load(os.path.join(d, "data.npy"))            # ld.tensor is truncated by create_batches
        assert full.size == 1115394 and ld.tensor.size == 1115000
        assert hashlib.sha256(full.astype("<i4").tobytes()).hexdigest() == ts["tensor_i4_sha256"]
        assert ts["tensor_i4_sha256"].startswith("deea0d6a1825f056")
        assert ld.num_batches == 446 == ts["num_batches"]
        assert ld.x_batches[0][0, :5].tolist() == [49, 9, 7, 6, 2] == ts["x0_row0"]
        assert ld.y_batches[0][0, :5].tolist() == [9, 7, 6, 2, 0] == ts["y0_row0"]
        assert ld.x_batches[1][0, :5].tolist() == ts["x1_row0"] and ld.x_batches[0][1, :5].tolist() == ts["x0_row1"]
        assert ld.y_batches[445][49, -1] == 49 == ts["y_last_wrap"]
        # workers on the shipped shards: 223 batches each
        for i in range(2):
            w = U.TextLoader(d, 50, 50, tensor_file=os.path.join(REFERENCE, "sharded_data/data-%d.npy" % i))
            assert w.num_batches == 223 and w.vocab_size == 65
        # our splitter reproduces the shipped shards
        od = os.path.join(d, "sh")
        DS.split(d, 2, od)
        for i in range(2):
            a = np.load(os.path.join(od, "data-%d.npy" % i))
            assert hashlib.sha256(a.astype("<i4").tobytes()).hexdigest() == ts["shipped_shard_sha256"][i]
        for bad in (4, 8):
            with pytest.raises(ValueError):
                DS.split(d, bad, os.path.join(d, "bad"))

    def test_file_bytes_and_objects_equal_reference(self, tmp_path):
        ref = self._ref()
        for name, g in GOLD["corpora"].items():
            d1, d2 = str(tmp_path / (name + "a")), str(tmp_path / (name + "b"))
            for d in (d1, d2):
                os.makedirs(d)
                with open(os.path.join(d, "input.txt"), "wb") as f:
                    f.write(g["text"].encode("utf-8"))
            a, b = U.TextLoader(d1, 3, 5), ref.TextLoader(d2, 3, 5)
            assert a.chars == b.chars and a.vocab == b.vocab and a.num_batches == b.num_batches
            assert a.tensor.dtype == b.tensor.dtype and (a.tensor == b.tensor).all()
            for k in range(a.num_batches):
                assert (a.x_batches[k] == b.x_batches[k]).all() and (a.y_batches[k] == b.y_batches[k]).all()
            for fn in ("vocab.pkl", "data.npy"):
                assert open(os.path.join(d1, fn), "rb").read() == open(os.path.join(d2, fn), "rb").read()

    def test_cli_splitter_equals_reference_script(self, tmp_path):
        g = GOLD["corpora"]["ascii"]
        outs = []
        for k, script in enumerate([os.path.join(ROOT, "data_splitter.py"), os.path.join(REFERENCE, "data_splitter.py")]):
            d = str(tmp_path / ("d%d" % k))
            os.makedirs(d)
            with open(os.path.join(d, "input.txt"), "wb") as f:
                f.write(g["text"].encode("utf-8"))
            subprocess.check_call([sys.executable, script, "--data_dir", d, "--num_parts", "3", "--out_dir",
                                   os.path.join(d, "o")], stdout=subprocess.DEVNULL, cwd=ROOT,
                                  env=dict(os.environ, PYTHONPATH=ROOT))
            outs.append([open(os.path.join(d, "o", "data-%d.npy" % i), "rb").read() for i in range(3)]
                        + [open(os.path.join(d, "vocab.pkl"), "rb").read()])
        assert outs[0] == outs[1]
