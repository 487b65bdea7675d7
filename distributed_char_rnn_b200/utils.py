"""Data contract of the char-RNN training path: vocabulary, id tensor, batches.

Host-side mirror of the reference's ``utils.py`` (same names, same attributes,
same files on disk), rebuilt around a code-point lookup table so that a 1 GB
corpus is handled as flat numpy arrays instead of one Python object per
character (reference: ``utils.py:48`` / ``data_splitter.py:29`` use
``list(map(vocab.get, data))``).

Everything in here is integer work and must be bit-identical to the reference:
``tests/test_data_contract.py`` imports the reference module and compares.
"""
import codecs
import os
import pickle

import numpy as np


def _code_points(text):
    """1-D uint32 array of the code points of ``text`` (one entry per ``str`` item)."""
    return np.frombuffer(text.encode("utf-32-le", "surrogatepass"), dtype=np.uint32)


def _vocab_from_text(data):
    """Vocabulary ordered by descending count, ties by first occurrence.

    Reference ``utils.py:10-14``: ``Counter(data)`` iterates in first-occurrence
    order and ``sorted(..., key=-count)`` is stable, so ties keep that order.
    Returns ``(chars, vocab, cp, cps_sorted)``.
    """
    cp = _code_points(data)
    if cp.size == 0:
        # the reference dies in ``zip(*count_pairs)`` with a ValueError on empty input
        raise ValueError("not enough values to unpack (expected 2, got 0)")
    uniq, first, counts = np.unique(cp, return_index=True, return_counts=True)
    order = np.lexsort((first, -counts.astype(np.int64)))
    cps_sorted = uniq[order]
    chars = tuple(chr(int(c)) for c in cps_sorted)
    vocab = dict(zip(chars, range(len(chars))))
    return chars, vocab, cp, cps_sorted


def encode_text(data, vocab=None, _cache=None):
    """ids of ``data`` under ``vocab`` as a 1-D int64 array (``utils.py:48``).

    ``np.array(list_of_python_ints)`` is int64 on Linux, which is what the
    reference writes to ``data.npy``; shipped shards are ``<i4`` (made on
    Windows) and both are accepted on load.
    """
    cp, cps_sorted = _cache if _cache is not None else (_code_points(data), None)
    if cps_sorted is None:
        cps_sorted = np.array([ord(c) for c in sorted(vocab, key=vocab.get)], dtype=np.uint32)
    lut = np.full(int(cps_sorted.max()) + 1, -1, dtype=np.int64)
    lut[cps_sorted] = np.arange(cps_sorted.size, dtype=np.int64)
    if cp.size and int(cp.max()) >= lut.size:
        raise KeyError("character outside the vocabulary")
    ids = lut[cp]
    if ids.size and ids.min() < 0:
        raise KeyError("character outside the vocabulary")
    return ids


def create_vocab_file(input_file, vocab_file, encoding="utf-8"):
    """Reference ``utils.py:7-17``: read text, build vocab, pickle ``chars``.

    Returns the same dict (``vocab``, ``chars``, ``vocab_size``, ``data``) plus a
    private ``_ids`` entry holding the already-encoded tensor.
    """
    with codecs.open(input_file, "r", encoding=encoding) as f:
        data = f.read()
    chars, vocab, cp, cps_sorted = _vocab_from_text(data)
    with open(vocab_file, "wb") as f:
        pickle.dump(chars, f)
    return {"vocab": vocab, "chars": chars, "vocab_size": len(chars), "data": data,
            "_ids": encode_text(data, vocab, _cache=(cp, cps_sorted))}


class TextLoader():
    """Reference ``utils.py:19-90`` — same constructor, attributes and methods.

    Attributes used by the drivers: ``vocab_size chars vocab tensor num_batches
    x_batches y_batches pointer``; methods ``next_batch`` and
    ``reset_batch_pointer``.  ``x_batches[k]`` is the ``[B, S]`` view
    ``X[:, k*S:(k+1)*S]`` of the row-major ``[B, num_batches*S]`` matrix, so row
    ``b`` of consecutive batches is one contiguous character stream — the
    property the carried RNN state relies on (``train.py:194-199``).
    """

    def __init__(self, data_dir, batch_size, seq_length, encoding="utf-8", tensor_file=None):
        self.data_dir = data_dir
        self.batch_size = batch_size
        self.seq_length = seq_length
        self.encoding = encoding

        input_file = os.path.join(data_dir, "input.txt")
        vocab_file = os.path.join(data_dir, "vocab.pkl")
        explicit_tensor_file = tensor_file is not None
        tensor_file = tensor_file or os.path.join(data_dir, "data.npy")

        if not (os.path.exists(vocab_file) and os.path.exists(tensor_file)):
            if explicit_tensor_file and os.path.exists(tensor_file):
                # the reference would re-preprocess and overwrite the
                # caller's shard with the full corpus.  Never write to --tensor_file.
                raise Exception("vocab.pkl is missing in %s: run data_splitter.py first "
                                "(a shard is never overwritten)" % data_dir)
            print("reading text file")
            self.preprocess(input_file, vocab_file, tensor_file)
        else:
            print("loading preprocessed files")
            self.load_preprocessed(vocab_file, tensor_file)
        self.create_batches()
        self.reset_batch_pointer()

    def preprocess(self, input_file, vocab_file, tensor_file):
        """``utils.py:40-49``."""
        res = create_vocab_file(input_file, vocab_file, encoding=self.encoding)
        self.chars = res["chars"]
        self.vocab_size = res["vocab_size"]
        self.vocab = res["vocab"]
        self.tensor = res["_ids"]
        np.save(tensor_file, self.tensor)

    def load_preprocessed(self, vocab_file, tensor_file):
        """``utils.py:53-60``."""
        with open(vocab_file, "rb") as f:
            self.chars = pickle.load(f)
        self.vocab_size = len(self.chars)
        self.vocab = dict(zip(self.chars, range(len(self.chars))))
        self.tensor = np.load(tensor_file)
        self.num_batches = int(self.tensor.size / (self.batch_size * self.seq_length))

    def create_batches(self):
        """``utils.py:62-82`` — truncate, labels = inputs shifted by one (wrapping), cut."""
        B, S = self.batch_size, self.seq_length
        nb = self.num_batches = int(self.tensor.size / (B * S))   # same float division
        assert nb != 0, "Not enough data. Make seq_length and batch_size small."
        self.tensor = self.tensor[:nb * B * S]
        X = self.tensor.reshape(B, nb * S)
        Y = np.roll(self.tensor, -1).reshape(B, nb * S)           # y[-1] == x[0]
        self.x_batches = [X[:, k * S:(k + 1) * S] for k in range(nb)]
        self.y_batches = [Y[:, k * S:(k + 1) * S] for k in range(nb)]
        self._XY = (X, Y)

    def next_batch(self):
        k = self.pointer
        self.pointer = k + 1
        return self.x_batches[k], self.y_batches[k]

    def reset_batch_pointer(self):
        self.pointer = 0

    # -- additions for the device path (not in the reference) -----------------
    def batch_matrices(self):
        """``(X, Y)`` as contiguous ``[B, num_batches*S]`` int32 matrices.

        Batch ``k`` is columns ``k*S:(k+1)*S``; uploading these once lets the
        training loop slice batches on the device instead of copying per step.
        """
        X, Y = self._XY
        return (np.ascontiguousarray(X, dtype=np.int32),
                np.ascontiguousarray(Y, dtype=np.int32))


def split_tensor(tensor, num_parts):
    """``data_splitter.py:31``: equal contiguous parts; ValueError unless divisible."""
    return np.split(tensor, num_parts)
