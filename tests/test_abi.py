"""The C-ABI library loads and exports exactly what include/crnn.h declares (no compute here)."""
import ctypes
import os
import re

import pytest

from conftest import ROOT
from distributed_char_rnn_b200 import _lib


def _declared():
    src = open(os.path.join(ROOT, "include", "crnn.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(crnn_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported(lib):
    names = _declared()
    assert "crnn_train_step" in names and "crnn_forward_backward" in names and len(names) >= 15
    for n in names:
        assert hasattr(lib, n), "libcrnn.so does not export %s" % n
    # and the ctypes prototypes cover the whole header
    assert sorted(_lib.SIGNATURES) == names


def test_no_torch_types_in_header():
    src = open(os.path.join(ROOT, "include", "crnn.h")).read()
    assert "torch" not in re.sub(r"/\*.*?\*/", "", src, flags=re.S) and "at::" not in src


def test_version_and_error_string(lib):
    assert b"sm_100a" in lib.crnn_version()
    assert isinstance(lib.crnn_last_error(), bytes)


def test_create_rejects_bad_config_and_has_no_cpu_path(lib):
    import torch
    h = ctypes.c_void_p()
    bad = _lib.CrnnConfig(7, 8, 1, 4, 2, 5, 0, 0)               # unknown model id, like model.py:25
    assert lib.crnn_create(ctypes.byref(bad), ctypes.byref(h)) != 0
    assert b"model type not supported" in lib.crnn_last_error()
    bad = _lib.CrnnConfig(2, 0, 1, 4, 2, 5, 0, 0)
    assert lib.crnn_create(ctypes.byref(bad), ctypes.byref(h)) != 0
    if not torch.cuda.is_available():
        ok = _lib.CrnnConfig(2, 8, 1, 4, 2, 5, 0, 0)
        assert lib.crnn_create(ctypes.byref(ok), ctypes.byref(h)) != 0
        assert b"no CPU path" in lib.crnn_last_error()
        from distributed_char_rnn_b200.model import Model
        import argparse
        with pytest.raises(_lib.CrnnError, match="no CPU path"):
            Model(argparse.Namespace(model="lstm", rnn_size=8, num_layers=1, seq_length=4, batch_size=2,
                                     vocab_size=5, grad_clip=5.0))


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(_lib.CrnnError, match="no CPU fallback"):
        _lib.load()


def test_python_constants_match_the_header():
    """Precision ids and return codes of the ctypes layer are the header's."""
    import re
    from distributed_char_rnn_b200 import _lib
    hdr = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "crnn.h")).read()
    defs = {m.group(1): int(m.group(2)) for m in re.finditer(r"#define\s+(CRNN_[A-Z0-9_]+)\s+(-?\d+)", hdr)}
    assert _lib.PREC_IDS == {"fp32": defs["CRNN_PREC_FP32"], "bf16": defs["CRNN_PREC_BF16"], "bf16x3": defs["CRNN_PREC_BF16X3"]}
    assert (_lib.CRNN_ERR_ARG, _lib.CRNN_ERR_CUDA, _lib.CRNN_ERR_UNSUP) == (defs["CRNN_ERR_ARG"], defs["CRNN_ERR_CUDA"], defs["CRNN_ERR_UNSUP"])
    assert _lib.MODEL_IDS == {"rnn": defs["CRNN_MODEL_RNN"], "gru": defs["CRNN_MODEL_GRU"], "lstm": defs["CRNN_MODEL_LSTM"]}
    assert _lib.GRAD_TAIL == defs["CRNN_GRAD_TAIL"]
