"""End-to-end through the reference's CLI surface on the GPU: train -> checkpoint -> resume -> sample
(the three smoke runs of the reference's .travis.yml:32-34), plus the replica protocol on CUDA."""
import argparse
import ctypes
import os
import pickle
import subprocess
import sys

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from conftest import ROOT                                           # noqa: E402
from distributed_char_rnn_b200 import parallel as PL                # noqa: E402
from oracle import char_rnn_oracle as O                             # noqa: E402


def _corpus(d):
    rng = np.random.default_rng(0)
    words = ["the", "quick", "brown", "fox", "jumps", "over", "lazy", "dog", "\n"]
    text = " ".join(words[i] for i in rng.integers(0, len(words), 3000))
    os.makedirs(d, exist_ok=True)
    with open(os.path.join(d, "input.txt"), "w") as f:
        f.write(text)


def _run(script, *argv):
    env = dict(os.environ, PYTHONPATH=ROOT)
    return subprocess.run([sys.executable, os.path.join(ROOT, script)] + list(argv), capture_output=True, text=True,
                          cwd=ROOT, env=env, timeout=300)


@pytest.mark.parametrize("model,precision", [("lstm", "fp32"), ("gru", "auto"), ("rnn", "bf16"), ("lstm", "bf16x3")])
def test_train_resume_sample(tmp_path, model, precision):
    data, s1, s2, logs = (str(tmp_path / n) for n in ("data", "s1", "s2", "logs"))
    _corpus(data)
    common = ["--data_dir", data, "--log_dir", logs, "--model", model, "--rnn_size", "32", "--num_layers", "2",
              "--seq_length", "20", "--batch_size", "10", "--num_epochs", "2", "--precision", precision, "--seed", "1"]
    r = _run("train.py", "--save_dir", s1, *common)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if "train_loss" in l]
    nb = int(lines[0].split("/")[1].split()[0]) // 2
    assert lines[0].startswith("1/%d (epoch 0), train_loss = " % (2 * nb)) and "time/batch" in lines[0]
    assert lines[-1].startswith("%d/%d (epoch 1)" % (2 * nb, 2 * nb))
    first, last = float(lines[0].split("train_loss = ")[1].split(",")[0]), float(lines[-1].split("train_loss = ")[1].split(",")[0])
    assert last < first
    assert os.path.getsize(os.path.join(s1, "model.ckpt-%d.index" % (2 * nb - 1))) > 0      # .travis.yml:32
    cfgp = pickle.load(open(os.path.join(s1, "config.pkl"), "rb"))
    assert cfgp.vocab_size > 0 and cfgp.model == model                                         # train.py:112-113
    chars, vocab = pickle.load(open(os.path.join(s1, "chars_vocab.pkl"), "rb"))
    assert vocab == {c: i for i, c in enumerate(chars)}
    r2 = _run("train.py", "--save_dir", s2, "--init_from", s1, *common)                        # .travis.yml:33
    assert r2.returncode == 0, r2.stderr[-2000:]
    l2 = [l for l in r2.stdout.splitlines() if "train_loss" in l]
    assert float(l2[0].split("train_loss = ")[1].split(",")[0]) < first                        # it resumed, not restarted
    assert l2[0].startswith("%d/" % (2 * nb + 1))                                              # global_step restored
    r3 = _run("sample.py", "--save_dir", s2, "-n", "40", "--prime", "the ")                    # .travis.yml:34
    assert r3.returncode == 0, r3.stderr[-2000:]
    out = r3.stdout.strip().splitlines()[-1]
    assert out.startswith("b'the ") and len(eval(out).decode("utf-8")) == 44
    bad = _run("train.py", "--save_dir", s2, "--init_from", s1, *(common[:-8] + ["--seq_length", "21"] + common[-6:]))
    assert bad.returncode != 0 and "disagree on 'seq_length'" in bad.stderr


def test_layout_matches_host_mirror():
    from distributed_char_rnn_b200.model import Model
    for model, H, L, V in [("lstm", 128, 2, 65), ("gru", 48, 3, 96), ("rnn", 20, 1, 7)]:
        ns = argparse.Namespace(model=model, rnn_size=H, num_layers=L, seq_length=3, batch_size=2, vocab_size=V, grad_clip=5.0)
        m = Model(ns, device="cuda:0")
        lay, P = PL.flat_layout(model, H, L, V)
        assert P == m.P
        assert [(n, o) for n, o, _ in lay] == [(n, o) for n, o, _, _ in m.layout]


@pytest.mark.parametrize("model,precision", [("lstm", "fp32"), ("gru", "fp32"), ("rnn", "fp32"), ("gru", "bf16"), ("lstm", "bf16x3")])
def test_two_replicas_on_cuda_equal_stacked_batch(model, precision):
    """Rank A and rank B run crnn_forward_backward on their own shard batch; their flat buffers are
    summed (what the NCCL all-reduce does) and both apply crnn_clip_adam with grad_mult = 1/2.
    Result == one process on the stacked [2B, S] batch (oracle), and both replicas stay identical."""
    from distributed_char_rnn_b200.model import Model
    from distributed_char_rnn_b200._lib import check, ptr
    H, L, S, B, V = 32, 2, 6, 4, 9
    cfg1 = O.Config(model, H, L, S, B, V, 0.05)
    cfg2 = O.Config(model, H, L, S, 2 * B, V, 0.05)
    params = O.init_params(cfg1, 3)
    ns = lambda: argparse.Namespace(model=model, rnn_size=H, num_layers=L, seq_length=S, batch_size=B, vocab_size=V, grad_clip=0.05)
    reps = [Model(ns(), device="cuda:0", precision=precision) for _ in range(2)]
    rng = np.random.default_rng(0)
    xs = [rng.integers(0, V, (B, S)).astype(np.int32) for _ in range(2)]
    ys = [rng.integers(0, V, (B, S)).astype(np.int32) for _ in range(2)]
    for m in reps:
        m.set_variables(params)
    for m, x, y in zip(reps, xs, ys):
        m.gradients(x, y, m.zero_state())
    total = reps[0].grads + reps[1].grads
    for m in reps:
        m.grads.copy_(total)
        check(m._lib.crnn_clip_adam(m._h, ptr(m.params), ptr(m.grads), ptr(m.adam_m), ptr(m.adam_v), 0.002, 0.05, 1, 0.5,
                                    ptr(m._gnorm), ptr(m.workspace), None))
    torch.cuda.synchronize()
    assert torch.equal(reps[0].params, reps[1].params)
    tr = O.Trainer(cfg2, params, np.float32, bf16_gemm=(precision == "bf16"))
    tr.begin_epoch(0)
    ref_loss = float(tr.step(np.concatenate(xs), np.concatenate(ys)))
    tol = 1e-2 if precision == "bf16" else (1e-4 if precision == "fp32" else 5e-4)   # bf16x3: parameters after Adam, see test_gpu_bf16x3
    assert abs(float(total[reps[0].P + 1]) / 2 - ref_loss) <= tol * ref_loss
    assert abs(float(reps[0]._gnorm) - tr.last["gnorm"]) <= tol * tr.last["gnorm"] and tr.last["gnorm"] > 0.05
    got = reps[0].variables_numpy()
    for k, v in tr.params.items():
        assert np.abs(got[k] - v).max() <= tol * max(np.abs(v).max(), 1e-2), k


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs (gpurun --gpus 2)")
def test_nccl_two_gpu_training_matches_single_process(tmp_path):
    """Config 2 in miniature: 2 shards, 2 workers launched with the reference's own flags."""
    data, sh = str(tmp_path / "data"), str(tmp_path / "sh")
    _corpus(data)
    n = len(open(os.path.join(data, "input.txt")).read())
    if n % 2:
        open(os.path.join(data, "input.txt"), "a").write(" ")
    assert _run("data_splitter.py", "--data_dir", data, "--num_parts", "2", "--out_dir", sh).returncode == 0
    port = 29000 + os.getpid() % 1000
    procs = []
    for i in range(2):
        env = dict(os.environ, PYTHONPATH=ROOT, CUDA_VISIBLE_DEVICES=str(i))
        procs.append(subprocess.Popen([sys.executable, os.path.join(ROOT, "train.py"), "--distributed", "--job_name", "worker",
                                       "--task_index", str(i), "--ps_hosts", "127.0.0.1:1", "--worker_hosts",
                                       "127.0.0.1:%d,127.0.0.1:%d" % (port, port + 1), "--tensor_file",
                                       os.path.join(sh, "data-%d.npy" % i), "--data_dir", data, "--save_dir", str(tmp_path / "s"),
                                       "--log_dir", str(tmp_path / "l"), "--rnn_size", "32", "--seq_length", "10", "--batch_size", "8",
                                       "--num_epochs", "1", "--seed", "5", "--max_steps", "6"],
                                      stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, cwd=ROOT, env=env))
    outs = [p.communicate(timeout=300) for p in procs]
    assert all(p.returncode == 0 for p in procs), outs[0][1][-1500:] + outs[1][1][-1500:]
    l0 = [l.split("train_loss = ")[1].split(",")[0] for l in outs[0][0].splitlines() if "train_loss" in l]
    l1 = [l.split("train_loss = ")[1].split(",")[0] for l in outs[1][0].splitlines() if "train_loss" in l]
    assert l0 == l1 and len(l0) == 6                     # every rank reports the same all-reduced cost
