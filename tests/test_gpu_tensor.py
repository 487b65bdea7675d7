"""tcgen05 GEMM (CRNN_PREC_BF16 building block) against fp64 products of the same bf16 inputs."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

CASES = [
    # a_mn, b_mn, M, N, K, bias, accumulate
    (0, 0, 128, 128, 64, False, 0),
    (0, 0, 128, 128, 256, True, 0),
    (0, 1, 256, 128, 128, True, 0),       # NN: x-projection shape (B stored [K,N])
    (1, 1, 128, 256, 192, False, 0),      # TN: weight-gradient shape (A stored [K,M])
    (1, 0, 128, 128, 128, False, 0),
    (0, 0, 200, 96, 520, True, 0),        # ragged M, N, K (TMA zero fill)
    (0, 1, 2560, 1536, 512, True, 0),     # config-3 x-projection, a tenth of the rows
    (0, 0, 2560, 512, 1536, False, 0),    # config-3 input gradients
    (1, 1, 512, 1536, 5120, False, 1),    # config-3 weight gradients: split-K + atomics
    (1, 1, 72, 40, 1000, False, 1),
    (0, 1, 100, 96, 512, True, 0),        # vocab projection
    (0, 0, 1, 8, 8, True, 0),
]


@pytest.mark.parametrize("a_mn,b_mn,M,N,K,bias,acc", CASES)
def test_gemm_bf16_tc(lib, a_mn, b_mn, M, N, K, bias, acc):
    from distributed_char_rnn_b200._lib import check, ptr
    rng = np.random.default_rng(M + 3 * N + 7 * K)
    pad = lambda n: (n + 7) // 8 * 8
    A = rng.normal(size=(M, K)).astype(np.float32)
    Bm = rng.normal(size=(K, N)).astype(np.float32)
    At = torch.from_numpy(A).cuda().to(torch.bfloat16)
    Bt = torch.from_numpy(Bm).cuda().to(torch.bfloat16)
    ref = At.double().cpu().numpy() @ Bt.double().cpu().numpy()
    # storage orientation + padded leading dimension
    if a_mn:
        As = torch.zeros((K, pad(M)), dtype=torch.bfloat16, device="cuda"); As[:, :M] = At.t()
    else:
        As = torch.zeros((M, pad(K)), dtype=torch.bfloat16, device="cuda"); As[:, :K] = At
    if b_mn:
        Bs = torch.zeros((K, pad(N)), dtype=torch.bfloat16, device="cuda"); Bs[:, :N] = Bt
    else:
        Bs = torch.zeros((N, pad(K)), dtype=torch.bfloat16, device="cuda"); Bs[:, :K] = Bt.t()
    C0 = rng.normal(size=(M, N)).astype(np.float32)
    C = torch.from_numpy(C0.copy()).cuda()
    b = rng.normal(size=N).astype(np.float32) if bias else None
    db = torch.from_numpy(b).cuda() if bias else None
    check(lib.crnn_gemm_bf16(a_mn, b_mn, M, N, K, ptr(As), As.stride(0), ptr(Bs), Bs.stride(0), ptr(C), N, ptr(db), acc, None))
    torch.cuda.synchronize()
    if bias:
        ref = ref + b
    if acc:
        ref = ref + C0
    err = np.abs(C.cpu().numpy() - ref).max()
    assert err <= 1e-4 * max(np.abs(ref).max(), 1.0), err


# ---------------------------------------------------------------------------------------
# CRNN_PREC_BF16: the whole step on tensor cores.  Comparator = the oracle evaluated with
# every matmul's inputs rounded to bf16 (fp64 accumulation): same rounding points as the
# kernels (bf16 operands, fp32 accumulate, fp32 state / gates / master weights).
# Tolerance 1e-2 relative per tensor (stated): a value that lands on the other side of a
# bf16 rounding boundary moves by 2^-8 = 3.9e-3 relative, and that propagates through S steps.
# ---------------------------------------------------------------------------------------
import argparse
import os

from oracle import char_rnn_oracle as O

BF16_RTOL = 1e-2


def _relerr(a, b, floor):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return np.abs(a - b).max() / max(np.abs(b).max(), floor)


def _bf16_model(model, H, L, S, B, V, seed=0, env=None):
    from distributed_char_rnn_b200.model import Model
    old = {}
    for k, v in (env or {}).items():
        old[k] = os.environ.get(k)
        os.environ[k] = str(v)
    try:
        cfg = O.Config(model, H, L, S, B, V, 5.0)
        params = O.init_params(cfg, seed)
        rng = np.random.default_rng(seed + 100)
        for k in params:
            if k.endswith("bias"):
                params[k] = (params[k] + rng.normal(size=params[k].shape) * 0.1).astype(np.float32)
        args = argparse.Namespace(model=model, rnn_size=H, num_layers=L, seq_length=S, batch_size=B, vocab_size=V,
                                  grad_clip=5.0, output_keep_prob=1.0, input_keep_prob=1.0)
        m = Model(args, device="cuda:0", precision="bf16")
        m.set_variables(params)
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    return cfg, params, m


BF16_SHAPES = [
    ("lstm", 32, 2, 6, 5, 11, {}), ("gru", 32, 2, 6, 5, 11, {}), ("rnn", 32, 2, 6, 5, 11, {}),
    ("lstm", 128, 2, 10, 50, 65, {}), ("gru", 64, 2, 8, 70, 96, {}), ("rnn", 48, 3, 7, 33, 40, {}),
    ("lstm", 64, 1, 5, 150, 20, {"CRNN_TC_TILES": 1}),      # BT = 150 -> two 128-row passes
    ("gru", 128, 1, 4, 260, 20, {"CRNN_TC_TILES": 2}),      # BT = 130 -> two passes, ragged
    ("rnn", 96, 2, 5, 40, 30, {"CRNN_TC_HC": 16}),
    # both MMA orientations forced (weights on M / batch rows on M)
    ("lstm", 32, 2, 6, 20, 11, {"CRNN_TC_SWAP": 1}), ("gru", 64, 2, 5, 40, 17, {"CRNN_TC_SWAP": 1}),
    ("rnn", 128, 1, 5, 24, 17, {"CRNN_TC_SWAP": 1}), ("gru", 64, 2, 5, 40, 17, {"CRNN_TC_SWAP": 0}),
    ("lstm", 32, 1, 4, 150, 9, {"CRNN_TC_SWAP": 1, "CRNN_TC_TILES": 1}),
    # layer wavefront: layers on concurrent streams, the sequence cut into chunks (uneven ones too)
    ("lstm", 32, 3, 7, 9, 11, {"CRNN_WAVEFRONT": 2, "CRNN_WAVEFRONT_CHUNKS": 3}),
    ("gru", 64, 2, 8, 20, 17, {"CRNN_WAVEFRONT": 2, "CRNN_WAVEFRONT_CHUNKS": 4}),
    ("rnn", 48, 3, 6, 12, 13, {"CRNN_WAVEFRONT": 3, "CRNN_WAVEFRONT_CHUNKS": 2}),
    ("gru", 64, 2, 8, 20, 17, {"CRNN_WAVEFRONT": 1}),
    ("lstm", 64, 2, 5, 8, 9, {"CRNN_WAVEFRONT": 2, "CRNN_WAVEFRONT_CHUNKS": 5}),
    # the same wavefront with layer 0's batch GEMMs hoisted out of it
    ("gru", 64, 2, 8, 20, 17, {"CRNN_WAVEFRONT": 2, "CRNN_WAVEFRONT_CHUNKS": 4}),
    ("lstm", 32, 3, 7, 9, 11, {"CRNN_WAVEFRONT": 2, "CRNN_WAVEFRONT_CHUNKS": 3}),
    ("gru", 64, 2, 8, 20, 17, {"CRNN_WAVEFRONT": 2, "CRNN_WAVEFRONT_CHUNKS": 4, "CRNN_WF_HOIST0": 1}),
    ("rnn", 48, 3, 6, 12, 13, {"CRNN_WAVEFRONT": 2, "CRNN_WAVEFRONT_CHUNKS": 3, "CRNN_WF_HOIST0": 1}),
    ("lstm", 128, 2, 6, 40, 13, {"CRNN_TC_HC": 16}), ("rnn", 96, 2, 6, 70, 13, {"CRNN_TC_HC": 16}),
    # direct exchange of the streamed rows (opt-in)
    ("lstm", 128, 2, 6, 40, 13, {"CRNN_TC_GATHER": 1}), ("gru", 64, 2, 8, 20, 17, {"CRNN_TC_GATHER": 1, "CRNN_WAVEFRONT_CHUNKS": 4}),
    ("rnn", 96, 2, 6, 70, 13, {"CRNN_TC_GATHER": 1}),
    ("gru", 128, 1, 5, 300, 13, {"CRNN_TC_HC": 16, "CRNN_TC_TILES": 2}),
    # K split over a cluster of 2 / 4 CTAs, partial sums exchanged through distributed shared memory
    ("lstm", 128, 2, 6, 40, 13, {"CRNN_TC_HC": 16, "CRNN_TC_KSPLIT": 2}),
    ("gru", 256, 1, 5, 20, 13, {"CRNN_TC_HC": 16, "CRNN_TC_KSPLIT": 4}),
    ("rnn", 256, 2, 5, 70, 13, {"CRNN_TC_HC": 32, "CRNN_TC_KSPLIT": 2}),
    ("lstm", 256, 1, 4, 150, 9, {"CRNN_TC_HC": 16, "CRNN_TC_KSPLIT": 4, "CRNN_TC_TILES": 1}),
    ("gru", 128, 2, 8, 24, 13, {"CRNN_TC_HC": 16, "CRNN_TC_KSPLIT": 2, "CRNN_WAVEFRONT_CHUNKS": 3}),
    # cluster step barrier (remote mbarrier arrives) at 16 slices, and switched off
    ("gru", 256, 1, 5, 20, 13, {"CRNN_TC_HC": 16, "CRNN_TC_CBAR": 16}), ("lstm", 128, 2, 6, 40, 13, {"CRNN_TC_HC": 16, "CRNN_TC_CBAR": 0}),
    # cluster exchange (default where all slices of a tile fit one cluster) across chunk boundaries, and the cluster step
    # barrier that it replaced (CRNN_TC_CX=0)
    ("lstm", 128, 2, 12, 50, 65, {"CRNN_WAVEFRONT_CHUNKS": 3}), ("gru", 128, 2, 9, 37, 30, {"CRNN_WAVEFRONT_CHUNKS": 3}),
    ("rnn", 64, 2, 9, 21, 13, {"CRNN_WAVEFRONT_CHUNKS": 2}), ("gru", 256, 1, 6, 90, 13, {}),
    ("lstm", 128, 2, 12, 50, 65, {"CRNN_TC_CX": 0}), ("gru", 64, 2, 8, 20, 17, {"CRNN_TC_CX": 0, "CRNN_WAVEFRONT_CHUNKS": 4}),
    ("lstm", 64, 2, 6, 20, 13, {"CRNN_TC_HC": 8}),          # LSTM slices of 8 units
    # one launch per layer and direction forced on (config 3's layers) / off (small layers)
    ("gru", 512, 2, 12, 64, 96, {"CRNN_WF_PERSIST": 1}), ("lstm", 128, 2, 12, 50, 65, {"CRNN_WF_PERSIST": 0, "CRNN_WAVEFRONT_CHUNKS": 3}),
]


@pytest.mark.parametrize("fast", [0, 1])
@pytest.mark.parametrize("model,H,L,S,B,V,env", BF16_SHAPES)
def test_bf16_gradients_match_bf16_oracle(model, H, L, S, B, V, env, fast):
    import torch
    cfg, params, m = _bf16_model(model, H, L, S, B, V, seed=H + L, env=dict(env, CRNN_FAST_ACT=fast))
    rng = np.random.default_rng(S * B)
    x, y = rng.integers(0, V, (B, S)), rng.integers(0, V, (B, S))
    st = [tuple((rng.normal(size=s.shape) * 0.5).astype(np.float32) for s in layer) for layer in O.zero_state(cfg)]
    fb = O.forward_backward(cfg, params, x, y, st, np.float64, bf16_gemm=True)
    sflat = torch.from_numpy(np.concatenate([np.stack(layer).reshape(-1) for layer in st]).astype(np.float32)).cuda()
    loss, sout = m.gradients(x.astype(np.int32), y.astype(np.int32), sflat)
    assert abs(loss - fb["loss"]) <= 2e-3 * abs(fb["loss"])
    got = m.gradients_numpy()
    gmax = max(np.abs(g).max() for g in fb["grads"].values())
    for k, g in fb["grads"].items():
        assert _relerr(got[k], g, 1e-2 * gmax) < BF16_RTOL, k
    ref_state = np.concatenate([np.stack(layer).reshape(-1) for layer in fb["final_state"]])
    assert _relerr(sout.cpu().numpy(), ref_state, 1e-2) < BF16_RTOL
    # and the bf16 path stays close to the exact fp32 graph (the reference's arithmetic)
    fb32 = O.forward_backward(cfg, params, x, y, st, np.float64)
    assert abs(loss - fb32["loss"]) <= 5e-3 * abs(fb32["loss"])


@pytest.mark.parametrize("model", ["lstm", "gru", "rnn"])
def test_bf16_training_tracks_fp32_oracle(model):
    """Loss curve of the bf16 tensor-core path against the fp32 oracle over 20 steps with state carry."""
    H, L, S, B, V = 64, 2, 12, 24, 30
    cfg, params, m = _bf16_model(model, H, L, S, B, V, seed=7)
    tr = O.Trainer(cfg, params, np.float32)
    tr.begin_epoch(0)
    m.assign_lr(0.002)
    rng = np.random.default_rng(2)
    # a learnable stream: next id = (id + 1) % V with noise
    base = rng.integers(0, V, (B, 1))
    for k in range(20):
        x = (base + np.arange(S)[None, :] + k * S) % V
        y = (x + 1) % V
        loss, _ = m.step(x.astype(np.int32), y.astype(np.int32))
        ref = float(tr.step(x, y))
        assert abs(loss - ref) <= 2e-2 * ref, (k, loss, ref)
    assert loss < 0.98 * np.log(V)            # it is learning
