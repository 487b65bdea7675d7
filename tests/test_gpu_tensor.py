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
