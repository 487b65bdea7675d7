// gemm_simt.cuh — fp32 FFMA GEMM for the parity mode (CRNN_PREC_FP32) and as the on-GPU
// cross-check of the tcgen05 path.  C[M,N] (+)= op(A)[M,K] * op(B)[K,N] (+ bias[N]).
//
// Element access is by strides (A(m,k) = A[m*sAm + k*sAk]) so that all three shapes the
// training step needs share one kernel:
//   NN  x-projection  XP = In * Wx + b        (model.py:72, the x half of [x,h]*W)
//   NT  input grads   dIn = DZ * Wx^T         (tf.gradients through the same matmul)
//   TN  weight grads  dW = In^T * DZ          (K = S*B is long: split-K + atomics)
#pragma once
#include "common.cuh"
#include <algorithm>

namespace crnn {

struct GemmArgs {
    const float* A; int64_t sAm, sAk;
    const float* B; int64_t sBk, sBn;
    float* C; int64_t ldc;
    const float* bias;
    int M, N, K;
    int accumulate;   // 1: C += ...
    int ksplit_len;   // K range per blockIdx.z (multiple of BK)
};

template <int BM, int BN, int BK, int TM, int TN>
__global__ void __launch_bounds__(256) gemm_f32_kernel(GemmArgs g) {
    static_assert((BM / TM) * (BN / TN) == 256, "256 threads");
    __shared__ float As[BK][BM + 4];
    __shared__ float Bs[BK][BN + 4];
    const int tid = threadIdx.x;
    const int tx = tid % (BN / TN), ty = tid / (BN / TN);
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    const int kbeg = blockIdx.z * g.ksplit_len;
    const int kend = min(g.K, kbeg + g.ksplit_len);

    float acc[TM][TN];
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

    const bool a_kfast = (g.sAk == 1);
    const bool b_nfast = (g.sBn == 1);

    for (int k0 = kbeg; k0 < kend; k0 += BK) {
        // A tile: BM x BK -> As[k][m]
#pragma unroll
        for (int i = 0; i < (BM * BK) / 256; ++i) {
            int idx = tid + i * 256;
            int m, k;
            if (a_kfast) { k = idx % BK; m = idx / BK; } else { m = idx % BM; k = idx / BM; }
            int gm = m0 + m, gk = k0 + k;
            float v = 0.f;
            if (gm < g.M && gk < kend) v = __ldg(g.A + (int64_t)gm * g.sAm + (int64_t)gk * g.sAk);
            As[k][m] = v;
        }
#pragma unroll
        for (int i = 0; i < (BN * BK) / 256; ++i) {
            int idx = tid + i * 256;
            int n, k;
            if (b_nfast) { n = idx % BN; k = idx / BN; } else { k = idx % BK; n = idx / BK; }
            int gn = n0 + n, gk = k0 + k;
            float v = 0.f;
            if (gn < g.N && gk < kend) v = __ldg(g.B + (int64_t)gk * g.sBk + (int64_t)gn * g.sBn);
            Bs[k][n] = v;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < BK; ++k) {
            float a[TM], b[TN];
#pragma unroll
            for (int i = 0; i < TM; i += 4) {
                float4 t = *reinterpret_cast<const float4*>(&As[k][ty * TM + i]);
                a[i] = t.x; a[i + 1] = t.y; a[i + 2] = t.z; a[i + 3] = t.w;
            }
#pragma unroll
            for (int j = 0; j < TN; j += 4) {
                float4 t = *reinterpret_cast<const float4*>(&Bs[k][tx * TN + j]);
                b[j] = t.x; b[j + 1] = t.y; b[j + 2] = t.z; b[j + 3] = t.w;
            }
#pragma unroll
            for (int i = 0; i < TM; ++i)
#pragma unroll
                for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        __syncthreads();
    }

    const bool split = gridDim.z > 1;
#pragma unroll
    for (int i = 0; i < TM; ++i) {
        int gm = m0 + ty * TM + i;
        if (gm >= g.M) continue;
#pragma unroll
        for (int j = 0; j < TN; ++j) {
            int gn = n0 + tx * TN + j;
            if (gn >= g.N) continue;
            float v = acc[i][j];
            if (g.bias && blockIdx.z == 0) v += __ldg(g.bias + gn);
            float* c = g.C + (int64_t)gm * g.ldc + gn;
            if (split) atomicAdd(c, v);
            else if (g.accumulate) *c += v;
            else *c = v;
        }
    }
}

// Host launcher.  With split-K the partial products are atomically added, so C must hold
// the value to accumulate onto (the gradient buffer is zeroed at the start of a step).
inline int gemm_f32(int ta, int tb, int M, int N, int K, const float* A, int64_t lda,
                    const float* B, int64_t ldb, float* C, int64_t ldc, const float* bias,
                    int accumulate, cudaStream_t st, int sm_count = 148) {
    if (M <= 0 || N <= 0) return CRNN_OK;
    GemmArgs g;
    g.A = A; g.sAm = ta ? 1 : lda; g.sAk = ta ? lda : 1;
    g.B = B; g.sBk = tb ? 1 : ldb; g.sBn = tb ? ldb : 1;
    g.C = C; g.ldc = ldc; g.bias = bias; g.M = M; g.N = N; g.K = K; g.accumulate = accumulate;
    const bool big = (int64_t)M * N >= (int64_t)128 * 128 * sm_count;
    const int BM = big ? 128 : 64, BN = big ? 128 : 64, BK = 16;
    dim3 grid(ceil_div(N, BN), ceil_div(M, BM), 1);
    int splits = 1;
    int64_t tiles = (int64_t)grid.x * grid.y;
    if (accumulate && tiles < 2 * sm_count && K >= 512) {
        splits = (int)std::min<int64_t>(ceil_div64(2 * sm_count, tiles), K / 256);
        if (splits < 1) splits = 1;
    }
    g.ksplit_len = (int)round_up64(ceil_div(K, splits), BK);
    grid.z = ceil_div(K, g.ksplit_len);
    if (K == 0) { grid.z = 1; g.ksplit_len = BK; }
    if (big) gemm_f32_kernel<128, 128, 16, 8, 8><<<grid, 256, 0, st>>>(g);
    else     gemm_f32_kernel<64, 64, 16, 4, 4><<<grid, 256, 0, st>>>(g);
    CRNN_LAUNCH_CHECK("gemm_f32_kernel");
    return CRNN_OK;
}

// Column sums: out[c] += sum_r X[r*ld + c]   (bias gradients: db = sum over (t,b) of dz).
// `part` == nullptr: row blocks are combined with atomics (order not fixed).  With a scratch array of
// gridDim.y * C floats the row blocks' sums are stored and added in block order by colsum_finish_kernel.
__global__ void __launch_bounds__(256) colsum_kernel(const float* __restrict__ X, int64_t ld, int R, int C,
                                                     float* __restrict__ out, int rows_per_block, float* __restrict__ part) {
    __shared__ float sh[8][33];
    int c = blockIdx.x * 32 + (threadIdx.x & 31);
    int ry = threadIdx.x >> 5;
    int r0 = blockIdx.y * rows_per_block;
    int r1 = min(R, r0 + rows_per_block);
    float s = 0.f;
    if (c < C)
        for (int r = r0 + ry; r < r1; r += 8) s += __ldg(X + (int64_t)r * ld + c);
    sh[ry][threadIdx.x & 31] = s;
    __syncthreads();
    if (ry == 0 && c < C) {
        float t = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) t += sh[i][threadIdx.x & 31];
        if (part) part[(int64_t)blockIdx.y * C + c] = t;
        else atomicAdd(out + c, t);
    }
}
__global__ void __launch_bounds__(256) colsum_finish_kernel(const float* __restrict__ part, int nblk, int C, float* __restrict__ out) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    float t = 0.f;
    for (int i = 0; i < nblk; ++i) t += part[(int64_t)i * C + c];
    out[c] += t;
}

inline int64_t colsum_scratch_floats(int64_t R, int64_t C) { return ceil_div64(R, 1024) * C; }
inline int colsum_f32(const float* X, int64_t ld, int R, int C, float* out, cudaStream_t st, float* part = nullptr) {
    if (R <= 0 || C <= 0) return CRNN_OK;
    int rows_per_block = 1024;
    dim3 grid(ceil_div(C, 32), ceil_div(R, rows_per_block));
    colsum_kernel<<<grid, 256, 0, st>>>(X, ld, R, C, out, rows_per_block, grid.y > 1 ? part : nullptr);
    CRNN_LAUNCH_CHECK("colsum_kernel");
    if (part && grid.y > 1) {
        colsum_finish_kernel<<<ceil_div(C, 256), 256, 0, st>>>(part, (int)grid.y, C, out);
        CRNN_LAUNCH_CHECK("colsum_finish_kernel");
    }
    return CRNN_OK;
}

}  // namespace crnn
