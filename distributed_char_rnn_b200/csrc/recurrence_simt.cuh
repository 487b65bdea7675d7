// recurrence_simt.cuh — persistent fp32 recurrence / BPTT kernels (CRNN_PREC_FP32).
//
// Replaces the S x L unrolled cell steps that legacy_seq2seq.rnn_decoder builds
// (reference model.py:72) and their reverse-mode mirror (model.py:88-91).
//
// One cooperative launch per layer and direction.  The x half of [x,h]*W is hoisted out
// as one batch GEMM (same math as the reference's concat+matmul, only the fp32 summation
// order changes), so only h_{t-1}*W_h is on the serial path.  CTA (slice, tile) owns
// hidden units [j0, j0+HC) for batch rows [b0, b0+BT): it loads its slice of W_h into
// shared memory ONCE and keeps it there for all S steps; per step it streams the
// previous step's rows from L2, runs the small contraction on FFMA, applies bias /
// sigmoid / tanh / cell update from shared memory, publishes h_t and joins a grid
// barrier (steps are sequential in t; slices of one step are independent).
#pragma once
#include "common.cuh"

namespace crnn {

constexpr int REC_THREADS = 256;
constexpr int REC_RB = 64;      // batch rows per pass
constexpr int REC_KC = 64;      // K chunk staged in shared memory
constexpr int REC_LDA = REC_KC + 4;

struct RecArgs {
    int B, H, S;
    int HC, BT, nslices, ntiles;
    int TXa, TXb;           // thread columns for contractions with G*HC / HC output columns
    int GS;                 // gate-stash row width in units of H
    const float* W;         // lstm/rnn kernel [2H, G*H]; gru gates/kernel [2H, 2H]
    const float* Wc;        // gru candidate/kernel [2H, H]
    float* gates;           // [S,B,GS*H]: in = x-projection (+bias); out = activations
    float* hx;              // [(S+1),B,H]  slot 0 = initial h, slot t+1 = h_t
    float* cx;              // lstm: [(S+1),B,H] cell states
    const float* dy;        // bwd: [S,B,H] gradient wrt this layer's outputs
    float* dz;              // bwd: [S,B,G*H] pre-activation gradients
    float* carry0;          // bwd scratch [B,H] (lstm: dc; gru: dh elementwise carry)
    float* carry1;          // bwd scratch [B,H] (gru: dh_t)
    unsigned* barrier;      // zeroed before launch
};

constexpr int REC_RM_MAX = 16;   // rows per thread (REC_RB / (REC_THREADS / 64))

__device__ __forceinline__ void cp_async16(float* smem_dst, const float* gsrc, int src_bytes) {
    unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(gsrc), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// Stage A[nrows][kc .. kc+kl) into one As buffer ([rows_pad][REC_LDA], zero padded).
// vec: rows are 16-byte aligned -> cp.async (L2 -> shared, no register staging, all in flight).
__device__ __forceinline__ void stage_a(float* __restrict__ As, const float* __restrict__ A, int64_t lda,
                                        int nrows, int rows_pad, int kc, int kl, bool vec) {
    const int tid = threadIdx.x;
    if (vec) {
        const int n4 = (kl + 3) >> 2;
        for (int idx = tid; idx < rows_pad * (REC_KC / 4); idx += REC_THREADS) {
            int r = idx / (REC_KC / 4), k4 = idx % (REC_KC / 4);
            if (k4 < n4) {
                const bool ok = r < nrows;
                cp_async16(As + r * REC_LDA + k4 * 4, ok ? A + (int64_t)r * lda + kc + k4 * 4 : A, ok ? 16 : 0);
            }
        }
    } else {
        const int kl4 = (kl + 3) & ~3;
        for (int idx = tid; idx < rows_pad * REC_KC; idx += REC_THREADS) {
            int r = idx / REC_KC, k = idx % REC_KC;
            if (k < kl4) {
                float v = 0.f;
                if (r < nrows && k < kl) v = __ldcg(A + (int64_t)r * lda + kc + k);
                As[r * REC_LDA + k] = v;
            }
        }
    }
}

// Z_s[r][lc] = sum_k A[r*lda + k] * W_s[k*ldw + wcol0 + lc]   for r < nrows (<= REC_RB), lc < NC.
// A is global (written by other CTAs in the previous phase -> read through L2, never L1).
// TX (power of two, 4..64) threads span the columns, 256/TX thread rows span the batch rows,
// RM rows per thread (compile-time so that no predicated-off FFMA is issued);
// the K loop is double-buffered with cp.async so the next chunk lands during the FFMAs.
template <int RM>
__device__ __forceinline__ void cta_gemm_rm(int TX, float* __restrict__ Zs, int ldz, const float* __restrict__ A,
                                            int64_t lda, int nrows, int K, const float* __restrict__ Ws,
                                            int ldw, int wcol0, int NC, float* __restrict__ As, bool vec) {
    const int TY = REC_THREADS / TX;
    const int tid = threadIdx.x;
    const int lc = tid % TX, ty = tid / TX;
    const int rows_pad = RM * TY;
    float acc[RM];
#pragma unroll
    for (int i = 0; i < RM; ++i) acc[i] = 0.f;
    const int nchunk = ceil_div(K, REC_KC);
    stage_a(As, A, lda, nrows, rows_pad, 0, min(REC_KC, K), vec);
    cp_async_commit();
    for (int c = 0; c < nchunk; ++c) {
        const int kc = c * REC_KC;
        const int kl = min(REC_KC, K - kc);
        const int kl4 = (kl + 3) & ~3;
        float* cur = As + (c & 1) * (REC_RB * REC_LDA);
        if (c + 1 < nchunk) {
            stage_a(As + ((c + 1) & 1) * (REC_RB * REC_LDA), A, lda, nrows, rows_pad, kc + REC_KC,
                    min(REC_KC, K - kc - REC_KC), vec);
            cp_async_commit();
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads();
        if (lc < NC) {
            const float* w = Ws + (int64_t)kc * ldw + wcol0 + lc;
            const float* arow = cur + ty * REC_LDA;
#pragma unroll 2
            for (int k = 0; k < kl4; k += 4) {
                float w0 = w[(k + 0) * ldw], w1 = w[(k + 1) * ldw], w2 = w[(k + 2) * ldw], w3 = w[(k + 3) * ldw];
#pragma unroll
                for (int i = 0; i < RM; ++i) {
                    float4 a = *reinterpret_cast<const float4*>(&arow[i * TY * REC_LDA + k]);
                    acc[i] = fmaf(a.x, w0, acc[i]);
                    acc[i] = fmaf(a.y, w1, acc[i]);
                    acc[i] = fmaf(a.z, w2, acc[i]);
                    acc[i] = fmaf(a.w, w3, acc[i]);
                }
            }
        }
        __syncthreads();
    }
    if (lc < NC) {
#pragma unroll
        for (int i = 0; i < RM; ++i) {
            int r = ty + i * TY;
            if (r < nrows) Zs[r * ldz + lc] = acc[i];
        }
    }
    __syncthreads();
}

__device__ __forceinline__ void cta_gemm(int TX, float* __restrict__ Zs, int ldz, const float* __restrict__ A,
                                         int64_t lda, int nrows, int K, const float* __restrict__ Ws,
                                         int ldw, int wcol0, int NC, float* __restrict__ As, bool vec) {
    const int nrm = ceil_div(nrows, REC_THREADS / TX);       // block-uniform
    if (nrm <= 1)      cta_gemm_rm<1>(TX, Zs, ldz, A, lda, nrows, K, Ws, ldw, wcol0, NC, As, vec);
    else if (nrm <= 2) cta_gemm_rm<2>(TX, Zs, ldz, A, lda, nrows, K, Ws, ldw, wcol0, NC, As, vec);
    else if (nrm <= 4) cta_gemm_rm<4>(TX, Zs, ldz, A, lda, nrows, K, Ws, ldw, wcol0, NC, As, vec);
    else if (nrm <= 8) cta_gemm_rm<8>(TX, Zs, ldz, A, lda, nrows, K, Ws, ldw, wcol0, NC, As, vec);
    else               cta_gemm_rm<16>(TX, Zs, ldz, A, lda, nrows, K, Ws, ldw, wcol0, NC, As, vec);
}

// Copy a slice of a row-major weight matrix into shared memory.
//   transpose == 0:  Ws[k][g*HC + jj] = W[(row0 + k) * ldW + g*gstride + j0 + jj]    k < K
//   transpose == 1:  Ws[k][jj]        = W[(row0 + j0 + jj) * ldW + k]                k < K
// Rows K..K4 and columns past the valid hidden units are zero.
__device__ __forceinline__ void load_w_slice(float* Ws, int ldw, const float* __restrict__ W, int64_t ldW,
                                             int row0, int K, int G, int gstride, int j0, int hc, int HC,
                                             int transpose) {
    const int K4 = (K + 3) & ~3;
    const int ncol = transpose ? HC : G * HC;
    for (int idx = threadIdx.x; idx < K4 * ncol; idx += REC_THREADS) {
        int k, c;
        if (transpose) { k = idx % K4; c = idx / K4; } else { c = idx % ncol; k = idx / ncol; }
        float v = 0.f;
        if (k < K) {
            if (transpose) {
                if (c < hc) v = __ldg(W + (int64_t)(row0 + j0 + c) * ldW + k);
            } else {
                int g = c / HC, jj = c % HC;
                if (jj < hc) v = __ldg(W + (int64_t)(row0 + k) * ldW + (int64_t)g * gstride + j0 + jj);
            }
        }
        Ws[k * ldw + c] = v;
    }
}

// ------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------
template <int MODEL>
__global__ void __launch_bounds__(REC_THREADS, 1) rec_fwd_kernel(RecArgs a) {
    extern __shared__ __align__(16) float smem[];
    const int B = a.B, H = a.H, HC = a.HC;
    const int slice = blockIdx.x % a.nslices, tile = blockIdx.x / a.nslices;
    const int j0 = slice * HC, hc = min(HC, H - j0);
    const int b0 = tile * a.BT, bt = min(a.BT, B - b0);
    const int K4 = (H + 3) & ~3;
    constexpr int G = MODEL == CRNN_MODEL_LSTM ? 4 : (MODEL == CRNN_MODEL_GRU ? 2 : 1);
    const int NCA = G * HC;                       // columns of the first contraction
    float* Wsa = smem;                            // [K4][NCA]
    float* Wsb = Wsa + (size_t)K4 * NCA;          // gru: [K4][HC]
    float* As = Wsb + (MODEL == CRNN_MODEL_GRU ? (size_t)K4 * HC : 0);
    float* Zs = As + 2 * REC_RB * REC_LDA;        // [REC_RB][NCA+1]
    const int ldz = NCA + 1;
    const int64_t gw = (int64_t)a.GS * H;         // gate-stash row width
    const int64_t BH = (int64_t)B * H;
    const bool vec = (H & 3) == 0;                // 16-byte aligned rows -> cp.async staging

    if (MODEL == CRNN_MODEL_LSTM) load_w_slice(Wsa, NCA, a.W, 4 * H, H, H, 4, H, j0, hc, HC, 0);
    if (MODEL == CRNN_MODEL_RNN)  load_w_slice(Wsa, NCA, a.W, H, H, H, 1, H, j0, hc, HC, 0);
    if (MODEL == CRNN_MODEL_GRU) {
        load_w_slice(Wsa, NCA, a.W, 2 * H, H, H, 2, H, j0, hc, HC, 0);
        load_w_slice(Wsb, HC, a.Wc, H, H, H, 1, H, j0, hc, HC, 0);
    }
    __syncthreads();

    unsigned bar = 0;
    for (int t = 0; t < a.S; ++t) {
        const float* hprev = a.hx + (int64_t)t * BH;
        float* hnext = a.hx + (int64_t)(t + 1) * BH;
        float* gt = a.gates + (int64_t)t * B * gw;
        for (int r0 = 0; r0 < bt; r0 += REC_RB) {
            const int nr = min(REC_RB, bt - r0);
            cta_gemm(a.TXa, Zs, ldz, hprev + (int64_t)(b0 + r0) * H, H, nr, H, Wsa, NCA, 0, NCA, As, vec);
            for (int idx = threadIdx.x; idx < nr * hc; idx += REC_THREADS) {
                const int r = idx / hc, jj = idx % hc;
                const int64_t b = b0 + r0 + r;
                const int j = j0 + jj;
                float* g = gt + b * gw + j;
                if (MODEL == CRNN_MODEL_LSTM) {
                    float zi = Zs[r * ldz + jj] + g[0];
                    float zj = Zs[r * ldz + HC + jj] + g[H];
                    float zf = Zs[r * ldz + 2 * HC + jj] + g[2 * H];
                    float zo = Zs[r * ldz + 3 * HC + jj] + g[3 * H];
                    float si = sigmoid_acc(zi), tj = tanh_acc(zj);
                    float sf = sigmoid_acc(zf + 1.0f), so = sigmoid_acc(zo);   // forget_bias = 1.0
                    float cp = a.cx[(int64_t)t * BH + b * H + j];
                    float cn = sf * cp + si * tj;
                    float hn = so * tanh_acc(cn);
                    g[0] = si; g[H] = tj; g[2 * H] = sf; g[3 * H] = so;
                    a.cx[(int64_t)(t + 1) * BH + b * H + j] = cn;
                    hnext[b * H + j] = hn;
                } else if (MODEL == CRNN_MODEL_RNN) {
                    hnext[b * H + j] = tanh_acc(Zs[r * ldz + jj] + g[0]);
                } else {
                    float r_ = sigmoid_acc(Zs[r * ldz + jj] + g[0]);
                    float u_ = sigmoid_acc(Zs[r * ldz + HC + jj] + g[H]);
                    float hp = __ldcg(hprev + b * H + j);
                    g[0] = r_; g[H] = u_;
                    g[3 * H] = r_ * hp;                                       // reset BEFORE the matmul
                }
            }
            __syncthreads();
        }
        if (MODEL == CRNN_MODEL_GRU) {
            grid_barrier(a.barrier, (++bar) * gridDim.x);
            for (int r0 = 0; r0 < bt; r0 += REC_RB) {
                const int nr = min(REC_RB, bt - r0);
                cta_gemm(a.TXb, Zs, ldz, gt + (int64_t)(b0 + r0) * gw + 3 * H, gw, nr, H, Wsb, HC, 0, HC, As, vec);
                for (int idx = threadIdx.x; idx < nr * hc; idx += REC_THREADS) {
                    const int r = idx / hc, jj = idx % hc;
                    const int64_t b = b0 + r0 + r;
                    const int j = j0 + jj;
                    float* g = gt + b * gw + j;
                    float cand = tanh_acc(Zs[r * ldz + jj] + g[2 * H]);
                    float u_ = g[H];
                    float hp = __ldcg(hprev + b * H + j);
                    g[2 * H] = cand;
                    hnext[b * H + j] = u_ * hp + (1.0f - u_) * cand;
                }
                __syncthreads();
            }
        }
        if (t + 1 < a.S) grid_barrier(a.barrier, (++bar) * gridDim.x);
    }
}

// ------------------------------------------------------------------------------------
// backward (BPTT).  dh_t = dy_t + (pre-activation grads of step t+1) * W_h^T, so the
// CTA keeps rows [j0, j0+HC) of W_h (transposed) resident instead.
// ------------------------------------------------------------------------------------
template <int MODEL>
__global__ void __launch_bounds__(REC_THREADS, 1) rec_bwd_kernel(RecArgs a) {
    extern __shared__ __align__(16) float smem[];
    const int B = a.B, H = a.H, HC = a.HC;
    const int slice = blockIdx.x % a.nslices, tile = blockIdx.x / a.nslices;
    const int j0 = slice * HC, hc = min(HC, H - j0);
    const int b0 = tile * a.BT, bt = min(a.BT, B - b0);
    constexpr int G = MODEL == CRNN_MODEL_LSTM ? 4 : (MODEL == CRNN_MODEL_GRU ? 2 : 1);
    constexpr int GZ = MODEL == CRNN_MODEL_LSTM ? 4 : (MODEL == CRNN_MODEL_GRU ? 3 : 1);
    const int KA = G * H, KA4 = (KA + 3) & ~3;    // contraction length over gate columns
    const int K4 = (H + 3) & ~3;
    float* Wsa = smem;                            // [KA4][HC]   W_h^T slice (gru: gates)
    float* Wsb = Wsa + (size_t)KA4 * HC;          // gru: [K4][HC] candidate W_h^T slice
    float* As = Wsb + (MODEL == CRNN_MODEL_GRU ? (size_t)K4 * HC : 0);
    float* Zs = As + 2 * REC_RB * REC_LDA;        // [REC_RB][HC+1]
    const int ldz = HC + 1;
    const int64_t gw = (int64_t)a.GS * H;
    const int64_t zw = (int64_t)GZ * H;
    const int64_t BH = (int64_t)B * H;
    const bool vec = (H & 3) == 0;

    load_w_slice(Wsa, HC, a.W, (int64_t)G * H, H, KA, 1, 0, j0, hc, HC, 1);
    if (MODEL == CRNN_MODEL_GRU) load_w_slice(Wsb, HC, a.Wc, H, H, H, 1, 0, j0, hc, HC, 1);
    // zero the CTA-private carries
    for (int idx = threadIdx.x; idx < bt * hc; idx += REC_THREADS) {
        int r = idx / hc, jj = idx % hc;
        int64_t o = (int64_t)(b0 + r) * H + j0 + jj;
        if (MODEL != CRNN_MODEL_RNN) a.carry0[o] = 0.f;
    }
    __syncthreads();

    unsigned bar = 0;
    for (int t = a.S - 1; t >= 0; --t) {
        const float* gt = a.gates + (int64_t)t * B * gw;
        float* dzt = a.dz + (int64_t)t * B * zw;
        const float* dzn = a.dz + (int64_t)(t + 1) * B * zw;
        const float* dyt = a.dy + (int64_t)t * BH;
        const bool has_next = (t + 1 < a.S);
        for (int r0 = 0; r0 < bt; r0 += REC_RB) {
            const int nr = min(REC_RB, bt - r0);
            if (has_next) cta_gemm(a.TXb, Zs, ldz, dzn + (int64_t)(b0 + r0) * zw, zw, nr, KA, Wsa, HC, 0, HC, As, vec);
            for (int idx = threadIdx.x; idx < nr * hc; idx += REC_THREADS) {
                const int r = idx / hc, jj = idx % hc;
                const int64_t b = b0 + r0 + r;
                const int j = j0 + jj;
                const int64_t o = b * H + j;
                float dht = __ldcg(dyt + o) + (has_next ? Zs[r * ldz + jj] : 0.f);
                const float* g = gt + b * gw + j;
                float* dz = dzt + b * zw + j;
                if (MODEL == CRNN_MODEL_LSTM) {
                    float si = g[0], tj = g[H], sf = g[2 * H], so = g[3 * H];
                    float cp = a.cx[(int64_t)t * BH + o];
                    float tc = tanh_acc(a.cx[(int64_t)(t + 1) * BH + o]);
                    float dct = a.carry0[o] + dht * so * (1.0f - tc * tc);
                    dz[0] = dct * tj * si * (1.0f - si);
                    dz[H] = dct * si * (1.0f - tj * tj);
                    dz[2 * H] = dct * cp * sf * (1.0f - sf);
                    dz[3 * H] = dht * tc * so * (1.0f - so);
                    a.carry0[o] = dct * sf;
                } else if (MODEL == CRNN_MODEL_RNN) {
                    float hn = a.hx[(int64_t)(t + 1) * BH + o];
                    dz[0] = dht * (1.0f - hn * hn);
                } else {
                    dht += a.carry0[o];
                    float u_ = g[H], cand = g[2 * H];
                    float hp = a.hx[(int64_t)t * BH + o];
                    float du = dht * (hp - cand);
                    dz[2 * H] = dht * (1.0f - u_) * (1.0f - cand * cand);
                    dz[H] = du * u_ * (1.0f - u_);
                    a.carry1[o] = dht;
                }
            }
            __syncthreads();
        }
        if (MODEL == CRNN_MODEL_GRU) {
            grid_barrier(a.barrier, (++bar) * gridDim.x);
            for (int r0 = 0; r0 < bt; r0 += REC_RB) {
                const int nr = min(REC_RB, bt - r0);
                cta_gemm(a.TXb, Zs, ldz, dzt + (int64_t)(b0 + r0) * zw + 2 * H, zw, nr, H, Wsb, HC, 0, HC, As, vec);
                for (int idx = threadIdx.x; idx < nr * hc; idx += REC_THREADS) {
                    const int r = idx / hc, jj = idx % hc;
                    const int64_t b = b0 + r0 + r;
                    const int j = j0 + jj;
                    const int64_t o = b * H + j;
                    const float* g = gt + b * gw + j;
                    float r_ = g[0], u_ = g[H];
                    float hp = a.hx[(int64_t)t * BH + o];
                    float drs = Zs[r * ldz + jj];
                    dzt[b * zw + j] = drs * hp * r_ * (1.0f - r_);
                    a.carry0[o] = a.carry1[o] * u_ + drs * r_;
                }
                __syncthreads();
            }
        }
        if (t > 0) grid_barrier(a.barrier, (++bar) * gridDim.x);
    }
}

}  // namespace crnn
