// proj_xent_tc.cuh — fused vocabulary projection + softmax + sequence cross-entropy (model.py:76-85) on tcgen05.
//
//   logits = top * softmax_w + softmax_b           [rows = S*B (time-major), V]
//   loss_i = logsumexp(logits_i) - logits_i[y_i]    (sequence_loss_by_example, one "timestep", weights 1)
//   dlogits = (softmax(logits) - onehot(y)) / (S*B) (what tf.gradients sends back into the two projection GEMMs)
//
// One CTA owns 128 rows and the WHOLE vocabulary (V <= 256: one TMEM allocation of 128 lanes x BN fp32 columns), so the
// logits never leave the chip: the accumulator is read back with tcgen05.ld by the thread that owns the row (three cheap
// passes over TMEM: max, sum of exponentials, gradient) and only
//   * dlogits as bf16 (hi, and the lo twin in split precision) - the operand of the two gradient GEMMs,
//   * one loss partial per CTA (double) and one column-sum partial per CTA (the softmax_b gradient), both summed
//     afterwards in CTA order (same bits every run),
//   * fp32 logits when the caller asks for them (crnn_forward: sampling needs probs),
// are written.  HBM bytes per row: H*2*(1+split) read, Vp*2*(1+split) written (was: + V*4 written, read, rewritten and
// read twice more).  Warp roles and the TMA/mbarrier ring are those of gemm_tc.cuh; B = softmax_w is MN-major ([H, Vp]
// as stored), fetched per K block as BN/64 boxes of 64 x 64.
#pragma once
#include "gemm_tc.cuh"

namespace crnn {
namespace tc {

struct ProjXentArgs {
    int M, V, Vp, K;              // rows (S*B), vocabulary, its bf16 row pitch, rnn_size
    int n_mma;                    // MMA N: V rounded up to 16
    int B, S;                     // row i = t*B + b -> label y[b*ldy + t]
    const int32_t* y; int64_t ldy;           // nullptr: forward only (no loss, no gradient)
    const float* bias;
    __nv_bfloat16* dl; int64_t dl_lo;        // [M, Vp] bf16 (+ lo twin dl_lo elements further, 0: none)
    double* loss_part;            // [gridDim.x]
    float* csum_part;             // [gridDim.x][V]
    float* logits;                // optional [M, V] fp32
};

template <int BN, int STAGES, bool SPLIT>
struct ProjXentSmem {
    static constexpr int A_BYTES = GT_BM * GT_BK * 2;
    static constexpr int B_BYTES = BN * GT_BK * 2;
    static constexpr int HALF_BYTES = A_BYTES + B_BYTES;
    static constexpr int STAGE_BYTES = HALF_BYTES * (SPLIT ? 2 : 1);
    static constexpr int BAR_OFF = STAGES * STAGE_BYTES;
    static constexpr int TOTAL = BAR_OFF + 256 + 1024 + BN * 4;   // + barriers/slot + alignment slack + bias
};

template <int BN, int STAGES, bool SPLIT>
__global__ void __launch_bounds__(GT_THREADS) proj_xent_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                                                                  const __grid_constant__ CUtensorMap tmAl, const __grid_constant__ CUtensorMap tmBl,
                                                                  ProjXentArgs g) {
    using L = ProjXentSmem<BN, STAGES, SPLIT>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::BAR_OFF);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tmem_full = empty_bar + STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);
    float* bias_s = reinterpret_cast<float*>(smem + L::BAR_OFF + 256);
    __shared__ double red[32];

    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
    const int m0 = blockIdx.x * GT_BM;
    const int nkb = ceil_div(g.K, GT_BK);

    for (int i = threadIdx.x; i < BN; i += GT_THREADS) bias_s[i] = i < g.V ? __ldg(g.bias + i) : 0.f;
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        for (int s = 0; s < STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        mbar_init(tmem_full, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(tmem_slot, BN);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            for (int i = 0; i < nkb; ++i) {
                const int s = i % STAGES;
                mbar_wait(&empty_bar[s], ((i / STAGES) & 1) ^ 1);
                uint8_t* sa = smem + s * L::STAGE_BYTES;
                mbar_expect_tx(&full_bar[s], L::STAGE_BYTES);
                const int k0 = i * GT_BK;
#pragma unroll
                for (int half = 0; half < (SPLIT ? 2 : 1); ++half) {
                    uint8_t* da = sa + half * L::HALF_BYTES;
                    uint8_t* db = da + L::A_BYTES;
                    tma_load_2d(da, half ? &tmAl : &tmA, &full_bar[s], k0, m0);
#pragma unroll
                    for (int j = 0; j < BN / 64; ++j) tma_load_2d(db + j * 8192, half ? &tmBl : &tmB, &full_bar[s], 64 * j, k0);
                }
            }
        }
    } else if (warp == 1) {
        const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
        const uint32_t idesc = idesc_bf16(GT_BM, g.n_mma, 0, 1);
        for (int i = 0; i < nkb; ++i) {
            const int s = i % STAGES;
            mbar_wait(&full_bar[s], (i / STAGES) & 1);
            tc_fence_after();
            const uint32_t sa = smem_u32(smem + s * L::STAGE_BYTES);
            const uint64_t ad = desc_kmajor(sa), bd = desc_mnmajor(sa + L::A_BYTES, 8192);
            if (elect_one()) {
#pragma unroll
                for (int k = 0; k < 4; ++k) mma_bf16(tmem_u, ad + 2 * k, bd + 128 * k, idesc, (uint32_t)((i | k) != 0));
                if (SPLIT) {                        // + lo(A)*hi(B) + hi(A)*lo(B)
                    const uint64_t lo = (uint64_t)(L::HALF_BYTES >> 4);
#pragma unroll
                    for (int k = 0; k < 4; ++k) mma_bf16(tmem_u, ad + lo + 2 * k, bd + 128 * k, idesc, 1u);
#pragma unroll
                    for (int k = 0; k < 4; ++k) mma_bf16(tmem_u, ad + 2 * k, bd + lo + 128 * k, idesc, 1u);
                }
                mma_commit(&empty_bar[s]);
                if (i + 1 == nkb) mma_commit(tmem_full);
            }
            __syncwarp();
        }
    } else {
        // ===== epilogue: the thread that owns row m reads its logits straight out of TMEM =====
        const int q = warp & 3;
        const int r = 32 * q + lane;
        const int m = m0 + r;
        const bool row_ok = m < g.M;
        const int nchunk = ceil_div(g.V, 32);
        const uint32_t trow = tmem_base + ((uint32_t)(32 * q) << 16);
        int label = -1;
        if (row_ok && g.y) {
            const int t = m / g.B, b = m - t * g.B;
            label = __ldg(g.y + (int64_t)b * g.ldy + t);
        }
        mbar_wait(tmem_full, 0);
        tc_fence_after();
        float v[32];
        // pass 1: row maximum (+ the fp32 logits when asked for)
        float mx = -INFINITY;
        for (int c = 0; c < nchunk; ++c) {
            tmem_ld32(trow + (uint32_t)(c * 32), v);
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                const int col = c * 32 + j;
                const float z = v[j] + bias_s[col];
                if (col < g.V) {
                    mx = fmaxf(mx, z);
                    if (g.logits && row_ok) g.logits[(int64_t)m * g.V + col] = z;
                }
            }
        }
        double loss_local = 0.0;
        if (g.y) {
            // pass 2: sum of exponentials and the label's logit
            float se = 0.f, zl = 0.f;
            for (int c = 0; c < nchunk; ++c) {
                tmem_ld32(trow + (uint32_t)(c * 32), v);
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const int col = c * 32 + j;
                    const float z = v[j] + bias_s[col];
                    if (col < g.V) {
                        se += expf(z - mx);
                        if (col == label) zl = z;
                    }
                }
            }
            const float inv = 1.0f / se, invN = 1.0f / (float)g.M;
            if (row_ok) loss_local = (double)((logf(se) + mx) - zl);
            // pass 3: dlogits as bf16 (hi, lo) and the CTA's column sums (softmax_b gradient)
            float* slab = reinterpret_cast<float*>(smem);                 // [128][33]: the pipeline stages are drained
            float* quart = slab + 128 * 33;                               // [4][32]
            const int npad = ceil_div(g.Vp, 32);
            for (int c = 0; c < npad; ++c) {
                if (c < nchunk) tmem_ld32(trow + (uint32_t)(c * 32), v);
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const int col = c * 32 + j;
                    float d = 0.f;
                    if (c < nchunk && col < g.V && row_ok) d = (expf(v[j] + bias_s[col] - mx) * inv - (col == label ? 1.0f : 0.0f)) * invN;
                    v[j] = d;
                    slab[r * 33 + j] = d;
                }
                if (row_ok) {
                    __nv_bfloat16* dst = g.dl + (int64_t)m * g.Vp + c * 32;
                    const int ncol = min(32, g.Vp - c * 32);              // multiple of 8
                    for (int j0 = 0; j0 < ncol; j0 += 8) {
                        uint32_t hi[4], lo[4];
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const __nv_bfloat162 h2 = __floats2bfloat162_rn(v[j0 + 2 * k], v[j0 + 2 * k + 1]);
                            const float2 f = __bfloat1622float2(h2);
                            const __nv_bfloat162 l2 = __floats2bfloat162_rn(v[j0 + 2 * k] - f.x, v[j0 + 2 * k + 1] - f.y);
                            hi[k] = *reinterpret_cast<const uint32_t*>(&h2);
                            lo[k] = *reinterpret_cast<const uint32_t*>(&l2);
                        }
                        *reinterpret_cast<uint4*>(dst + j0) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                        if (g.dl_lo) *reinterpret_cast<uint4*>(dst + g.dl_lo + j0) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
                    }
                }
                asm volatile("bar.sync 1, 128;" ::: "memory");
                {   // column `lane` over this warp's 32 rows, then the four quarters in order
                    float t = 0.f;
#pragma unroll 8
                    for (int rr = 0; rr < 32; ++rr) t += slab[(32 * q + rr) * 33 + lane];
                    quart[q * 32 + lane] = t;
                }
                asm volatile("bar.sync 1, 128;" ::: "memory");
                if (q == 0 && c * 32 + lane < g.V)
                    g.csum_part[(int64_t)blockIdx.x * g.V + c * 32 + lane] = ((quart[lane] + quart[32 + lane]) + quart[64 + lane]) + quart[96 + lane];
            }
        }
        // the CTA's loss: warp sums in lane order, then the four epilogue warps in order
        loss_local = warp_sum_d(loss_local);
        if (lane == 0) red[q] = loss_local;
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (q == 0 && lane == 0 && g.loss_part) g.loss_part[blockIdx.x] = ((red[0] + red[1]) + red[2]) + red[3];
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, BN);
    }
}

// Host launcher.  top: [M, K] bf16 K-major (lo twin optional); wsm: softmax_w as [K, Vp] bf16 (MN-major B).
// Returns CRNN_ERR_UNSUP for V > 256 (the caller then takes the unfused path).
inline int proj_xent_tc(int M, int V, int Vp, int K, int B, int S, const __nv_bfloat16* top, const __nv_bfloat16* top_lo,
                        const __nv_bfloat16* wsm, const __nv_bfloat16* wsm_lo, const float* bias, const int32_t* y, int64_t ldy,
                        __nv_bfloat16* dl, int64_t dl_lo, double* loss_part, float* csum_part, float* logits, cudaStream_t st,
                        int* nblocks_out) {
    if (V > 256) return CRNN_ERR_UNSUP;
    const bool split = top_lo != nullptr;
    CUtensorMap tmA, tmB, tmAl, tmBl;
    int rc = make_tmap_bf16(&tmA, top, M, K, K, GT_BM);
    if (rc == CRNN_OK) rc = make_tmap_bf16(&tmB, wsm, K, V, Vp, 64);
    tmAl = tmA; tmBl = tmB;
    if (rc == CRNN_OK && split) {
        rc = make_tmap_bf16(&tmAl, top_lo, M, K, K, GT_BM);
        if (rc == CRNN_OK) rc = make_tmap_bf16(&tmBl, wsm_lo, K, V, Vp, 64);
    }
    if (rc != CRNN_OK) return rc;
    ProjXentArgs g;
    g.M = M; g.V = V; g.Vp = Vp; g.K = K; g.n_mma = (V + 15) & ~15; g.B = B; g.S = S; g.y = y; g.ldy = ldy; g.bias = bias;
    g.dl = dl; g.dl_lo = dl_lo; g.loss_part = loss_part; g.csum_part = csum_part; g.logits = logits;
    const int grid = ceil_div(M, GT_BM);
    if (nblocks_out) *nblocks_out = grid;
    static bool attr_done[64] = {};
    int dev = 0;
    CRNN_CUDA(cudaGetDevice(&dev));
    if (!attr_done[dev & 63]) {
        CRNN_CUDA(cudaFuncSetAttribute(proj_xent_tc_kernel<128, 3, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, ProjXentSmem<128, 3, false>::TOTAL));
        CRNN_CUDA(cudaFuncSetAttribute(proj_xent_tc_kernel<128, 3, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, ProjXentSmem<128, 3, true>::TOTAL));
        CRNN_CUDA(cudaFuncSetAttribute(proj_xent_tc_kernel<256, 3, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, ProjXentSmem<256, 3, false>::TOTAL));
        CRNN_CUDA(cudaFuncSetAttribute(proj_xent_tc_kernel<256, 2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, ProjXentSmem<256, 2, true>::TOTAL));
        attr_done[dev & 63] = true;
    }
    if (V <= 128) {
        if (split) proj_xent_tc_kernel<128, 3, true><<<grid, GT_THREADS, ProjXentSmem<128, 3, true>::TOTAL, st>>>(tmA, tmB, tmAl, tmBl, g);
        else       proj_xent_tc_kernel<128, 3, false><<<grid, GT_THREADS, ProjXentSmem<128, 3, false>::TOTAL, st>>>(tmA, tmB, tmAl, tmBl, g);
    } else {
        if (split) proj_xent_tc_kernel<256, 2, true><<<grid, GT_THREADS, ProjXentSmem<256, 2, true>::TOTAL, st>>>(tmA, tmB, tmAl, tmBl, g);
        else       proj_xent_tc_kernel<256, 3, false><<<grid, GT_THREADS, ProjXentSmem<256, 3, false>::TOTAL, st>>>(tmA, tmB, tmAl, tmBl, g);
    }
    CRNN_LAUNCH_CHECK("proj_xent_tc_kernel");
    return CRNN_OK;
}

}  // namespace tc
}  // namespace crnn
