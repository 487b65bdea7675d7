// gemm_tc.cuh — bf16 x bf16 -> fp32 GEMM on the 5th-gen tensor cores (tcgen05 + TMEM + TMA).
//
//   C[M,N] (fp32, row-major) (+)= op(A)[M,K] * op(B)[K,N] (+ bias[N])
//
// The batch-parallel contractions of the training step (CRNN_PREC_BF16): the hoisted
// x-projection of every layer (NN), the input gradients dIn = dZ * Wx^T (NT) and the weight
// gradients dW = In^T * dZ (TN, split-K with fp32 atomics).  Operands are bf16 in global
// memory in whichever orientation their producer wrote them: "K-major" (the contraction
// index is contiguous) or "MN-major" (the M/N index is contiguous); both are fed to
// tcgen05.mma directly through the descriptor's major bits, so nothing is transposed in HBM.
//
// CTA = 6 warps: warp 0 TMA producer, warp 1 MMA issuer (one thread), warps 2-5 epilogue
// (TMEM -> registers -> global).  128 x BN accumulator tile in TMEM, BLOCK_K = 64
// (128-byte swizzled rows), STAGES-deep mbarrier ring between TMA and MMA.
#pragma once
#include "tc_common.cuh"
#include <algorithm>

namespace crnn {
namespace tc {

constexpr int GT_BM = 128;
constexpr int GT_BK = 64;
constexpr int GT_THREADS = 192;

struct GemmTcArgs {
    float* C; int64_t ldc;
    const float* bias;
    int M, N, K;
    int kblocks_per_split;   // BLOCK_K blocks per blockIdx.z
    int mode;                // 0 store, 1 C += (no atomics), 2 atomicAdd (split-K)
    int tma_out;             // 1: epilogue stages 128x32 slabs in shared memory and writes them with TMA
                             //    (store, or reduce-add for modes 1 and 2) instead of per-thread stores
    int a_mn, b_mn;          // 1: operand is MN-major in global memory
};

// SPLIT: both operands come as (hi, lo) bf16 pairs (see store_bf16_split); a stage then holds
// A_hi, B_hi, A_lo, B_lo and every K step issues hi*hi + lo*hi + hi*lo into the same accumulator.
template <int BN, int STAGES, bool SPLIT>
struct GemmTcSmem {
    static constexpr int A_BYTES = GT_BM * GT_BK * 2;       // 16 KB
    static constexpr int B_BYTES = BN * GT_BK * 2;
    static constexpr int HALF_BYTES = A_BYTES + B_BYTES;
    static constexpr int STAGE_BYTES = HALF_BYTES * (SPLIT ? 2 : 1);
    static constexpr int BAR_OFF = STAGES * STAGE_BYTES;
    static constexpr int TOTAL = BAR_OFF + 256 + 1024;       // + barriers/slot + alignment slack
};

template <int BN, int STAGES, bool SPLIT>
__global__ void __launch_bounds__(GT_THREADS) gemm_bf16_tc_kernel(const __grid_constant__ CUtensorMap tmA,
                                                                  const __grid_constant__ CUtensorMap tmB,
                                                                  const __grid_constant__ CUtensorMap tmAl,
                                                                  const __grid_constant__ CUtensorMap tmBl,
                                                                  const __grid_constant__ CUtensorMap tmC,
                                                                  GemmTcArgs g) {
    using L = GemmTcSmem<BN, STAGES, SPLIT>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L::BAR_OFF);
    uint64_t* empty_bar = full_bar + STAGES;
    uint64_t* tmem_full = empty_bar + STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;   // warp-uniform for ptxas
    const int m0 = blockIdx.y * GT_BM, n0 = blockIdx.x * BN;
    const int total_kb = ceil_div(g.K, GT_BK);
    const int kb0 = blockIdx.z * g.kblocks_per_split;
    const int nkb = min(g.kblocks_per_split, total_kb - kb0);

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
        // ===== TMA producer =====
        if (lane == 0) {
            for (int i = 0; i < nkb; ++i) {
                const int s = i % STAGES;
                const uint32_t ph = (i / STAGES) & 1;
                mbar_wait(&empty_bar[s], ph ^ 1);
                uint8_t* sa = smem + s * L::STAGE_BYTES;
                uint8_t* sb = sa + L::A_BYTES;
                mbar_expect_tx(&full_bar[s], L::STAGE_BYTES);
                const int k0 = (kb0 + i) * GT_BK;
#pragma unroll
                for (int half = 0; half < (SPLIT ? 2 : 1); ++half) {
                    const CUtensorMap* ma = half ? &tmAl : &tmA;
                    const CUtensorMap* mb = half ? &tmBl : &tmB;
                    uint8_t* da = sa + half * L::HALF_BYTES;
                    uint8_t* db = sb + half * L::HALF_BYTES;
                    if (!g.a_mn) {
                        tma_load_2d(da, ma, &full_bar[s], k0, m0);
                    } else {
#pragma unroll
                        for (int j = 0; j < GT_BM / 64; ++j) tma_load_2d(da + j * 8192, ma, &full_bar[s], m0 + 64 * j, k0);
                    }
                    if (!g.b_mn) {
                        tma_load_2d(db, mb, &full_bar[s], k0, n0);
                    } else {
#pragma unroll
                        for (int j = 0; j < BN / 64; ++j) tma_load_2d(db + j * 8192, mb, &full_bar[s], n0 + 64 * j, k0);
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer: the whole warp runs the loop, one elected lane issues (operands stay in uniform registers) =====
        {
            const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem_base, 0);
            const uint32_t idesc = idesc_bf16(GT_BM, BN, g.a_mn, g.b_mn);
            for (int i = 0; i < nkb; ++i) {
                const int s = i % STAGES;
                const uint32_t ph = (i / STAGES) & 1;
                mbar_wait(&full_bar[s], ph);
                tc_fence_after();
                const uint32_t sa = smem_u32(smem + s * L::STAGE_BYTES);
                const uint32_t sb = sa + L::A_BYTES;
                // one descriptor per operand per stage, advanced by adds (K step: 32 B K-major, 2048 B MN-major)
                const uint64_t ad = g.a_mn ? desc_mnmajor(sa, 8192) : desc_kmajor(sa);
                const uint64_t bd = g.b_mn ? desc_mnmajor(sb, 8192) : desc_kmajor(sb);
                const uint64_t as = g.a_mn ? 128u : 2u, bs = g.b_mn ? 128u : 2u;
                if (elect_one()) {
                    mma_bf16(tmem_u, ad, bd, idesc, (uint32_t)(i != 0));
                    mma_bf16(tmem_u, ad + as, bd + bs, idesc, 1u);
                    mma_bf16(tmem_u, ad + 2 * as, bd + 2 * bs, idesc, 1u);
                    mma_bf16(tmem_u, ad + 3 * as, bd + 3 * bs, idesc, 1u);
                    if (SPLIT) {                    // + lo(A)*hi(B) + hi(A)*lo(B)
                        const uint64_t lo = (uint64_t)(L::HALF_BYTES >> 4);
#pragma unroll
                        for (int k = 0; k < 4; ++k) mma_bf16(tmem_u, ad + lo + k * as, bd + k * bs, idesc, 1u);
#pragma unroll
                        for (int k = 0; k < 4; ++k) mma_bf16(tmem_u, ad + k * as, bd + lo + k * bs, idesc, 1u);
                    }
                    mma_commit(&empty_bar[s]);          // frees the smem slot once these MMAs retire
                    if (i + 1 == nkb) mma_commit(tmem_full);   // accumulator complete
                }
                __syncwarp();
            }
            if (nkb <= 0 && elect_one()) mma_commit(tmem_full);
        }
    } else {
        // ===== epilogue: TMEM -> registers -> global =====
        const int q = warp & 3;                     // TMEM lane quarter this warp may access
        const int m = m0 + 32 * q + lane;
        mbar_wait(tmem_full, 0);
        tc_fence_after();
        float v[32];
        if (g.tma_out) {
            // All TMA loads have been consumed and all MMAs have retired (tmem_full): the pipeline stages
            // are free, two of them become 128x32 fp32 staging slabs (128-byte swizzled rows).  A thread owns
            // one accumulator row; it writes eight 16-byte pieces XOR-swizzled by (row & 7) - the layout
            // the output tensor map expects - then ONE thread hands the slab to the TMA unit, which clips
            // at the matrix edges and either stores or adds in L2.
            const int r = 32 * q + lane;
            const int epi_tid = (warp - 2) * 32 + lane;
            const bool add_bias = g.bias != nullptr && blockIdx.z == 0;
#pragma unroll 1
            for (int c = 0; c < BN / 32; ++c) {
                const int nb = n0 + c * 32;
                if (nb >= g.N) break;               // warp-uniform
                uint8_t* slab = smem + (c & 1) * (128 * 128);
                if (c >= 2) {                       // the store issued from this slab two rounds ago has read it
                    if (epi_tid == 0) bulk_wait_read<1>();
                    asm volatile("bar.sync 1, 128;" ::: "memory");
                }
                tmem_ld32(tmem_base + ((uint32_t)(32 * q) << 16) + (uint32_t)(c * 32), v);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    float4 o = make_float4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                    if (add_bias) {
                        const int n = nb + 4 * j;
                        if (n + 3 < g.N) {
                            float4 b4 = __ldg(reinterpret_cast<const float4*>(g.bias + n));
                            o.x += b4.x; o.y += b4.y; o.z += b4.z; o.w += b4.w;
                        } else {
                            if (n < g.N) o.x += __ldg(g.bias + n);
                            if (n + 1 < g.N) o.y += __ldg(g.bias + n + 1);
                            if (n + 2 < g.N) o.z += __ldg(g.bias + n + 2);
                        }
                    }
                    *reinterpret_cast<float4*>(slab + r * 128 + ((j ^ (r & 7)) << 4)) = o;
                }
                fence_proxy_async();                // generic writes -> visible to the async proxy
                asm volatile("bar.sync 1, 128;" ::: "memory");
                if (epi_tid == 0 && nkb > 0) {
                    if (g.mode == 0) tma_store_2d(&tmC, slab, nb, m0);
                    else             tma_reduce_add_2d(&tmC, slab, nb, m0);
                    bulk_commit();
                }
            }
            if (epi_tid == 0) bulk_wait<0>();
        } else {
#pragma unroll 1
        for (int c = 0; c < BN / 32; ++c) {
            const int nb = n0 + c * 32;
            if (nb >= g.N) break;                   // warp-uniform
            tmem_ld32(tmem_base + ((uint32_t)(32 * q) << 16) + (uint32_t)(c * 32), v);
            if (m < g.M && nkb > 0) {
                float* crow = g.C + (int64_t)m * g.ldc + nb;
                const bool add_bias = g.bias != nullptr && blockIdx.z == 0;
                const bool vec = (nb + 32 <= g.N) && ((g.ldc & 3) == 0) && (((uintptr_t)g.C & 15) == 0) && g.mode != 2;
                if (vec) {
#pragma unroll
                    for (int j = 0; j < 32; j += 4) {
                        float4 o = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
                        if (add_bias) {
                            float4 b4 = __ldg(reinterpret_cast<const float4*>(g.bias + nb + j));
                            o.x += b4.x; o.y += b4.y; o.z += b4.z; o.w += b4.w;
                        }
                        float4* dst = reinterpret_cast<float4*>(crow + j);
                        if (g.mode == 1) { float4 old = *dst; o.x += old.x; o.y += old.y; o.z += old.z; o.w += old.w; }
                        *dst = o;
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        if (nb + j < g.N) {
                            float o = v[j];
                            if (add_bias) o += __ldg(g.bias + nb + j);
                            if (g.mode == 2) atomicAdd(crow + j, o);
                            else if (g.mode == 1) crow[j] += o;
                            else crow[j] = o;
                        }
                    }
                }
            }
        }
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, BN);
    }
}

inline bool env_splits_round_up() { static int v = -1; if (v < 0) { const char* e = getenv("CRNN_GEMM_SPLITS_UP"); v = (e && *e == '1') ? 1 : 0; } return v == 1; }

// Host launcher.  A: op(A) is [M,K]; a_mn = 0 -> stored [M,K] row-major (lda), a_mn = 1 -> stored
// [K,M] row-major.  B: op(B) is [K,N]; b_mn = 0 -> stored [N,K] row-major (i.e. B^T), b_mn = 1 ->
// stored [K,N] row-major.  accumulate = 1 adds into C (split-K with atomics when that fills the GPU).
template <int BN, int STAGES>
inline int gemm_bf16_tc_impl(int M, int N, int K, const __nv_bfloat16* A, int64_t lda, int a_mn,
                        const __nv_bfloat16* B, int64_t ldb, int b_mn, float* C, int64_t ldc,
                        const float* bias, int accumulate, cudaStream_t st, int sm_count,
                        const __nv_bfloat16* A_lo, const __nv_bfloat16* B_lo) {
    if (M <= 0 || N <= 0) return CRNN_OK;
    if (K <= 0) return set_err(CRNN_ERR_ARG, "gemm_bf16_tc: K must be positive");
    if ((A_lo == nullptr) != (B_lo == nullptr)) return set_err(CRNN_ERR_ARG, "gemm_bf16_tc: both operands or neither must be split");
    const bool split = A_lo != nullptr;
    CUtensorMap tmA, tmB, tmAl, tmBl;
    int rc = a_mn ? make_tmap_bf16(&tmA, A, K, M, lda, 64) : make_tmap_bf16(&tmA, A, M, K, lda, GT_BM);
    if (rc != CRNN_OK) return rc;
    rc = b_mn ? make_tmap_bf16(&tmB, B, K, N, ldb, 64) : make_tmap_bf16(&tmB, B, N, K, ldb, BN);
    if (rc != CRNN_OK) return rc;
    tmAl = tmA; tmBl = tmB;
    if (split) {
        rc = a_mn ? make_tmap_bf16(&tmAl, A_lo, K, M, lda, 64) : make_tmap_bf16(&tmAl, A_lo, M, K, lda, GT_BM);
        if (rc != CRNN_OK) return rc;
        rc = b_mn ? make_tmap_bf16(&tmBl, B_lo, K, N, ldb, 64) : make_tmap_bf16(&tmBl, B_lo, N, K, ldb, BN);
        if (rc != CRNN_OK) return rc;
    }
    CUtensorMap tmC = tmA;                               // placeholder when the TMA epilogue is not usable
    const int tma_out = ((ldc & 3) == 0 && ((uintptr_t)C & 15) == 0 && ((uintptr_t)bias & 15) == 0) ? 1 : 0;
    if (tma_out) {
        rc = make_tmap_f32(&tmC, C, M, N, ldc, GT_BM);
        if (rc != CRNN_OK) return rc;
    }
    GemmTcArgs g;
    g.tma_out = tma_out;
    g.C = C; g.ldc = ldc; g.bias = bias; g.M = M; g.N = N; g.K = K; g.a_mn = a_mn; g.b_mn = b_mn;
    dim3 grid(ceil_div(N, BN), ceil_div(M, GT_BM), 1);
    const int total_kb = ceil_div(K, GT_BK);
    int splits = 1;
    const int64_t tiles = (int64_t)grid.x * grid.y;
    // Split K so that the grid is ONE wave: the split kernel (192 KB of stages) is resident once per SM, the plain one twice.
    // (Rounding the split count up instead - 336 CTAs on 148 slots for config 3's weight gradients - runs a nearly empty
    //  third wave.)
    const int64_t slots = (int64_t)sm_count * (split ? 1 : 2);
    if (accumulate && tiles < slots && total_kb >= 16)
        splits = (int)std::min<int64_t>(std::max<int64_t>(1, env_splits_round_up() ? ceil_div64(slots, tiles) : slots / tiles), total_kb / 8);
    splits = std::max(1, splits);
    g.kblocks_per_split = ceil_div(total_kb, splits);
    grid.z = ceil_div(total_kb, g.kblocks_per_split);
    g.mode = accumulate ? (grid.z > 1 ? 2 : 1) : 0;
    static bool attr_done[2][64] = {};                   // per variant and device (function attributes are per context)
    int dev = 0;
    CRNN_CUDA(cudaGetDevice(&dev));
    bool& attr_set = attr_done[split ? 1 : 0][dev & 63];
    if (!attr_set) {                                     // (only the variant that is launched: <256, 3, split> would not fit)
        if (split) CRNN_CUDA(cudaFuncSetAttribute(gemm_bf16_tc_kernel<BN, STAGES, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                  GemmTcSmem<BN, STAGES, true>::TOTAL));
        else       CRNN_CUDA(cudaFuncSetAttribute(gemm_bf16_tc_kernel<BN, STAGES, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                  GemmTcSmem<BN, STAGES, false>::TOTAL));
        attr_set = true;
    }
    if (split) gemm_bf16_tc_kernel<BN, STAGES, true><<<grid, GT_THREADS, GemmTcSmem<BN, STAGES, true>::TOTAL, st>>>(tmA, tmB, tmAl, tmBl, tmC, g);
    else       gemm_bf16_tc_kernel<BN, STAGES, false><<<grid, GT_THREADS, GemmTcSmem<BN, STAGES, false>::TOTAL, st>>>(tmA, tmB, tmAl, tmBl, tmC, g);
    CRNN_LAUNCH_CHECK("gemm_bf16_tc_kernel");
    return CRNN_OK;
}

inline int gemm_bf16_tc(int M, int N, int K, const __nv_bfloat16* A, int64_t lda, int a_mn,
                        const __nv_bfloat16* B, int64_t ldb, int b_mn, float* C, int64_t ldc,
                        const float* bias, int accumulate, cudaStream_t st, int sm_count = 148,
                        const __nv_bfloat16* A_lo = nullptr, const __nv_bfloat16* B_lo = nullptr) {
    // Split operands: a stage of the 128 x 128 tile (A hi+lo 32 KB, B hi+lo 32 KB) feeds twelve MMAs of 64 cycles - the TMA
    // stream (64 KB per 768 tensor cycles) and the tensor pipe are then about level and the kernel reaches ~0.7 of the tensor
    // peak; 128 x 256 tiles (96 KB per stage, twelve MMAs of 128 cycles, two stages) are tensor-bound.  CRNN_GEMM_BN256=0: off.
    static const int bn256 = [] { const char* e = getenv("CRNN_GEMM_BN256"); return (e && *e == '0') ? 0 : 1; }();
    // Only where the wider tiles still fill the GPU (measured, ms per step: config 5 113.8 -> 109.8 with them, config 3's chunk
    // GEMMs - 80 tiles on the SMs the recurrence kernels leave - 4.09 -> 4.14, config 1 0.58 -> 0.62): at least two waves of
    // tiles, or a long-K accumulating GEMM whose split K fills the rest.
    const int64_t tiles256 = (int64_t)ceil_div(M, GT_BM) * (N / 256);
    const bool big = tiles256 >= 2LL * sm_count || (accumulate && K >= 256 * GT_BK && tiles256 >= 64);
    // (plain bf16 operands keep 128 x 128 with two CTAs per SM: 128 x 256 x 3 stages, one CTA per SM, measured slower -
    //  config 4's x-projection 0.86 -> 1.14 ms, config 5's 4.19 -> 5.19)
    if (A_lo != nullptr && bn256 && N >= 256 && N % 256 == 0 && big)
        return gemm_bf16_tc_impl<256, 2>(M, N, K, A, lda, a_mn, B, ldb, b_mn, C, ldc, bias, accumulate, st, sm_count, A_lo, B_lo);
    return gemm_bf16_tc_impl<128, 3>(M, N, K, A, lda, a_mn, B, ldb, b_mn, C, ldc, bias, accumulate, st, sm_count, A_lo, B_lo);
}

// fp32 -> bf16 cast of a row-major matrix (optionally re-pitched): the bf16 shadow copies.
__global__ void __launch_bounds__(256) cast_bf16_kernel(const float* __restrict__ src, int64_t lds, int64_t rows, int cols,
                                                        __nv_bfloat16* __restrict__ dst, int64_t ldd, int64_t lo_ofs = 0) {
    const int64_t n = rows * cols;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        int64_t r = i / cols;
        int c = (int)(i % cols);
        store_bf16_split(dst + r * ldd + c, lo_ofs, src[r * lds + c]);
    }
}

}  // namespace tc
}  // namespace crnn
