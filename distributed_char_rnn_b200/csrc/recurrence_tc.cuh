// recurrence_tc.cuh — persistent tensor-core recurrence / BPTT kernels (CRNN_PREC_BF16).
//
// The serial part of the reference's unrolled MultiRNNCell (model.py:72) and of its
// reverse-mode mirror (model.py:88-91): per time step only h_{t-1} * W_h (forward) or
// dz_{t+1} * W_h^T (backward) remains once the x halves are hoisted into batch GEMMs.
//
// One cooperative launch per layer and direction.  CTA (slice, tile) owns hidden units
// [j0, j0+HC) for batch rows [b0, b0+BT).  Its slice of the recurrent weights is loaded by TMA
// ONCE into shared memory as a K-major, 128-byte-swizzled UMMA operand and stays resident
// for all S steps.  Per step: warp 0 streams the previous step's bf16 rows (all CTAs' output,
// via L2) through a 4-stage TMA ring, one thread of warp 1 issues tcgen05.mma (M=128 batch
// rows x N=gate columns, fp32 accumulators in TMEM), warps 2-5 read the accumulators with
// tcgen05.ld, fuse bias(x-projection) + sigmoid/tanh + the cell update in registers, publish
// h_t (fp32 for the optimiser side, bf16 for the next step's MMA) and join a grid barrier.
// c_t / gate activations are stashed for BPTT on the way.
#pragma once
#include "tc_common.cuh"

namespace crnn {
namespace tc {

constexpr int RT_THREADS = 192;
constexpr int RT_STAGES = 4;
constexpr int RT_ASTAGE = 128 * 128;      // bytes: 128 rows x 64 bf16 (one K block of the A tile)

struct RecTcArgs {
    int B, H, S;
    int HC, BT, nslices, ntiles, npass;
    int GS;                     // gate-stash row width / H
    int box_rows;               // rows per A-tile TMA box
    int fast_act;               // 1: MUFU.TANH activations
    float* gates;               // [S,B,GS*H]  in: x-projection + bias, out: activations
    float* hx;                  // [(S+1),B,H]
    float* cx;                  // lstm [(S+1),B,H]
    const float* dy;            // bwd [S,B,H]
    float* dz;                  // bwd [S,B,GZ*H]
    float* carry0;              // bwd [B,H]
    float* carry1;              // bwd [B,H]
    __nv_bfloat16* hxb;         // [(S+1),B,H] bf16 copy of hx (fwd out; MMA operand)
    __nv_bfloat16* rsb;         // gru [S,B,H] bf16 r*h (fwd out; MMA operand)
    __nv_bfloat16* dzb;         // bwd [(S+1),B,GZ*H] bf16 copy of dz (MMA operand)
    unsigned* barrier;
};

__device__ __forceinline__ float tanh_mufu(float x) {
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// `fast` (block-uniform): one MUFU.TANH per activation (|err| ~ 5e-4, well under bf16's 4e-3
// operand rounding) instead of the accurate expf/tanhf sequences.
#define fast_tanh(x) (a.fast_act ? tanh_mufu(x) : tanhf(x))
#define fast_sigmoid(x) (a.fast_act ? fmaf(0.5f, tanh_mufu(0.5f * (x)), 0.5f) : 1.0f / (1.0f + expf(-(x))))

__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
}

struct RtSmem {
    uint8_t* wsa; uint8_t* wsb; uint8_t* ring;
    uint64_t *a_full, *a_empty, *z_full, *z_empty, *w_bar;
    uint32_t* tmem_slot;
};

__device__ __forceinline__ RtSmem rt_carve(uint8_t* smem_raw, size_t wsa_bytes, size_t wsb_bytes) {
    RtSmem s;
    uint8_t* p = reinterpret_cast<uint8_t*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    s.wsa = p; p += wsa_bytes;
    s.wsb = p; p += wsb_bytes;
    s.ring = p; p += RT_STAGES * RT_ASTAGE;
    s.a_full = reinterpret_cast<uint64_t*>(p);
    s.a_empty = s.a_full + RT_STAGES;
    s.z_full = s.a_empty + RT_STAGES;
    s.z_empty = s.z_full + 1;
    s.w_bar = s.z_empty + 1;
    s.tmem_slot = reinterpret_cast<uint32_t*>(s.w_bar + 1);
    return s;
}

inline size_t rt_smem_bytes(size_t wsa_bytes, size_t wsb_bytes) {
    return wsa_bytes + wsb_bytes + RT_STAGES * RT_ASTAGE + 256 + 1024;
}

// Role-private pipeline state.
struct RtPipe {
    uint32_t ring_i;      // A-ring blocks issued / consumed so far
    uint32_t z_i;         // accumulator hand-offs so far
};

// One contraction Z[128 x NC] = A_tile[rows x K] * W_region^T, split across the three roles.
// `arow`: first row of the A tile in the 2-D tensor view; `wbase`: resident weight region.
__device__ __forceinline__ void rt_produce(const RtSmem& s, RtPipe& p, const CUtensorMap* tmA, int arow, int nkb, int box_rows) {
    for (int kb = 0; kb < nkb; ++kb, ++p.ring_i) {
        const int st = p.ring_i % RT_STAGES;
        mbar_wait(&s.a_empty[st], ((p.ring_i / RT_STAGES) & 1) ^ 1);
        mbar_expect_tx(&s.a_full[st], (uint32_t)box_rows * 128u);
        tma_load_2d(s.ring + st * RT_ASTAGE, tmA, &s.a_full[st], kb * 64, arow);
    }
}
__device__ __forceinline__ void rt_mma(const RtSmem& s, RtPipe& p, uint32_t tmem, const uint8_t* wbase, int nc, int nkb) {
    mbar_wait(s.z_empty, (p.z_i & 1) ^ 1);          // epilogue has drained the previous accumulator
    tc_fence_after();
    const uint32_t idesc = idesc_bf16(128, nc, 0, 0);
    const uint32_t w0 = smem_u32(wbase);
    for (int kb = 0; kb < nkb; ++kb, ++p.ring_i) {
        const int st = p.ring_i % RT_STAGES;
        mbar_wait(&s.a_full[st], (p.ring_i / RT_STAGES) & 1);
        tc_fence_after();
        const uint32_t sa = smem_u32(s.ring + st * RT_ASTAGE);
        const uint32_t sb = w0 + (uint32_t)kb * (uint32_t)nc * 128u;
#pragma unroll
        for (int k = 0; k < 4; ++k)
            mma_bf16(tmem, desc_kmajor(sa + k * 32), desc_kmajor(sb + k * 32), idesc, (kb | k) != 0);
        mma_commit(&s.a_empty[st]);
    }
    mma_commit(s.z_full);
    ++p.z_i;
}

// Grid barrier, split: epilogue threads publish, the TMA producer waits before touching
// rows written by other CTAs (generic-proxy writes -> async-proxy reads need the proxy fence).
__device__ __forceinline__ void rt_publish(unsigned* counter, int epi_tid) {
    __threadfence();
    epi_bar_sync();
    if (epi_tid == 0) {
        __threadfence();
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(counter) : "memory");
    }
}
__device__ __forceinline__ void rt_await(const unsigned* counter, unsigned target) {
    while (ld_acquire_u32(counter) < target) { }
    asm volatile("fence.proxy.async;" ::: "memory");
}

// ------------------------------------------------------------------------------------------
// forward.  Packed weights (pack_rec_weights_kernel): row (slice*NCW + lc) of Wr is the column of
// W_h this CTA's local column lc contracts with, K contiguous.  lc = jj*G1 + g for the first
// contraction (all gates of one unit adjacent -> one tcgen05.ld chunk holds whole units);
// GRU candidate rows follow at lc = 2*HC + jj.
// ------------------------------------------------------------------------------------------
template <int MODEL>
__global__ void __launch_bounds__(RT_THREADS, 1) rec_fwd_tc_kernel(const __grid_constant__ CUtensorMap tmH,
                                                                   const __grid_constant__ CUtensorMap tmRS,
                                                                   const __grid_constant__ CUtensorMap tmW0,
                                                                   const __grid_constant__ CUtensorMap tmW1,
                                                                   RecTcArgs a) {
    extern __shared__ uint8_t smem_raw[];
    constexpr int G1 = MODEL == CRNN_MODEL_LSTM ? 4 : (MODEL == CRNN_MODEL_GRU ? 2 : 1);
    constexpr int GW = MODEL == CRNN_MODEL_GRU ? 3 : G1;
    const int B = a.B, H = a.H, HC = a.HC;
    const int NCA = G1 * HC;
    const int nkb = ceil_div(H, 64);
    const RtSmem s = rt_carve(smem_raw, (size_t)nkb * NCA * 128, MODEL == CRNN_MODEL_GRU ? (size_t)nkb * HC * 128 : 0);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int slice = blockIdx.x % a.nslices, tile = blockIdx.x / a.nslices;
    const int j0 = slice * HC;
    const int b0 = tile * a.BT, bt = min(a.BT, B - b0);
    const int64_t gw = (int64_t)a.GS * H, BH = (int64_t)B * H;
    uint32_t tcols = 32;
    while ((int)tcols < NCA) tcols <<= 1;

    // zero the ring once: rows past the TMA box are never written and must not hold NaNs
    for (int i = threadIdx.x; i < RT_STAGES * RT_ASTAGE / 16; i += RT_THREADS)
        reinterpret_cast<uint4*>(s.ring)[i] = make_uint4(0, 0, 0, 0);
    fence_proxy_async();
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmH); tma_prefetch_desc(&tmW0);
        for (int i = 0; i < RT_STAGES; ++i) { mbar_init(&s.a_full[i], 1); mbar_init(&s.a_empty[i], 1); }
        mbar_init(s.z_full, 1); mbar_init(s.z_empty, 128); mbar_init(s.w_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(s.tmem_slot, tcols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *s.tmem_slot;

    RtPipe p{0, 0};
    unsigned phases_done = 0;
    const int nphase = MODEL == CRNN_MODEL_GRU ? 2 : 1;

    if (warp == 0) {
        if (lane == 0) {
            // resident weights: once
            uint32_t wbytes = (uint32_t)nkb * NCA * 128u + (MODEL == CRNN_MODEL_GRU ? (uint32_t)nkb * HC * 128u : 0u);
            mbar_expect_tx(s.w_bar, wbytes);
            for (int kb = 0; kb < nkb; ++kb) {
                tma_load_2d(s.wsa + (size_t)kb * NCA * 128, &tmW0, s.w_bar, kb * 64, slice * GW * HC);
                if (MODEL == CRNN_MODEL_GRU)
                    tma_load_2d(s.wsb + (size_t)kb * HC * 128, &tmW1, s.w_bar, kb * 64, slice * GW * HC + NCA);
            }
            for (int t = 0; t < a.S; ++t) {
                for (int ph = 0; ph < nphase; ++ph) {
                    if (phases_done > 0) rt_await(a.barrier, phases_done * gridDim.x);
                    for (int ps = 0; ps < a.npass; ++ps) {
                        const int row = (ph == 0 ? t * B : t * B) + b0 + ps * 128;
                        rt_produce(s, p, ph == 0 ? &tmH : &tmRS, row, nkb, a.box_rows);
                    }
                    ++phases_done;
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            mbar_wait(s.w_bar, 0);
            for (int t = 0; t < a.S; ++t)
                for (int ph = 0; ph < nphase; ++ph)
                    for (int ps = 0; ps < a.npass; ++ps)
                        rt_mma(s, p, tmem, ph == 0 ? s.wsa : s.wsb, ph == 0 ? NCA : HC, nkb);
        }
    } else {
        const int q = warp & 3;
        const int epi_tid = (warp - 2) * 32 + lane;
        const uint32_t tlane = tmem + ((uint32_t)(32 * q) << 16);
        float v[32];
        for (int t = 0; t < a.S; ++t) {
            for (int ph = 0; ph < nphase; ++ph) {
                for (int ps = 0; ps < a.npass; ++ps) {
                    const int r = ps * 128 + 32 * q + lane;       // row within the batch tile
                    const bool row_ok = r < bt;
                    const int64_t b = b0 + r;
                    float* g = a.gates + ((int64_t)t * B + b) * gw;
                    const float* hprev = a.hx + (int64_t)t * BH + b * H;
                    float* hnext = a.hx + (int64_t)(t + 1) * BH + b * H;
                    __nv_bfloat16* hnb = a.hxb + ((int64_t)(t + 1) * B + b) * H;
                    mbar_wait(s.z_full, p.z_i & 1);
                    tc_fence_after();
                    const int nc = ph == 0 ? NCA : HC;
                    for (int c0 = 0; c0 < nc; c0 += 32) {
                        tmem_ld32(tlane + (uint32_t)c0, v);
                        if (!row_ok) continue;
                        if (MODEL == CRNN_MODEL_LSTM) {
                            // 8 units x (i,j,f,o)
                            const int jb = j0 + c0 / 4;
                            float hn[8];
#pragma unroll
                            for (int u = 0; u < 8; ++u) {
                                const int j = jb + u;
                                hn[u] = 0.f;
                                if (j < H) {
                                    float si = fast_sigmoid(v[4 * u] + g[j]);
                                    float tj = fast_tanh(v[4 * u + 1] + g[H + j]);
                                    float sf = fast_sigmoid(v[4 * u + 2] + g[2 * H + j] + 1.0f);
                                    float so = fast_sigmoid(v[4 * u + 3] + g[3 * H + j]);
                                    float cp = a.cx[(int64_t)t * BH + b * H + j];
                                    float cn = fmaf(sf, cp, si * tj);
                                    hn[u] = so * fast_tanh(cn);
                                    g[j] = si; g[H + j] = tj; g[2 * H + j] = sf; g[3 * H + j] = so;
                                    a.cx[(int64_t)(t + 1) * BH + b * H + j] = cn;
                                    hnext[j] = hn[u];
                                }
                            }
                            if (jb < H) {   // H % 8 == 0: the 8 units are all valid or all past the end
                                uint4 pk = make_uint4(pack_bf16x2(hn[0], hn[1]), pack_bf16x2(hn[2], hn[3]),
                                                      pack_bf16x2(hn[4], hn[5]), pack_bf16x2(hn[6], hn[7]));
                                *reinterpret_cast<uint4*>(hnb + jb) = pk;
                            }
                        } else if (MODEL == CRNN_MODEL_RNN) {
#pragma unroll
                            for (int u = 0; u < 32; u += 8) {
                                const int jb = j0 + c0 + u;
                                if (c0 + u < HC && jb < H) {
                                    float hn[8];
#pragma unroll
                                    for (int e = 0; e < 8; ++e) {
                                        hn[e] = fast_tanh(v[u + e] + g[jb + e]);
                                        hnext[jb + e] = hn[e];
                                    }
                                    *reinterpret_cast<uint4*>(hnb + jb) = make_uint4(pack_bf16x2(hn[0], hn[1]), pack_bf16x2(hn[2], hn[3]),
                                                                                    pack_bf16x2(hn[4], hn[5]), pack_bf16x2(hn[6], hn[7]));
                                }
                            }
                        } else if (ph == 0) {
                            // GRU gates: 16 units x (r,u); publish r*h for the candidate contraction
                            __nv_bfloat16* rsb = a.rsb + ((int64_t)t * B + b) * H;
#pragma unroll
                            for (int u = 0; u < 16; u += 8) {
                                const int jb = j0 + c0 / 2 + u;
                                if (jb < H) {
                                    float rs[8];
#pragma unroll
                                    for (int e = 0; e < 8; ++e) {
                                        const int j = jb + e;
                                        float r_ = fast_sigmoid(v[2 * (u + e)] + g[j]);
                                        float u_ = fast_sigmoid(v[2 * (u + e) + 1] + g[H + j]);
                                        rs[e] = r_ * hprev[j];
                                        g[j] = r_; g[H + j] = u_; g[3 * H + j] = rs[e];
                                    }
                                    *reinterpret_cast<uint4*>(rsb + jb) = make_uint4(pack_bf16x2(rs[0], rs[1]), pack_bf16x2(rs[2], rs[3]),
                                                                                    pack_bf16x2(rs[4], rs[5]), pack_bf16x2(rs[6], rs[7]));
                                }
                            }
                        } else {
#pragma unroll
                            for (int u = 0; u < 32; u += 8) {
                                const int jb = j0 + c0 + u;
                                if (c0 + u < HC && jb < H) {
                                    float hn[8];
#pragma unroll
                                    for (int e = 0; e < 8; ++e) {
                                        const int j = jb + e;
                                        float cand = fast_tanh(v[u + e] + g[2 * H + j]);
                                        float u_ = g[H + j];
                                        hn[e] = fmaf(u_, hprev[j] - cand, cand);     // u*h + (1-u)*cand
                                        g[2 * H + j] = cand;
                                        hnext[j] = hn[e];
                                    }
                                    *reinterpret_cast<uint4*>(hnb + jb) = make_uint4(pack_bf16x2(hn[0], hn[1]), pack_bf16x2(hn[2], hn[3]),
                                                                                    pack_bf16x2(hn[4], hn[5]), pack_bf16x2(hn[6], hn[7]));
                                }
                            }
                        }
                    }
                    tc_fence_before();
                    mbar_arrive(s.z_empty);
                    ++p.z_i;
                }
                rt_publish(a.barrier, epi_tid);
            }
        }
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem, tcols);
    }
}

// ------------------------------------------------------------------------------------------
// backward.  dh_t = dy_t + dz_{t+1} * W_h^T: the B operand is rows H+j0.. of the bf16 copy of
// the cell kernel itself (K = gate columns, contiguous) - no packing needed.
// ------------------------------------------------------------------------------------------
template <int MODEL>
__global__ void __launch_bounds__(RT_THREADS, 1) rec_bwd_tc_kernel(const __grid_constant__ CUtensorMap tmZ0,   // dzb, first contraction
                                                                   const __grid_constant__ CUtensorMap tmZ1,   // gru: dzb candidate columns
                                                                   const __grid_constant__ CUtensorMap tmW0,
                                                                   const __grid_constant__ CUtensorMap tmW1,
                                                                   RecTcArgs a) {
    extern __shared__ uint8_t smem_raw[];
    constexpr int G1 = MODEL == CRNN_MODEL_LSTM ? 4 : (MODEL == CRNN_MODEL_GRU ? 2 : 1);
    constexpr int GZ = MODEL == CRNN_MODEL_LSTM ? 4 : (MODEL == CRNN_MODEL_GRU ? 3 : 1);
    const int B = a.B, H = a.H, HC = a.HC;
    const int nkb0 = ceil_div(G1 * H, 64);
    const int nkb1 = ceil_div(H, 64);
    const RtSmem s = rt_carve(smem_raw, (size_t)nkb0 * HC * 128, MODEL == CRNN_MODEL_GRU ? (size_t)nkb1 * HC * 128 : 0);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int slice = blockIdx.x % a.nslices, tile = blockIdx.x / a.nslices;
    const int j0 = slice * HC;
    const int b0 = tile * a.BT, bt = min(a.BT, B - b0);
    const int64_t gw = (int64_t)a.GS * H, zw = (int64_t)GZ * H, BH = (int64_t)B * H;
    uint32_t tcols = 32;
    while ((int)tcols < HC) tcols <<= 1;

    for (int i = threadIdx.x; i < RT_STAGES * RT_ASTAGE / 16; i += RT_THREADS)
        reinterpret_cast<uint4*>(s.ring)[i] = make_uint4(0, 0, 0, 0);
    fence_proxy_async();
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmZ0); tma_prefetch_desc(&tmW0);
        for (int i = 0; i < RT_STAGES; ++i) { mbar_init(&s.a_full[i], 1); mbar_init(&s.a_empty[i], 1); }
        mbar_init(s.z_full, 1); mbar_init(s.z_empty, 128); mbar_init(s.w_bar, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(s.tmem_slot, tcols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *s.tmem_slot;

    RtPipe p{0, 0};
    unsigned phases_done = 0;

    if (warp == 0) {
        if (lane == 0) {
            uint32_t wbytes = (uint32_t)nkb0 * HC * 128u + (MODEL == CRNN_MODEL_GRU ? (uint32_t)nkb1 * HC * 128u : 0u);
            mbar_expect_tx(s.w_bar, wbytes);
            for (int kb = 0; kb < nkb0; ++kb) tma_load_2d(s.wsa + (size_t)kb * HC * 128, &tmW0, s.w_bar, kb * 64, H + j0);
            if (MODEL == CRNN_MODEL_GRU)
                for (int kb = 0; kb < nkb1; ++kb) tma_load_2d(s.wsb + (size_t)kb * HC * 128, &tmW1, s.w_bar, kb * 64, H + j0);
            for (int t = a.S - 1; t >= 0; --t) {
                // contraction 0 reads dz_{t+1} (all CTAs, previous step)
                if (t + 1 < a.S) {
                    rt_await(a.barrier, phases_done * gridDim.x);
                    for (int ps = 0; ps < a.npass; ++ps)
                        rt_produce(s, p, &tmZ0, (t + 1) * B + b0 + ps * 128, nkb0, a.box_rows);
                }
                ++phases_done;
                if (MODEL == CRNN_MODEL_GRU) {
                    rt_await(a.barrier, phases_done * gridDim.x);
                    for (int ps = 0; ps < a.npass; ++ps)
                        rt_produce(s, p, &tmZ1, t * B + b0 + ps * 128, nkb1, a.box_rows);
                    ++phases_done;
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            mbar_wait(s.w_bar, 0);
            for (int t = a.S - 1; t >= 0; --t) {
                if (t + 1 < a.S)
                    for (int ps = 0; ps < a.npass; ++ps) rt_mma(s, p, tmem, s.wsa, HC, nkb0);
                if (MODEL == CRNN_MODEL_GRU)
                    for (int ps = 0; ps < a.npass; ++ps) rt_mma(s, p, tmem, s.wsb, HC, nkb1);
            }
        }
    } else {
        const int q = warp & 3;
        const int epi_tid = (warp - 2) * 32 + lane;
        const uint32_t tlane = tmem + ((uint32_t)(32 * q) << 16);
        float v[32];
        // zero the CTA-private carries
        if (MODEL != CRNN_MODEL_RNN) {
            for (int ps = 0; ps < a.npass; ++ps) {
                const int r = ps * 128 + 32 * q + lane;
                if (r < bt)
                    for (int jj = 0; jj < HC && j0 + jj < H; ++jj) a.carry0[(int64_t)(b0 + r) * H + j0 + jj] = 0.f;
            }
        }
        for (int t = a.S - 1; t >= 0; --t) {
            const bool has_next = t + 1 < a.S;
            for (int ps = 0; ps < a.npass; ++ps) {
                const int r = ps * 128 + 32 * q + lane;
                const bool row_ok = r < bt;
                const int64_t b = b0 + r;
                const int64_t o0 = b * H;
                const float* g = a.gates + ((int64_t)t * B + b) * gw;
                float* dz = a.dz + ((int64_t)t * B + b) * zw;
                __nv_bfloat16* dzb = a.dzb + ((int64_t)t * B + b) * zw;
                const float* dyt = a.dy + (int64_t)t * BH + o0;
                if (has_next) { mbar_wait(s.z_full, p.z_i & 1); tc_fence_after(); }
                for (int c0 = 0; c0 < HC; c0 += 32) {
                    if (has_next) tmem_ld32(tlane + (uint32_t)c0, v);
                    if (!row_ok) continue;
#pragma unroll
                    for (int u = 0; u < 32; u += 8) {
                        const int jb = j0 + c0 + u;
                        if (c0 + u >= HC || jb >= H) continue;
                        float o_a[8], o_b[8], o_c[8], o_d[8];
#pragma unroll
                        for (int e = 0; e < 8; ++e) {
                            const int j = jb + e;
                            float dht = dyt[j] + (has_next ? v[u + e] : 0.f);
                            if (MODEL == CRNN_MODEL_LSTM) {
                                float si = g[j], tj = g[H + j], sf = g[2 * H + j], so = g[3 * H + j];
                                float cp = a.cx[(int64_t)t * BH + o0 + j];
                                float tc_ = fast_tanh(a.cx[(int64_t)(t + 1) * BH + o0 + j]);
                                float dct = a.carry0[o0 + j] + dht * so * (1.0f - tc_ * tc_);
                                o_a[e] = dct * tj * si * (1.0f - si);
                                o_b[e] = dct * si * (1.0f - tj * tj);
                                o_c[e] = dct * cp * sf * (1.0f - sf);
                                o_d[e] = dht * tc_ * so * (1.0f - so);
                                a.carry0[o0 + j] = dct * sf;
                                dz[j] = o_a[e]; dz[H + j] = o_b[e]; dz[2 * H + j] = o_c[e]; dz[3 * H + j] = o_d[e];
                            } else if (MODEL == CRNN_MODEL_RNN) {
                                float hn = a.hx[(int64_t)(t + 1) * BH + o0 + j];
                                o_a[e] = dht * (1.0f - hn * hn);
                                dz[j] = o_a[e];
                            } else {
                                dht += a.carry0[o0 + j];
                                float u_ = g[H + j], cand = g[2 * H + j];
                                float hp = a.hx[(int64_t)t * BH + o0 + j];
                                o_c[e] = dht * (1.0f - u_) * (1.0f - cand * cand);      // d candidate pre-activation
                                o_b[e] = dht * (hp - cand) * u_ * (1.0f - u_);          // d update-gate pre-activation
                                dz[2 * H + j] = o_c[e]; dz[H + j] = o_b[e];
                                a.carry1[o0 + j] = dht;
                            }
                        }
#define CRNN_ST8(ptr, arr) *reinterpret_cast<uint4*>(ptr) = make_uint4(pack_bf16x2(arr[0], arr[1]), pack_bf16x2(arr[2], arr[3]), \
                                                                       pack_bf16x2(arr[4], arr[5]), pack_bf16x2(arr[6], arr[7]))
                        if (MODEL == CRNN_MODEL_LSTM) {
                            CRNN_ST8(dzb + jb, o_a); CRNN_ST8(dzb + H + jb, o_b); CRNN_ST8(dzb + 2 * H + jb, o_c); CRNN_ST8(dzb + 3 * H + jb, o_d);
                        } else if (MODEL == CRNN_MODEL_RNN) {
                            CRNN_ST8(dzb + jb, o_a);
                        } else {
                            CRNN_ST8(dzb + 2 * H + jb, o_c); CRNN_ST8(dzb + H + jb, o_b);
                        }
                    }
                }
                if (has_next) { tc_fence_before(); mbar_arrive(s.z_empty); ++p.z_i; }
            }
            rt_publish(a.barrier, epi_tid);
            if (MODEL == CRNN_MODEL_GRU) {
                for (int ps = 0; ps < a.npass; ++ps) {
                    const int r = ps * 128 + 32 * q + lane;
                    const bool row_ok = r < bt;
                    const int64_t b = b0 + r;
                    const int64_t o0 = b * H;
                    const float* g = a.gates + ((int64_t)t * B + b) * gw;
                    float* dz = a.dz + ((int64_t)t * B + b) * zw;
                    __nv_bfloat16* dzb = a.dzb + ((int64_t)t * B + b) * zw;
                    mbar_wait(s.z_full, p.z_i & 1);
                    tc_fence_after();
                    for (int c0 = 0; c0 < HC; c0 += 32) {
                        tmem_ld32(tlane + (uint32_t)c0, v);
                        if (!row_ok) continue;
#pragma unroll
                        for (int u = 0; u < 32; u += 8) {
                            const int jb = j0 + c0 + u;
                            if (c0 + u >= HC || jb >= H) continue;
                            float o_a[8];
#pragma unroll
                            for (int e = 0; e < 8; ++e) {
                                const int j = jb + e;
                                float r_ = g[j], u_ = g[H + j];
                                float hp = a.hx[(int64_t)t * BH + o0 + j];
                                float drs = v[u + e];
                                o_a[e] = drs * hp * r_ * (1.0f - r_);
                                dz[j] = o_a[e];
                                a.carry0[o0 + j] = fmaf(a.carry1[o0 + j], u_, drs * r_);
                            }
                            CRNN_ST8(dzb + jb, o_a);
                        }
                    }
                    tc_fence_before();
                    mbar_arrive(s.z_empty);
                    ++p.z_i;
                }
                rt_publish(a.barrier, epi_tid);
            }
#undef CRNN_ST8
        }
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem, tcols);
    }
}

#undef fast_tanh
#undef fast_sigmoid

// Pack the recurrent halves of the cell kernels for the forward kernel (bf16, K-major rows).
//   out[(slice*GW*HC + lc) * Kp + k]:  first G1*HC rows lc = jj*G1 + g -> W[(H+k), g*H + j0+jj];
//   GRU: rows 2*HC + jj -> Wc[(H+k), j0+jj].  Rows of units past H and k >= H are zero.
__global__ void __launch_bounds__(256) pack_rec_weights_kernel(const float* __restrict__ W, const float* __restrict__ Wc,
                                                               int H, int HC, int nslices, int G1, int GW, int Kp,
                                                               __nv_bfloat16* __restrict__ out) {
    const int64_t n = (int64_t)nslices * GW * HC * Kp;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int k = (int)(i % Kp);
        const int64_t row = i / Kp;
        const int lc = (int)(row % (GW * HC));
        const int slice = (int)(row / (GW * HC));
        float v = 0.f;
        if (k < H) {
            if (lc < G1 * HC) {
                const int jj = lc / G1, g = lc % G1, j = slice * HC + jj;
                if (j < H) v = W[(int64_t)(H + k) * (G1 * H) + (int64_t)g * H + j];
            } else {
                const int j = slice * HC + (lc - G1 * HC);
                if (j < H) v = Wc[(int64_t)(H + k) * H + j];
            }
        }
        out[i] = __float2bfloat16_rn(v);
    }
}

}  // namespace tc
}  // namespace crnn
