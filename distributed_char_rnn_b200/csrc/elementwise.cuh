// elementwise.cuh — the HBM-bound pieces of the step: embedding gather / scatter-add,
// softmax + sequence cross-entropy rows, global-norm clip + Adam.
#pragma once
#include "common.cuh"

namespace crnn {

// ---- embedding (model.py:54-55,62-63) ---------------------------------------------------
// out[t,b,:] = table[x[b,t],:]; one float4 per thread, rows of the (tiny, L2-resident) table.
__global__ void __launch_bounds__(256) embed_gather_kernel(const float* __restrict__ table, int V, int H,
                                                           const int32_t* __restrict__ x, int64_t ldx,
                                                           int B, int S, float* __restrict__ out,
                                                           __nv_bfloat16* __restrict__ out_bf16, int64_t lo_ofs = 0) {
    const int64_t n = (int64_t)S * B * H;
    const bool vec = (H & 3) == 0;
    if (vec) {
        const int H4 = H >> 2;
        const int64_t n4 = n >> 2;
        for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
            int64_t row = i / H4;
            int c = (int)(i % H4);
            int t = (int)(row / B), b = (int)(row % B);
            int id = __ldg(x + (int64_t)b * ldx + t);
            id = min(max(id, 0), V - 1);
            float4 v = __ldg(reinterpret_cast<const float4*>(table + (int64_t)id * H) + c);
            reinterpret_cast<float4*>(out)[i] = v;
            if (out_bf16) {
                const __nv_bfloat162 h0 = __floats2bfloat162_rn(v.x, v.y), h1 = __floats2bfloat162_rn(v.z, v.w);
                reinterpret_cast<__nv_bfloat162*>(out_bf16)[2 * i] = h0;
                reinterpret_cast<__nv_bfloat162*>(out_bf16)[2 * i + 1] = h1;
                if (lo_ofs) {                       // lo_ofs is a multiple of 4: the twin keeps the 8-byte alignment
                    const float2 f0 = __bfloat1622float2(h0), f1 = __bfloat1622float2(h1);
                    reinterpret_cast<__nv_bfloat162*>(out_bf16 + lo_ofs)[2 * i] = __floats2bfloat162_rn(v.x - f0.x, v.y - f0.y);
                    reinterpret_cast<__nv_bfloat162*>(out_bf16 + lo_ofs)[2 * i + 1] = __floats2bfloat162_rn(v.z - f1.x, v.w - f1.y);
                }
            }
        }
    } else {
        for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
            int64_t row = i / H;
            int c = (int)(i % H);
            int t = (int)(row / B), b = (int)(row % B);
            int id = __ldg(x + (int64_t)b * ldx + t);
            id = min(max(id, 0), V - 1);
            out[i] = __ldg(table + (int64_t)id * H + c);
            if (out_bf16) store_bf16_split(out_bf16 + i, lo_ofs, out[i]);
        }
    }
}

// dtable[x[b,t],:] += dout[t,b,:]  (IndexedSlices -> dense) and sqnorm += sum dout^2.
// Atomic form: used by the single-operator entry point (no workspace) and for vocabularies whose rows do not fit the
// shared-memory table of the ordered form below.  The summation order of the atomics is not fixed.
__global__ void __launch_bounds__(256) embed_scatter_kernel(float* __restrict__ dtable, int V, int H,
                                                            const int32_t* __restrict__ x, int64_t ldx,
                                                            int B, int S, const float* __restrict__ dout,
                                                            double* __restrict__ sqnorm) {
    __shared__ double sh[32];
    const int64_t n = (int64_t)S * B * H;
    double ss = 0.0;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        int64_t row = i / H;
        int c = (int)(i % H);
        int t = (int)(row / B), b = (int)(row % B);
        int id = __ldg(x + (int64_t)b * ldx + t);
        id = min(max(id, 0), V - 1);
        float g = __ldg(dout + i);
        ss += (double)g * (double)g;
        atomicAdd(dtable + (int64_t)id * H + c, g);
    }
    double tot = block_sum_d(ss, sh);
    if (threadIdx.x == 0 && sqnorm) atomicAdd(sqnorm, tot);
}

// Ordered form (the training step): no atomics, the same bits on every run and on every replica.
// Pass 1, CTA (column chunk, token segment): a [V][128] table in shared memory; the segment's rows (time-major order
// t*B + b) are added one after the other, thread c owning column c of the chunk - a fixed order, no conflicts, the
// hot rows never leave the SM.  The rows stream through a cp.async ring (ES_STAGES x 32 rows x 128 columns), so the
// serial part of the loop is the shared-memory read-add-write of one table entry (~35 cycles per row) and not a DRAM
// round trip.  The table goes to part[seg][V][H]; the chunk's share of sum dout^2 to sqpart.
// Pass 2: dtable[v][c] = sum over segments in order; block 0 also sums sqpart in order into *sqnorm.
constexpr int ES_COLS = 128;
constexpr int ES_ROWS = 32;          // rows per ring stage
constexpr int ES_STAGES = 4;
inline size_t es_smem_bytes(int V) { return (size_t)V * ES_COLS * 4 + (size_t)ES_STAGES * ES_ROWS * ES_COLS * 4; }
__global__ void __launch_bounds__(ES_COLS) embed_scatter_part_kernel(int V, int H, const int32_t* __restrict__ x, int64_t ldx, int B, int S,
                                                                     const float* __restrict__ dout, int tokens_per_seg,
                                                                     float* __restrict__ part, double* __restrict__ sqpart) {
    extern __shared__ __align__(16) float es_smem[];
    float* es_tab = es_smem;                          // [V][ES_COLS]
    float* ring = es_smem + V * ES_COLS;              // [ES_STAGES][ES_ROWS][ES_COLS]
    __shared__ double sh[32];
    const int tid = threadIdx.x, lane = tid & 31;
    const int c0 = blockIdx.x * ES_COLS;
    const int c = c0 + tid;
    const bool live = c < H;
    const int64_t N = (int64_t)S * B;
    const int64_t r0 = (int64_t)blockIdx.y * tokens_per_seg, r1 = min(N, r0 + tokens_per_seg);
    const int nrows = (int)(r1 - r0), nstage = (nrows + ES_ROWS - 1) / ES_ROWS;
    for (int i = tid; i < V * ES_COLS; i += ES_COLS) es_tab[i] = 0.f;
    // copy role: thread -> (row in stage = tid / 32 + 4k, 16-byte piece = tid % 32); H % 4 == 0 (host)
    auto issue = [&](int sidx) {
        if (sidx < nstage) {
            float* dst = ring + (sidx % ES_STAGES) * (ES_ROWS * ES_COLS);
#pragma unroll
            for (int k = 0; k < ES_ROWS / 4; ++k) {
                const int rr = (tid >> 5) + 4 * k;
                const int64_t row = r0 + (int64_t)sidx * ES_ROWS + rr;
                const int col = c0 + 4 * lane;
                if (row < r1 && col < H) {
                    const unsigned sa = (unsigned)__cvta_generic_to_shared(dst + rr * ES_COLS + 4 * lane);
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(dout + row * H + col) : "memory");
                }
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
#pragma unroll
    for (int i = 0; i < ES_STAGES - 1; ++i) issue(i);
    double ss = 0.0;
    auto load_id = [&](int sidx) -> int {             // a stage's ids: lane l holds row l's
        const int64_t myrow = r0 + (int64_t)sidx * ES_ROWS + lane;
        if (myrow >= r1) return 0;
        const int t = (int)(myrow / B), b = (int)(myrow % B);
        return min(max(__ldg(x + (int64_t)b * ldx + t), 0), V - 1);
    };
    int nextid = load_id(0);
    for (int sidx = 0; sidx < nstage; ++sidx) {
        issue(sidx + ES_STAGES - 1);                  // (overwrites the stage consumed in the previous iteration)
        const int myid = nextid;
        nextid = load_id(sidx + 1);                   // in flight while this stage is consumed
        asm volatile("cp.async.wait_group %0;" ::"n"(ES_STAGES - 1) : "memory");
        __syncthreads();                              // every thread's pieces of the stage have landed
        const float* src = ring + (sidx % ES_STAGES) * (ES_ROWS * ES_COLS) + tid;
        const int nr = min(ES_ROWS, nrows - sidx * ES_ROWS);
#pragma unroll 4
        for (int rr = 0; rr < nr; ++rr) {             // (all lanes take part in the shuffle, also those past column H)
            const int id = __shfl_sync(0xffffffffu, myid, rr);
            if (live) {
                const float g = src[rr * ES_COLS];
                es_tab[id * ES_COLS + tid] += g;
                ss += (double)g * (double)g;
            }
        }
        __syncthreads();                              // the stage may be overwritten
    }
    if (live) {
        float* dst = part + ((int64_t)blockIdx.y * V) * H + c;
        for (int v = 0; v < V; ++v) dst[(int64_t)v * H] = es_tab[v * ES_COLS + tid];
    }
    const double tot = block_sum_d(ss, sh);
    if (tid == 0) sqpart[blockIdx.y * gridDim.x + blockIdx.x] = tot;
}
__global__ void __launch_bounds__(256) embed_scatter_reduce_kernel(float* __restrict__ dtable, int64_t VH, int nseg,
                                                                   const float* __restrict__ part, const double* __restrict__ sqpart,
                                                                   int nsq, double* __restrict__ sqnorm) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < VH; i += (int64_t)gridDim.x * blockDim.x) {
        float acc = 0.f;
        int sg = 0;
        for (; sg + 8 <= nseg; sg += 8) {             // eight loads in flight, added in segment order
            float v[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = __ldg(part + (int64_t)(sg + k) * VH + i);
#pragma unroll
            for (int k = 0; k < 8; ++k) acc += v[k];
        }
        for (; sg < nseg; ++sg) acc += __ldg(part + (int64_t)sg * VH + i);
        dtable[i] = acc;
    }
    if (blockIdx.x == 0 && sqnorm) {                  // fixed order: strided per thread, then the block's fixed tree
        __shared__ double sh[32];
        double t = 0.0;
        for (int i = threadIdx.x; i < nsq; i += blockDim.x) t += sqpart[i];
        t = block_sum_d(t, sh);
        if (threadIdx.x == 0) *sqnorm = t;
    }
}

// Sorted form of the ordered scatter (the default of the training step): the token ids are known when the step starts, so
// the tokens are first sorted stably by id (three small kernels that run beside the forward pass), which turns the scatter
// into a segmented sum over contiguous runs of the index array:
//   es_rank_kernel     block of 1024 tokens (time-major order i = t*B + b): per-block histogram, and for every token the
//                      number of EARLIER tokens of its block with the same id;
//   es_offsets_kernel  one block: exclusive sums over blocks (per id) and over ids, and the list of "pieces" - runs of at
//                      most 64 (or 32) tokens of one id;
//   es_place_kernel    order[offset(id) + base(block, id) + rank] = i;
//   es_sum_kernel      CTA (piece, 512-column chunk): thread c adds column quad c of the piece's rows in a register, in
//                      time order - independent 16-byte loads, no shared memory, no atomics, every byte of dout read once;
//   es_finish_kernel   dtable[v] = sum of v's pieces in order; block 0 adds the pieces' sums of squares in order.
// Same bits on every run and every replica: every sum has a fixed order that does not depend on the grid.
constexpr int ES_PIECE_MAX = 64;     // tokens per piece: 64, or 32 where that is needed to fill the GPU (es_sum_kernel<PIECE>)
constexpr int ES_SORT_BLOCK = 1024;
__global__ void __launch_bounds__(ES_SORT_BLOCK) es_rank_kernel(const int32_t* __restrict__ x, int64_t ldx, int B, int64_t N, int V,
                                                                int* __restrict__ hist, int* __restrict__ lrank) {
    __shared__ int ids[ES_SORT_BLOCK];
    extern __shared__ int es_hist_s[];                 // [V]
    const int tid = threadIdx.x;
    const int64_t i = (int64_t)blockIdx.x * ES_SORT_BLOCK + tid;
    for (int v = tid; v < V; v += ES_SORT_BLOCK) es_hist_s[v] = 0;
    int id = -1;
    if (i < N) {
        const int t = (int)(i / B), b = (int)(i % B);
        id = min(max(__ldg(x + (int64_t)b * ldx + t), 0), V - 1);
    }
    ids[tid] = id;
    __syncthreads();
    if (i < N) {
        atomicAdd(&es_hist_s[id], 1);                  // (a count: the order of the adds does not matter)
        int r = 0;
        for (int j = 0; j < tid; ++j) r += ids[j] == id;
        lrank[i] = r;
    }
    __syncthreads();
    for (int v = tid; v < V; v += ES_SORT_BLOCK) hist[(int64_t)blockIdx.x * V + v] = es_hist_s[v];
}
// hist[blk][v] becomes the number of tokens with id v in EARLIER blocks; off[v] the first position of id v in `order`;
// poff[v] the first piece of id v; piece[p] = {k0, k1} positions in `order`; *npieces their number.
__global__ void __launch_bounds__(1024) es_offsets_kernel(int nblk, int V, int ES_PIECE, int* __restrict__ hist, int* __restrict__ off,
                                                          int* __restrict__ poff, int2* __restrict__ piece, int* __restrict__ npieces) {
    extern __shared__ int es_cnt[];                    // [V] counts, then [V] piece counts
    int* cnt = es_cnt;
    int* pc = es_cnt + V;
    for (int v = threadIdx.x; v < V; v += blockDim.x) {
        int run = 0;
        for (int b = 0; b < nblk; ++b) { const int t = hist[(int64_t)b * V + v]; hist[(int64_t)b * V + v] = run; run += t; }
        cnt[v] = run;
        pc[v] = (run + ES_PIECE - 1) / ES_PIECE;
    }
    __syncthreads();
    if (threadIdx.x == 0) {                            // V is a vocabulary of characters: a serial scan is a few hundred adds
        int o = 0, p = 0;
        for (int v = 0; v < V; ++v) { off[v] = o; poff[v] = p; o += cnt[v]; p += pc[v]; }
        off[V] = o; poff[V] = p;
        *npieces = p;
    }
    __syncthreads();
    for (int v = threadIdx.x; v < V; v += blockDim.x) {
        const int o = off[v], n = cnt[v], p0 = poff[v];
        for (int j = 0; j < pc[v]; ++j) piece[p0 + j] = make_int2(o + j * ES_PIECE, o + min(n, (j + 1) * ES_PIECE));
    }
}
__global__ void __launch_bounds__(256) es_place_kernel(const int32_t* __restrict__ x, int64_t ldx, int B, int64_t N, int V,
                                                       const int* __restrict__ hist, const int* __restrict__ off,
                                                       const int* __restrict__ lrank, int* __restrict__ order) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < N; i += (int64_t)gridDim.x * blockDim.x) {
        const int t = (int)(i / B), b = (int)(i % B);
        const int id = min(max(__ldg(x + (int64_t)b * ldx + t), 0), V - 1);
        order[off[id] + hist[(i / ES_SORT_BLOCK) * V + id] + lrank[i]] = (int)i;
    }
}
constexpr int ES_SUM_THREADS = 128;  // x 4 columns
template <int ES_PIECE>
__global__ void __launch_bounds__(ES_SUM_THREADS) es_sum_kernel(int H, const float* __restrict__ dout, const int* __restrict__ order,
                                                                const int2* __restrict__ piece, const int* __restrict__ npieces,
                                                                float* __restrict__ partial, double* __restrict__ sqpart) {
    __shared__ double sh[32];
    const int p = blockIdx.y;
    if (p >= *npieces) return;
    const int col = (blockIdx.x * ES_SUM_THREADS + threadIdx.x) * 4;
    const bool live = col < H;                         // H % 4 == 0 (host)
    const int2 pc = piece[p];
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    double ss = 0.0;
    if (live) {
        const float* src = dout + col;
        if (ES_PIECE == 32) {
            // short pieces (small inputs: latency matters): the piece's row indices first - one round trip for all of them -,
            // then its rows in two batches of sixteen independent 16-byte loads, added in time order
            int idx[32];
            const int n = pc.y - pc.x;
#pragma unroll
            for (int u = 0; u < 32; ++u) idx[u] = u < n ? __ldg(order + pc.x + u) : -1;
#pragma unroll
            for (int h0 = 0; h0 < 32; h0 += 16) {
                float4 g[16];
#pragma unroll
                for (int u = 0; u < 16; ++u)
                    g[u] = idx[h0 + u] >= 0 ? __ldg(reinterpret_cast<const float4*>(src + (int64_t)idx[h0 + u] * H)) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int u = 0; u < 16; ++u) {
                    acc.x += g[u].x; acc.y += g[u].y; acc.z += g[u].z; acc.w += g[u].w;
                    ss += (double)g[u].x * g[u].x + (double)g[u].y * g[u].y + (double)g[u].z * g[u].z + (double)g[u].w * g[u].w;
                }
            }
        } else {
            // long pieces (large inputs: bandwidth matters, many CTAs per SM hide the latency): eight rows in flight
            int k = pc.x;
            for (; k + 8 <= pc.y; k += 8) {
                float4 g[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) g[u] = __ldg(reinterpret_cast<const float4*>(src + (int64_t)__ldg(order + k + u) * H));
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    acc.x += g[u].x; acc.y += g[u].y; acc.z += g[u].z; acc.w += g[u].w;
                    ss += (double)g[u].x * g[u].x + (double)g[u].y * g[u].y + (double)g[u].z * g[u].z + (double)g[u].w * g[u].w;
                }
            }
            for (; k < pc.y; ++k) {
                const float4 g = __ldg(reinterpret_cast<const float4*>(src + (int64_t)__ldg(order + k) * H));
                acc.x += g.x; acc.y += g.y; acc.z += g.z; acc.w += g.w;
                ss += (double)g.x * g.x + (double)g.y * g.y + (double)g.z * g.z + (double)g.w * g.w;
            }
        }
        *reinterpret_cast<float4*>(partial + (int64_t)p * H + col) = acc;
    }
    const double tot = block_sum_d(ss, sh);
    if (threadIdx.x == 0) sqpart[(int64_t)p * gridDim.x + blockIdx.x] = tot;
}
__global__ void __launch_bounds__(256) es_finish_kernel(float* __restrict__ dtable, int V, int H, const int* __restrict__ poff,
                                                        const float* __restrict__ partial, const double* __restrict__ sqpart,
                                                        const int* __restrict__ npieces, int nchunk, double* __restrict__ sqnorm) {
    const int64_t VH = (int64_t)V * H;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < VH; i += (int64_t)gridDim.x * blockDim.x) {
        const int v = (int)(i / H), c = (int)(i % H);
        float acc = 0.f;
        int p = poff[v];
        const int p1 = poff[v + 1];
        for (; p + 8 <= p1; p += 8) {                  // eight loads in flight, added in piece order
            float g[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) g[u] = __ldg(partial + (int64_t)(p + u) * H + c);
#pragma unroll
            for (int u = 0; u < 8; ++u) acc += g[u];
        }
        for (; p < p1; ++p) acc += __ldg(partial + (int64_t)p * H + c);
        dtable[i] = acc;
    }
    if (blockIdx.x == 0 && sqnorm) {                  // fixed order: strided per thread, then the block's fixed tree
        __shared__ double sh[32];
        const int nsq = *npieces * nchunk;
        double t = 0.0;
        for (int i = threadIdx.x; i < nsq; i += blockDim.x) t += sqpart[i];
        t = block_sum_d(t, sh);
        if (threadIdx.x == 0) *sqnorm = t;
    }
}

// ---- dropout (model.py:31-34,58-59; train.py:58-61) ----------------------------------------
// tf.nn.dropout: x * floor(keep + u) / keep, a fresh mask per element and per step.  Masks are not
// stored: they are a pure function of (seed, training-call index, stage, element index), regenerated
// in the backward pass.  Stages: 0 = embedding, 1+2l = input of cell l, 2+2l = output of cell l.
// A layer input carries the product of two masks (what produced it, and the cell's input wrapper).
__device__ __forceinline__ float dropout_mult(uint64_t seed, uint32_t stage, uint64_t idx, float keep) {
    if (keep >= 1.0f) return 1.0f;
    uint64_t z = seed + 0x9E3779B97F4A7C15ull * (uint64_t)(stage + 1) + idx * 0xD1B54A32D192ED03ull;
    z ^= z >> 33; z *= 0xFF51AFD7ED558CCDull; z ^= z >> 33; z *= 0xC4CEB9FE1A85EC53ull; z ^= z >> 33;
    const float u = (float)(uint32_t)(z >> 40) * (1.0f / 16777216.0f);        // uniform [0,1)
    return u < keep ? 1.0f / keep : 0.0f;
}

struct DropArgs {
    uint64_t seed;
    uint32_t stage_a, stage_b;
    float keep_a, keep_b;
    int64_t idx0;          // global element index of src[0] inside the [S,B,H] tensor (chunked calls)
};

// dst = src * mask_a * mask_b (dst may equal src); optional bf16 copy.
__global__ void __launch_bounds__(256) dropout_apply_kernel(const float* __restrict__ src, float* __restrict__ dst,
                                                            __nv_bfloat16* __restrict__ dst_bf16, int64_t n, DropArgs d,
                                                            int64_t lo_ofs = 0) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const uint64_t gi = (uint64_t)(d.idx0 + i);
        const float v = (src ? src[i] : 1.0f) * dropout_mult(d.seed, d.stage_a, gi, d.keep_a) * dropout_mult(d.seed, d.stage_b, gi, d.keep_b);
        if (dst) dst[i] = v;
        if (dst_bf16) store_bf16_split(dst_bf16 + i, lo_ofs, v);
    }
}

// ---- softmax + sequence cross-entropy (model.py:77,80-85) ---------------------------------
// One warp per row of logits [N, V] (row i = t*B + b).  In place: logits -> (softmax-onehot)/N.
__global__ void __launch_bounds__(256) softmax_xent_kernel(float* __restrict__ logits, int B, int S, int V,
                                                           const int32_t* __restrict__ y, int64_t ldy,
                                                           double* __restrict__ loss_sum,
                                                           __nv_bfloat16* __restrict__ dl_bf16, int64_t ld_bf16,
                                                           int64_t lo_ofs = 0, double* __restrict__ loss_part = nullptr) {
    __shared__ double sh[32];
    const int lane = threadIdx.x & 31;
    const int64_t N = (int64_t)B * S;
    const float invN = 1.0f / (float)N;
    double local = 0.0;
    for (int64_t row = blockIdx.x * 8 + (threadIdx.x >> 5); row < N; row += (int64_t)gridDim.x * 8) {
        float* z = logits + row * V;
        int t = (int)(row / B), b = (int)(row % B);
        int label = __ldg(y + (int64_t)b * ldy + t);
        float m = -INFINITY;
        for (int c = lane; c < V; c += 32) m = fmaxf(m, z[c]);
        m = warp_max(m);
        float s = 0.f;
        for (int c = lane; c < V; c += 32) s += expf(z[c] - m);
        s = warp_sum(s);
        float lse = logf(s) + m;
        float inv = 1.0f / s;
        float zl = 0.f;
        for (int c = lane; c < V; c += 32) {
            float v = z[c];
            if (c == label) zl = v;
            float p = expf(v - m) * inv;
            const float d = (p - (c == label ? 1.0f : 0.0f)) * invN;
            z[c] = d;
            if (dl_bf16) store_bf16_split(dl_bf16 + row * ld_bf16 + c, lo_ofs, d);
        }
        zl = warp_sum(zl);
        if (lane == 0) local += (double)(lse - zl);
    }
    double tot = block_sum_d(local, sh);
    // the training step hands in one slot per block and sums them in block order (finalize_scalars_kernel): same bits every run
    if (threadIdx.x == 0 && loss_part) loss_part[blockIdx.x] = tot;
    else if (threadIdx.x == 0 && loss_sum) atomicAdd(loss_sum, tot);
}

// probs[b*S+t,:] = softmax(logits_tb[t*B+b,:]); also writes the re-ordered logits.
__global__ void __launch_bounds__(256) logits_out_kernel(const float* __restrict__ logits_tb, int B, int S, int V,
                                                         float* __restrict__ logits_bs, float* __restrict__ probs_bs) {
    const int lane = threadIdx.x & 31;
    const int64_t N = (int64_t)B * S;
    for (int64_t row = blockIdx.x * 8 + (threadIdx.x >> 5); row < N; row += (int64_t)gridDim.x * 8) {
        const float* z = logits_tb + row * V;
        int t = (int)(row / B), b = (int)(row % B);
        int64_t orow = (int64_t)b * S + t;
        float m = -INFINITY;
        for (int c = lane; c < V; c += 32) m = fmaxf(m, z[c]);
        m = warp_max(m);
        float s = 0.f;
        for (int c = lane; c < V; c += 32) s += expf(z[c] - m);
        s = warp_sum(s);
        float inv = 1.0f / s;
        for (int c = lane; c < V; c += 32) {
            float v = z[c];
            if (logits_bs) logits_bs[orow * V + c] = v;
            if (probs_bs) probs_bs[orow * V + c] = expf(v - m) * inv;
        }
    }
}

// grads[P] = sum |dX|^2 ; grads[P+1] = cost = loss_sum / (B*S)
__global__ void finalize_scalars_kernel(const double* __restrict__ sq, const double* __restrict__ loss_sum,
                                        double n, float* __restrict__ tail, const double* __restrict__ loss_part = nullptr,
                                        int nloss = 0) {
    __shared__ double sh[32];
    double l = 0.0;                                   // fixed order: strided per thread, then the block's fixed tree
    for (int i = threadIdx.x; i < nloss; i += blockDim.x) l += loss_part[i];
    l = block_sum_d(l, sh);
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        tail[0] = (float)(*sq);
        tail[1] = (float)((l + *loss_sum) / n);
    }
}

// ---- clip_by_global_norm + Adam (model.py:91-98) ------------------------------------------
constexpr int NORM_BLOCKS = 592;   // 4 x 148

// partial[blk] = sum of g^2 over this block's share of [0,P) minus the embedding range
// (the embedding enters the norm through its un-deduplicated slices, grads[P]).
__global__ void __launch_bounds__(256) sqnorm_partials_kernel(const float* __restrict__ g, int64_t P,
                                                              int64_t skip_beg, int64_t skip_end,
                                                              double* __restrict__ partial) {
    __shared__ double sh[32];
    double ss = 0.0;
    const int64_t P4 = P >> 2;   // P is a multiple of 64
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < P4; i += (int64_t)gridDim.x * blockDim.x) {
        int64_t e = i << 2;
        float4 v = __ldg(reinterpret_cast<const float4*>(g) + i);
        if (e + 3 < skip_beg || e >= skip_end) {
            ss += (double)v.x * v.x + (double)v.y * v.y + (double)v.z * v.z + (double)v.w * v.w;
        } else {
            if (e + 0 < skip_beg || e + 0 >= skip_end) ss += (double)v.x * v.x;
            if (e + 1 < skip_beg || e + 1 >= skip_end) ss += (double)v.y * v.y;
            if (e + 2 < skip_beg || e + 2 >= skip_end) ss += (double)v.z * v.z;
            if (e + 3 < skip_beg || e + 3 >= skip_end) ss += (double)v.w * v.w;
        }
    }
    double tot = block_sum_d(ss, sh);
    if (threadIdx.x == 0) partial[blockIdx.x] = tot;
}

struct AdamArgs {
    float* p; const float* g; float* m; float* v;
    __nv_bfloat16* p_bf16;       // optional bf16 shadow of the parameters (perf mode)
    int64_t P;
    const double* partial; int npartial;
    float lr_t;                  // lr * sqrt(1-beta2^t) / (1-beta1^t)
    float beta1, beta2, eps;
    float clip, mult;
    float* gnorm_out;
};

__global__ void __launch_bounds__(256) clip_adam_kernel(AdamArgs a) {
    __shared__ double sh[32];
    __shared__ float s_scale;
    // every block reduces the partials in the same fixed order -> identical scale everywhere
    double ss = 0.0;
    for (int i = threadIdx.x; i < a.npartial; i += blockDim.x) ss += a.partial[i];
    double tot = block_sum_d(ss, sh);
    if (threadIdx.x == 0) {
        tot += (double)a.g[a.P];                                   // slices' sum of squares
        float gn = a.mult * (float)sqrt(tot);
        // tf.clip_by_global_norm: scale = clip * min(1/norm, 1/clip)
        s_scale = a.clip * fminf(1.0f / gn, 1.0f / a.clip);
        if (blockIdx.x == 0 && a.gnorm_out) *a.gnorm_out = gn;
    }
    __syncthreads();
    const float gs = s_scale * a.mult;
    const float omb1 = 1.0f - a.beta1, omb2 = 1.0f - a.beta2;
    const int64_t P4 = a.P >> 2;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < P4; i += (int64_t)gridDim.x * blockDim.x) {
        float4 g = __ldg(reinterpret_cast<const float4*>(a.g) + i);
        float4 m = reinterpret_cast<float4*>(a.m)[i];
        float4 v = reinterpret_cast<float4*>(a.v)[i];
        float4 p = reinterpret_cast<float4*>(a.p)[i];
        float gg[4] = {g.x * gs, g.y * gs, g.z * gs, g.w * gs};
        float mm[4] = {m.x, m.y, m.z, m.w}, vv[4] = {v.x, v.y, v.z, v.w}, pp[4] = {p.x, p.y, p.z, p.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            // ApplyAdam: m += (g-m)(1-b1); v += (g^2-v)(1-b2); p -= lr_t*m / (sqrt(v)+eps)
            mm[k] += (gg[k] - mm[k]) * omb1;
            vv[k] += (gg[k] * gg[k] - vv[k]) * omb2;
            pp[k] -= (a.lr_t * mm[k]) / (sqrtf(vv[k]) + a.eps);
        }
        reinterpret_cast<float4*>(a.m)[i] = make_float4(mm[0], mm[1], mm[2], mm[3]);
        reinterpret_cast<float4*>(a.v)[i] = make_float4(vv[0], vv[1], vv[2], vv[3]);
        reinterpret_cast<float4*>(a.p)[i] = make_float4(pp[0], pp[1], pp[2], pp[3]);
        if (a.p_bf16) {
            __nv_bfloat162 lo = __floats2bfloat162_rn(pp[0], pp[1]);
            __nv_bfloat162 hi = __floats2bfloat162_rn(pp[2], pp[3]);
            reinterpret_cast<__nv_bfloat162*>(a.p_bf16)[2 * i] = lo;
            reinterpret_cast<__nv_bfloat162*>(a.p_bf16)[2 * i + 1] = hi;
        }
    }
}

}  // namespace crnn
