// recurrence_tc.cuh — persistent tensor-core recurrence / BPTT kernels (CRNN_PREC_BF16).
//
// The serial part of the reference's unrolled MultiRNNCell (model.py:72) and of its
// reverse-mode mirror (model.py:88-91): per time step only h_{t-1} * W_h (forward) or
// dz_{t+1} * W_h^T (backward) remains once the x halves are hoisted into batch GEMMs.
//
// One launch per layer, direction and sequence chunk (or per layer and direction, see below).  CTA (slice, tile) owns hidden units [j0, j0+HC) for batch rows [b0, b0+BT).
// Its slice of the recurrent weights is loaded by TMA ONCE into shared memory as a K-major, 128-byte-swizzled
// UMMA operand and stays resident for all steps.  Per step and contraction:
//   warp 0, one thread: waits (acquire polls) until every slice of the tile has released the previous step,
//           then streams that step's bf16 rows from L2 - one 3-D TMA instruction for the whole K extent
//           where it fits the ring, else a ring of ~32 KB stages;
//   warp 1, converged: runs the MMA loop with warp-uniform operands; one elected lane issues the stage's
//           tcgen05.mma (M = 64/128 batch rows x N = gate columns, fp32 accumulators in TMEM) back to back;
//   warps 2-9: have every global operand of the cell update in flight before the accumulator arrives, move
//           it TMEM -> shared, then each thread updates quads of adjacent units of one row (16-byte loads
//           and stores), stores the re-quantised bf16 rows, releases the step (bar.sync + red.release.gpu)
//           and only then writes what no other CTA waits for (fp32 state, gate stash for BPTT).
// CRNN_PREC_BF16X3 ("split"): every bf16 array has a lo twin; weights are resident as [hi ; lo] rows per K
// block and each K step issues A_hi*[W_hi|W_lo] and A_lo*W_hi (see rt_mma).
// Variants the planner (crnn_api.cu) chooses between, all in these two kernels:
//   sub-chains      a tile's rows as two passes with their own step barrier, so one pass's exchange hides behind the other;
//   K split         the contraction index over 2 / 4 CTAs of a cluster, partial sums through distributed shared memory;
//   cluster exchange (xchg)  all slices of a tile in one cluster of <= 8: finished rows are written with st.async straight
//                   into every slice's operand buffer, nothing of the step hand-shake touches L2 (cx_push, rt_mma_cx);
//   one launch per layer (pchunks)  the launch covers the whole sequence and gates its chunks on flags that stream memory
//                   operations of EARLIER-enqueued work write (RecTcArgs::pchunks);
//   direct exchange (gather)  no step barrier, sentinel-filled slots checked in shared memory (opt-in, slower in situ).
#pragma once
#include "tc_common.cuh"

namespace crnn {
namespace tc {

constexpr int RT_EPI_WARPS = 8;             // two per SM sub-partition: one hides the other's latencies
constexpr int RT_EPI_THREADS = RT_EPI_WARPS * 32;
constexpr int RT_THREADS = 64 + RT_EPI_THREADS;
constexpr int RT_RING_BYTES = 64 * 1024;  // TMA ring for the streamed operand (default; RecTcArgs::ring_bytes may be larger)
constexpr int RT_MAX_STAGES = 32;
constexpr int RT_MROWS_BYTES = 128 * 128; // bytes an M=128 operand reads from a tile base (128 rows x 64 bf16)
// Ring geometry.  A tile is one K block (64 bf16 = 128 B per row) of `box_rows` rows; a STAGE is
// `chunk_kb` consecutive tiles whose TMA loads are issued back to back on ONE mbarrier (one wait /
// expect_tx / commit per chunk instead of per tile).  Tiles are packed, so an M=128 read that runs past a short
// tile sees later tiles / the slack at the ring's end; those products fall in accumulator rows nobody
// reads.  Small batch tiles get a whole phase's K extent in one or two instructions.
__host__ __device__ inline int rt_tile_bytes(int box_rows) { return ((box_rows + 7) & ~7) * 128; }
// Split operands (RecTcArgs::split): a stage holds `chunk_kb` hi tiles followed by their `chunk_kb` lo tiles.
__host__ __device__ inline int rt_chunk_kb(int box_rows, int nkb, int split = 0) {
    // ~32 KB per stage: both the producer and the MMA thread pay a few hundred cycles of barrier
    // hand-shake per stage, which is what bounds the large configurations (measured, config 4:
    // 1/2/3/4 tiles of 8 KB per stage -> 28.6 / 21.9 / 19.7 / 18.5 ms per step)
    int c = (32 * 1024) / (rt_tile_bytes(box_rows) * (1 + split));
    if (c < 1) c = 1;
    if (c > nkb) c = nkb;
    if (c > 256) c = 256;
    if (c == nkb && nkb >= 4 && (nkb & 1) == 0) c = nkb / 2;     // two instructions: MMAs start on the first half
    return c;
}
__host__ __device__ inline int rt_num_stages(int box_rows, int chunk_kb, int split = 0, int ring_bytes = RT_RING_BYTES) {
    int n = (ring_bytes - RT_MROWS_BYTES + rt_tile_bytes(box_rows)) / (rt_tile_bytes(box_rows) * chunk_kb * (1 + split));
    if (n < 1) n = 1;            // (the launcher shrinks chunk_kb until a stage fits)
    return n > RT_MAX_STAGES ? RT_MAX_STAGES : n;
}

struct RecTcArgs {
    int B, H, S;
    int t_begin, t_end;         // steps this launch covers (absolute; chunks of one sequence run as separate launches)
    int HC, BT, nslices, ntiles, npass;
    int pr;                     // rows per pass (128, or less when a tile is cut into sub-chains)
    int sub;                    // 1: every pass of a tile is a chain of its own (own step barrier word, published per pass): while
                                // one pass waits for the other slices, the CTA works on the next - the exchange latency of a
                                // pass hides behind the other passes' MMAs and cell updates (same resident weights for all)
    int GS;                     // gate-stash row width / H
    int box_rows;               // rows per A-tile TMA box
    int ckb, nst;               // ring geometry, fixed per launch: K tiles per stage, number of stages
    int use3d;                  // streamed operand maps are [kblock][row][64]: one TMA instruction per stage
    int fast_act;               // 1: MUFU.TANH activations
    int swap;                   // 1: weights are the MMA A operand (M), batch rows the N dimension
    int ksplit;                 // Q: the contraction index is split over Q CTAs of a cluster; each contracts the columns of Q*HC
                                // units over 1/Q of K and the partial sums are exchanged through distributed shared memory
    // One launch per layer and direction (pchunks > 0): the launch covers the whole sequence, cut into pchunks chunks for
    // the layer wavefront.  Before the first step of chunk k (in processing order) the kernel waits until *in_flag >= k + 1 -
    // written by a stream memory operation that follows the chunk's batch GEMM (hoisted x-projection forward, input gradients
    // of the layer above backward) on a stream whose work was enqueued BEFORE this launch -, and after the last step of a
    // chunk k every CTA adds 1 to out_flag[k] - one counter per chunk: batch tiles are independent chains and drift apart, a
    // single sum over all chunks would reach "chunk k done" while some tile is still inside it -, which the next layer's GEMM
    // stream waits for (cuStreamWaitValue32(out_flag + k, >= CTAs)).  No launch ever waits for work enqueued after it; the
    // waits are bounded and report through `err`.
    int pchunks;
    const unsigned* in_flag;
    unsigned* out_flag;
    int xchg;                   // 1: cluster exchange - the cluster holds ALL slices of a batch tile and every slice writes the bf16 rows it
                                // finishes straight into the operand buffers of all slices (distributed shared memory), see cx_push
    int cbar;                   // 1: the cluster holds ALL slices of a batch tile: steps are ordered by remote mbarrier arrives, no global counter
    int csize;                  // CTAs per cluster along the slice index (1: no cluster): the K-split group or the slices of one batch tile (cbar)
    int m64;                    // 1: tiles of <= 64 batch rows use M=64 MMAs (half the A-operand read); rows 16i..16i+15 -> TMEM lanes 32i..32i+15
    int split;                  // 1: every bf16 operand is a (hi, lo) pair (CRNN_PREC_BF16X3): hi*hi + lo*hi + hi*lo per K step
    long long hxb_lo, rsb_lo, dzb_lo;   // element offsets of the lo twins of hxb / rsb / dzb (0 when not split)
    int expt;                   // timing experiments only (CRNN_TC_EXPT; results are wrong when set): 1 no critical stores, 2 no math, 4 no deferred stores (fwd)
    // Direct exchange (gather != 0): no step barrier at all.  The bf16 rows a phase streams are copied from global memory
    // into the ring by the epilogue threads themselves, and every 16-bit half of the copied words is checked against a
    // sentinel (0xFFFF, a NaN pattern no conversion produces) the slots were filled with before: a word is either still
    // the sentinel or final, so data needs no flag, no fence and no poll of a counter - see rt_validate.
    int gather;
    int g_rp_log2;              // log2 of the tile's rows padded to a power of two (>= 32)
    int g_ckb_log2;             // log2(ckb): K blocks per ring stage
    int fill_t0, fill_t1;       // steps whose output slots this launch fills with the sentinel for the NEXT launch of the chain
    unsigned* err;              // set to 1 when a wait ran into its bound (host-mapped; the host turns it into CRNN_ERR_CUDA)
    int acq_poll;               // 1: the step barrier is polled with acquire loads (no fence after the poll loop)
    int defer;                  // 1: stores only this CTA reads later are issued after the step barrier's release
    int rotate;                 // 1: slices of a tile walk the K blocks from different starting stages, so they do not all
                                // ask L2 for the same rows at the same moment
    int ring_bytes;             // size of the TMA ring (>= RT_RING_BYTES, multiple of 1024)
    int zs_floats;              // shared staging for one accumulator pass: rows x (columns + 1)
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
    long long* dbg;             // optional (CRNN_TC_DEBUG): four CTAs record clock64 stamps [cta][phase 20..27][16] in shared memory and dump them at exit
    unsigned* barrier;          // one zeroed counter per batch tile: rows of different tiles never interact
};

__device__ __forceinline__ float tanh_mufu(float x) {
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// `fast` (block-uniform): one MUFU.TANH per activation (|err| ~ 5e-4, well under bf16's 4e-3
// operand rounding) instead of the accurate expf/tanhf sequences.
// !FAST (the 1e-4 parity bar of CRNN_PREC_BF16X3): ex2.approx + rcp.approx forms, absolute error < 1e-6 -
// a tenth of the split-operand contraction error - at a fifth of the instructions of expf / tanhf, which
// sit on the serial chain of every step.  Saturate correctly: e^x -> inf gives 1/inf = 0.
__device__ __forceinline__ float sigmoid_ex2(float x) { return __fdividef(1.0f, 1.0f + __expf(-x)); }
__device__ __forceinline__ float tanh_ex2(float x) { return 1.0f - __fdividef(2.0f, 1.0f + __expf(2.0f * x)); }
#define fast_tanh(x) (FAST ? tanh_mufu(x) : tanh_ex2(x))
#define fast_sigmoid(x) (FAST ? fmaf(0.5f, tanh_mufu(0.5f * (x)), 0.5f) : sigmoid_ex2(x))

__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, %0;" ::"n"(RT_EPI_THREADS) : "memory"); }

// ---- cluster exchange (RecTcArgs::xchg) ------------------------------------------------------------------------------
// When all slices of a batch tile sit in ONE cluster (<= 8 CTAs: small hidden sizes, the reference's default 128 among
// them), the rows a slice finishes do not go through L2 at all: the epilogue threads write them, already in the K-major
// 128-byte-swizzled operand layout, into the operand buffer of EVERY slice of the cluster with st.async, whose bytes
// complete the transaction count of the buffer's mbarrier in the receiving CTA, on which that CTA's MMA warp sleeps.
// No store to global memory, no MEMBAR.GPU, no flag in L2, no poll, no TMA load - not even a fence or an arrive of the
// writer - are left on the chain.  Buffers are double-buffered by step (or by phase for the GRU), which the data dependencies make safe
// without an "empty" hand-shake: a slice can only produce rows for use k+1 of a buffer after every peer's MMAs of use k
// have completed (it needed their results).  Measured (tools/ubench_dsmem.cu, B200): an all-to-all of 8 x 2 KB inside a
// cluster takes 1.4 k cycles against 3.0-3.8 k for store + release + acquire poll + bulk load through L2; at 16 x 4 KB
// (config 3's geometry) the cluster network (~16 B/cycle/SM in) loses to L2: 4.3 k against 3.5 k.
constexpr int RT_CX_MAX = 8;
__device__ __forceinline__ void st_dsmem_v2(uint32_t cluster_addr, uint32_t x, uint32_t y) {
    asm volatile("st.shared::cluster.v2.b32 [%0], {%1, %2};" ::"r"(cluster_addr), "r"(x), "r"(y) : "memory");
}
// 8 bytes into a peer's shared memory; their arrival is counted (complete_tx) on the peer's mbarrier `cluster_bar`
__device__ __forceinline__ void st_async_v2(uint32_t cluster_addr, uint32_t x, uint32_t y, uint32_t cluster_bar) {
    asm volatile("st.async.shared::cluster.mbarrier::complete_tx::bytes.v2.b32 [%0], {%1, %2}, [%3];"
                 ::"r"(cluster_addr), "r"(x), "r"(y), "r"(cluster_bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_cnt(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
// Four values v[] = K indices kcol .. kcol+3 of row r of the operand buffer at byte offset buf_ofs of every peer's ring
// (pbase[p] = the ring's address in CTA p's window, pbar[p] = the buffer's barrier there).  Buffer layout: nkb tiles of
// [box rows][128 B] (K block kb = kcol / 64, 16-byte chunk c of a row stored at chunk c ^ (r % 8)), the lo tiles lo_ofs bytes
// after the hi tiles (0: not split).  st.async: the bytes count themselves on the receiver's barrier when they land, so the
// writer needs no fence and no arrive, and the receiver only states how many bytes a complete buffer is (rt_mma_cx).
__device__ __forceinline__ void cx_push(const uint32_t (&pbase)[RT_CX_MAX], const uint32_t (&pbar)[RT_CX_MAX], int csize, int bidx,
                                        uint32_t buf_ofs, uint32_t lo_ofs, int tile, int r, int kcol, const float* v) {
    const uint32_t byte = (uint32_t)(kcol & 63) * 2u;
    const uint32_t off = buf_ofs + (uint32_t)(kcol >> 6) * (uint32_t)tile + (uint32_t)r * 128u + ((((byte >> 4) ^ (uint32_t)r) & 7u) << 4) + (byte & 15u);
    const __nv_bfloat162 h0 = __floats2bfloat162_rn(v[0], v[1]), h1 = __floats2bfloat162_rn(v[2], v[3]);
    const float2 f0 = __bfloat1622float2(h0), f1 = __bfloat1622float2(h1);
    const __nv_bfloat162 l0 = __floats2bfloat162_rn(v[0] - f0.x, v[1] - f0.y), l1 = __floats2bfloat162_rn(v[2] - f1.x, v[3] - f1.y);
#pragma unroll
    for (int p = 0; p < RT_CX_MAX; ++p)
        if (p < csize) {
            const uint32_t bar = pbar[p] + 8u * (uint32_t)bidx;
            st_async_v2(pbase[p] + off, *reinterpret_cast<const uint32_t*>(&h0), *reinterpret_cast<const uint32_t*>(&h1), bar);
            if (lo_ofs) st_async_v2(pbase[p] + off + lo_ofs, *reinterpret_cast<const uint32_t*>(&l0), *reinterpret_cast<const uint32_t*>(&l1), bar);
        }
}

__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&v);
}

#define RT_STAMP(slot)                                                                     \
    do {                                                                                   \
        if (dbgs && dbg_phase >= 20 && dbg_phase < 28) dbgs[(dbg_phase - 20) * 24 + (slot)] = clock64(); \
    } while (0)


// ---- direct exchange of the streamed rows --------------------------------------------------------------------------
// Producers store final bf16 words straight to L2 (.cg), consumers read them from L2 (.cg: never from this SM's L1, which
// could hold a stale copy of the line) until no 16-bit half is the sentinel.  Each 32-bit word is written exactly once
// between two fills.  (Measured: the same loads as ld.relaxed.gpu are issued one at a time - sixteen of them took 7 k
// cycles, one L2 round trip each - while .cg loads pipeline.)
constexpr int RT_GATHER_SLOTS = 16;            // 16-byte chunks one thread keeps in flight
constexpr int RT_GATHER_SPINS = 1 << 22;       // re-reads of one chunk before the wait is declared dead (seconds)
__device__ __forceinline__ uint4 ld_relaxed_gpu_v4(const void* p) {     // coherent at gpu scope (LDG.STRONG.GPU): always from L2
    uint4 v;
    asm volatile("ld.relaxed.gpu.global.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
// First attempt: a weak load that does not allocate in L1.  Strong loads of one thread complete one at a time (measured:
// sixteen took 7 k cycles, one L2 round trip each); these pipeline.  Should L1 ever serve one of them from a stale line,
// the stale value can only be the sentinel (every word is written once per fill), and the retry below uses strong loads.
__device__ __forceinline__ uint4 ld_weak_nol1_v4(const void* p) {
    uint4 v;
    asm volatile("ld.global.L1::no_allocate.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ void st_relaxed_gpu_v2(void* p, uint32_t a, uint32_t b) {
    asm volatile("st.global.cg.v2.b32 [%0], {%1, %2};" ::"l"(p), "r"(a), "r"(b) : "memory");
}
__device__ __forceinline__ void st_relaxed_gpu_b16(void* p, uint16_t a) {
    asm volatile("st.global.cg.b16 [%0], %1;" ::"l"(p), "h"(a) : "memory");
}
__device__ __forceinline__ bool rt_has_sentinel(const uint4& v) {
    return (__vcmpeq2(v.x, 0xFFFFFFFFu) | __vcmpeq2(v.y, 0xFFFFFFFFu) | __vcmpeq2(v.z, 0xFFFFFFFFu) | __vcmpeq2(v.w, 0xFFFFFFFFu)) != 0u;
}
__device__ __forceinline__ void st_shared_v4(uint32_t saddr, const uint4& v) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(saddr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

struct RtSmem {
    uint8_t* wsa; uint8_t* wsb; uint8_t* ring; float* zs;
    uint64_t *a_full, *a_empty, *a_valid, *z_full, *z_empty, *w_bar, *step_bar, *zs_ready, *zs_free;
    uint32_t* tmem_slot;
    long long* dbg;          // 8 phases x 24 stamps
};

__device__ __forceinline__ RtSmem rt_carve(uint8_t* smem_raw, size_t wsa_bytes, size_t wsb_bytes, int zs_floats, int ring_bytes) {
    RtSmem s;
    uint8_t* p = reinterpret_cast<uint8_t*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    s.wsa = p; p += wsa_bytes;
    s.wsb = p; p += wsb_bytes;
    s.ring = p; p += ring_bytes;
    s.zs = reinterpret_cast<float*>(p); p += (size_t)((zs_floats + 3) & ~3) * 4;
    s.a_full = reinterpret_cast<uint64_t*>(p);
    s.a_empty = s.a_full + RT_MAX_STAGES;
    s.a_valid = s.a_empty + RT_MAX_STAGES;
    s.z_full = s.a_valid + RT_MAX_STAGES;
    s.z_empty = s.z_full + 1;
    s.w_bar = s.z_empty + 1;
    s.step_bar = s.w_bar + 1;
    s.zs_ready = s.step_bar + 1;
    s.zs_free = s.zs_ready + 1;
    s.tmem_slot = reinterpret_cast<uint32_t*>(s.zs_free + 1);
    s.dbg = reinterpret_cast<long long*>(s.tmem_slot + 2);
    return s;
}

// (alignment slack + barriers and TMEM slot + the debug stamps, which sit last and exist only in debug runs)
inline size_t rt_smem_bytes(size_t wsa_bytes, size_t wsb_bytes, int zs_floats, int ring_bytes = RT_RING_BYTES, int dbg = 1) {
    return wsa_bytes + wsb_bytes + ring_bytes + (size_t)((zs_floats + 3) & ~3) * 4 + 1024 + 1024 + (dbg ? 2048 : 0);
}

// Role-private pipeline state.
struct RtPipe {
    uint32_t ring_i;      // ring stages issued / consumed so far
    uint32_t z_i;         // accumulator hand-offs so far
    uint32_t st;          // current ring stage  (= ring_i % nst, kept incrementally)
    uint32_t ph;          // current ring phase bit (= (ring_i / nst) & 1)
    __device__ __forceinline__ void advance(int nst) {
        ++ring_i;
        if (++st == (uint32_t)nst) { st = 0; ph ^= 1u; }
    }
};

// The streamed operand of one contraction, copied by the RT_EPI_THREADS epilogue threads (RecTcArgs::gather): rows
// [row0, row0 + nrows) x columns [k0, k0 + nkb*64) of `src` (row pitch `ld` elements, lo twin `lo_ofs` elements
// further when nsp == 2) go into the ring stage by stage, each K block as a K-major tile of 128-byte rows whose
// 16-byte chunks are XOR-swizzled by (row & 7) - what a SWIZZLE_128B TMA load would have produced.  Chunk index
// inside a stage: [half][row (padded to 2^rp_log2)][K block][chunk of 8 elements]; thread i owns chunks i, i + 256, ...
// so a warp reads whole contiguous row segments and its shared-memory stores cover all banks.  The low 8 bits of a
// chunk index are the thread's (column, K block, low row bits - fixed for the whole kernel), the high bits the slot's
// (block-uniform), so per chunk only a handful of instructions remain.  Up to RT_GATHER_SLOTS chunks per thread are in
// flight at once; a chunk is re-read until it holds no sentinel half.  Columns >= kvalid are written as zeros; rows
// >= nrows are never written (the ring starts zeroed).
__device__ __forceinline__ bool rt_word_has_sentinel(uint32_t w) {      // a 16-bit half of w is 0xFFFF
    const uint32_t x = ~w;
    return ((x - 0x00010001u) & w & 0x80008000u) != 0u;               // "has a zero halfword" on ~w
}
__device__ __forceinline__ bool rt_chunk_has_sentinel(const uint4& v) {
    return rt_word_has_sentinel(v.x) | rt_word_has_sentinel(v.y) | rt_word_has_sentinel(v.z) | rt_word_has_sentinel(v.w);
}
// Everything about one phase's check that does not change from call to call inside a launch.
struct RtGatherGeom {
    uint32_t ring0, a_full0, a_valid0;   // shared-memory addresses: ring, first full / valid barrier
    int nsp, tile, ckb_log2, rp_log2, nst, expt;
    unsigned* err;
};
__device__ __forceinline__ uint4 ld_shared_v4(uint32_t saddr) {
    uint4 v;
    asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(saddr) : "memory");
    return v;
}
// Direct exchange, consumer side.  The rows a phase streams are still moved by TMA (one instruction per ring stage; the
// TMA unit reads L2 on its own path, so the copy does not queue behind the epilogue's loads and stores - a version that
// copied with cp.async from the epilogue threads ran 2x slower next to the other layer's kernel for that reason), but
// nothing orders those loads after the other slices' stores: no release, no flag, no fence.  Instead every slot was
// filled with a sentinel (0xFFFF in each 16-bit half: a NaN pattern no float->bf16 conversion produces) before, every
// word is written exactly once, and the RT_EPI_THREADS epilogue threads read the landed stage back from shared memory:
// a 16-byte chunk with a sentinel half was loaded too early and is re-read from L2 with coherent loads until complete,
// then patched in place.  Chunk index inside a stage: [half][row (padded to 2^rp_log2)][K block][chunk of 8 elements];
// thread i owns chunks i, i + 256, ...  Rows >= nrows belong to other tiles and are not looked at (their products land
// in accumulator rows nobody reads); columns >= kvalid were zero-filled by TMA.
// `stph`: the caller's ring position (stage | phase bit << 16), returned advanced.
__device__ __forceinline__ uint32_t rt_validate(const RtGatherGeom& G, uint32_t stph, const __nv_bfloat16* __restrict__ src, int64_t lo_ofs,
                                                int64_t row0, int nrows, int ld, int k0, int kvalid, int nkb, int epi_tid, int lane,
                                                long long* stamps) {
    const int nsp = G.nsp, tile = G.tile, ckb_log2 = G.ckb_log2, rp_log2 = G.rp_log2, nst = G.nst;
    uint32_t st = stph & 0xFFFFu, ph = stph >> 16;
    const int hbit = nsp - 1;
    const int L = hbit + rp_log2 + ckb_log2 + 3;             // log2(chunks per stage), L >= 8 (host)
    const int ckb = 1 << ckb_log2;
    const int qs = 1 << (L - 8);                             // chunks per stage and thread
    const int nstages = (nkb + ckb - 1) >> ckb_log2;
    // the thread's part of every chunk it owns: low 8 bits of the chunk index
    const int c = epi_tid & 7, kbl = (epi_tid >> 3) & (ckb - 1), r_low = epi_tid >> (3 + ckb_log2);
    const int kcol_t = kbl * 64 + c * 8;
    const __nv_bfloat16* gsrc = src + (row0 + r_low) * (int64_t)ld + k0 + kcol_t;
    const uint32_t s_thr = G.ring0 + (uint32_t)(kbl * tile + r_low * 128);
    const uint32_t stage_bytes = (uint32_t)(ckb * tile * nsp), half_bytes = (uint32_t)(ckb * tile);
    if (stamps && epi_tid == 0) stamps[19] = clock64();
    long long nretry = 0;
#pragma unroll 1
    for (int sv = 0; sv < nstages; ++sv) {
        {   // the stage's TMA loads have landed
            const uint32_t bar = G.a_full0 + st * 8u;
            uint32_t ok;
            do {
                asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                             : "=r"(ok) : "r"(bar), "r"(ph) : "memory");
            } while (!ok);
        }
        if (stamps && epi_tid == 0 && sv == 0) stamps[20] = clock64();
        const int kb0 = sv << ckb_log2;
        const uint32_t sbase = s_thr + st * stage_bytes;
        if ((kb0 + kbl) < nkb && (kb0 * 64 + kcol_t) < kvalid && !(G.expt & 8)) {
#pragma unroll 2
            for (int i = 0; i < qs; ++i) {
                const int hi_bits = i << 8;                           // the slot's part of the chunk index
                const int half = hbit ? (hi_bits >> (L - 1)) : 0;
                const int rq = (hi_bits >> (ckb_log2 + 3)) & ((1 << rp_log2) - 1);
                const int r = r_low + rq;
                if (r >= nrows) continue;
                const uint32_t sa = sbase + (half ? half_bytes : 0u) + (uint32_t)(rq * 128 + ((c ^ (r & 7)) << 4));
                uint4 v = ld_shared_v4(sa);
                if (!rt_chunk_has_sentinel(v)) continue;
                const __nv_bfloat16* g = gsrc + (int64_t)rq * ld + (half ? lo_ofs : 0) + kb0 * 64;
                int spins = 0;
                do {
                    v = ld_relaxed_gpu_v4(g);
                    // bounded: a dead wait raises the host-visible error word, and every other wait of the launch then gives up too
                    if ((++spins & 1023) == 0 && G.err && (spins > RT_GATHER_SPINS || *reinterpret_cast<volatile unsigned*>(G.err))) {
                        *reinterpret_cast<volatile unsigned*>(G.err) = 1u;
                        break;
                    }
                } while (rt_chunk_has_sentinel(v));
                st_shared_v4(sa, v);
                nretry += spins;
            }
        }
        fence_proxy_async();                                  // patches -> visible to the tensor core's reads
        __syncwarp();
        if (lane == 0)                                        // one arrival per epilogue warp
            asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(G.a_valid0 + st * 8u) : "memory");
        if (++st == (uint32_t)nst) { st = 0; ph ^= 1u; }
        if (stamps && epi_tid == 0 && sv == 0) stamps[21] = clock64();
    }
    if (stamps && epi_tid == 0) { stamps[22] = clock64(); stamps[23] = nretry; }
    return st | (ph << 16);
}

// Direct exchange, trigger of the TMA loads (warp 0, all lanes): wait until a word from the END of every slice's share of
// the rows (last row, last two units; the lo twin when there is one - it is stored after the hi word) is no longer the
// sentinel.  Only a hint that the rows are about to be complete - it keeps the number of chunks rt_validate has to patch
// small; nothing depends on it for correctness.
__device__ __forceinline__ void rt_probe(const __nv_bfloat16* __restrict__ src, int64_t lo_ofs, int64_t row_last, int ld, int col0, int HC,
                                         int ncols, int nslices, int lane, unsigned* err) {
    for (int sl = lane; sl < nslices; sl += 32) {
        const int col = min(col0 + sl * HC + HC, ncols) - 2;
        const uint32_t* w = reinterpret_cast<const uint32_t*>(src + lo_ofs + row_last * (int64_t)ld + col);
        int spins = 0;
        uint32_t v;
        do {
            asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(w) : "memory");
            if ((++spins & 1023) == 0 && err && (spins > RT_GATHER_SPINS || *reinterpret_cast<volatile unsigned*>(err))) {
                *reinterpret_cast<volatile unsigned*>(err) = 1u;
                break;
            }
        } while (rt_word_has_sentinel(v));
    }
    __syncwarp();
}

// Sentinel over the bf16 slots rows [row0, row0 + nrows) x columns [c0, c0 + ncols) of `dst` (ncols % 8 == 0, 16-byte aligned).
__device__ __forceinline__ void rt_fill_sentinel(__nv_bfloat16* __restrict__ dst, int64_t lo_ofs, int64_t row0, int nrows, int ld,
                                                 int c0, int ncols, int epi_tid) {
    const int per_row = ncols >> 3;
    const uint4 ff = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu);
    for (int i = epi_tid; i < nrows * per_row; i += RT_EPI_THREADS) {
        const int r = i / per_row, c = i - r * per_row;
        __nv_bfloat16* q = dst + (row0 + r) * (int64_t)ld + c0 + c * 8;
        *reinterpret_cast<uint4*>(q) = ff;
        if (lo_ofs) *reinterpret_cast<uint4*>(q + lo_ofs) = ff;
    }
}

// One contraction Z[128 x NC] = A_tile[rows x K] * W_region^T, split across the three roles.
// `arow`: first row of the A tile in the 2-D tensor view; `wbase`: resident weight region.
__device__ __forceinline__ void rt_produce(const RtSmem& s, RtPipe& p, const CUtensorMap* tmA, int arow, int nkb, int box_rows,
                                           int ckb, int nst, int use3d, int kb_ofs = 0,
                                           const CUtensorMap* tmAlo = nullptr, int rot = 0, long long* stamps = nullptr,
                                           const void* lin = nullptr, const void* lin_lo = nullptr) {
    const int tile = rt_tile_bytes(box_rows);
    // lin != nullptr (CRNN_TC_EXPT=64, timing only - the rows land unswizzled, results are wrong): the same number of bytes
    // as ONE linear bulk copy per operand half instead of a tensor-map load
    auto bulk1d = [&](void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
    };
    if (tmAlo) {                                       // split operands: hi tiles then lo tiles, one barrier per stage
        const int stage = 2 * ckb * tile;
        for (int kb = 0; kb < nkb; kb += ckb, p.advance(nst)) {
            const int st = (int)p.st;
            const int n = min(ckb, nkb - kb);
            mbar_wait(&s.a_empty[st], p.ph ^ 1u);
            mbar_expect_tx(&s.a_full[st], 2u * (uint32_t)box_rows * 128u * (uint32_t)(use3d ? ckb : n));
            uint8_t* hi = s.ring + st * stage;
            uint8_t* lo = hi + ckb * tile;
            const int kbr = kb + rot >= nkb ? kb + rot - nkb : kb + rot;      // rotated start (rot is a multiple of ckb)
            if (lin && use3d) {
                bulk1d(hi, lin, (uint32_t)box_rows * 128u * (uint32_t)ckb, &s.a_full[st]);
                bulk1d(lo, lin_lo, (uint32_t)box_rows * 128u * (uint32_t)ckb, &s.a_full[st]);
            } else if (use3d) {
                tma_load_3d(hi, tmA, &s.a_full[st], 0, arow, kb_ofs + kbr);
                tma_load_3d(lo, tmAlo, &s.a_full[st], 0, arow, kb_ofs + kbr);
            } else {
                for (int j = 0; j < n; ++j) {
                    tma_load_2d(hi + j * tile, tmA, &s.a_full[st], (kb_ofs + kbr + j) * 64, arow);
                    tma_load_2d(lo + j * tile, tmAlo, &s.a_full[st], (kb_ofs + kbr + j) * 64, arow);
                }
            }
        }
        return;
    }
    for (int kb = 0; kb < nkb; kb += ckb, p.advance(nst)) {
        const int st = (int)p.st;
        const int n = min(ckb, nkb - kb);
        mbar_wait(&s.a_empty[st], p.ph ^ 1u);
        if (stamps && kb <= ckb) stamps[kb == 0 ? 19 : 21] = clock64();
        const uint32_t bytes = (uint32_t)box_rows * 128u * (uint32_t)(use3d ? ckb : n);
        mbar_expect_tx(&s.a_full[st], bytes);
        const int kbr = kb + rot >= nkb ? kb + rot - nkb : kb + rot;
        if (lin && use3d) {
            bulk1d(s.ring + st * ckb * tile, lin, bytes, &s.a_full[st]);
        } else if (use3d) {                            // whole stage in one instruction (out-of-range k-blocks read zero)
            tma_load_3d(s.ring + st * ckb * tile, tmA, &s.a_full[st], 0, arow, kb_ofs + kbr);
            if (stamps && kb <= ckb) stamps[kb == 0 ? 20 : 22] = clock64();
        } else {
            for (int j = 0; j < n; ++j)
                tma_load_2d(s.ring + (st * ckb + j) * tile, tmA, &s.a_full[st], (kb_ofs + kbr + j) * 64, arow);
        }
    }
}
// swap == 0: D[128 batch rows x nc gate columns]   (A = streamed rows, B = resident weights)
// swap == 1: D[128 gate columns x nrows16 batch rows] (A = resident weights, B = streamed rows):
//            the MMA cost is max(M,128)*N/256 cycles per K=16, so the smaller of (nc, rows) goes to N.
//            Weight regions hold nc < 128 rows; the rows the M=128 operand reads past them are other
//            resident data whose products land in accumulator lanes nobody reads.
// Executed by ALL lanes of the (converged) MMA warp with warp-uniform arguments; one elected lane issues.
__device__ __forceinline__ void rt_mma(const RtSmem& s, RtPipe& p, uint32_t tmem, const uint8_t* wbase, int nc, int nkb,
                                       int swap, int nrows16, int box_rows, int ckb, int nst, int m64,
                                       long long* stamps = nullptr, int split = 0, int rot = 0, int checked = 0, int lite = 0) {
    const int tile = rt_tile_bytes(box_rows);
    const bool lane0 = (threadIdx.x & 31) == 0;
    // split: lo tiles sit ckb tiles after their hi tiles in the stage, the lo weight region right after the hi one
    const uint64_t a_lo = (uint64_t)((ckb * tile) >> 4), w_lo = (uint64_t)(((uint32_t)nkb * (uint32_t)nc * 128u) >> 4);
    mbar_wait(s.z_empty, (p.z_i & 1) ^ 1);          // epilogue has drained the previous accumulator
    tc_fence_after();
    if (stamps && lane0) stamps[10] = clock64();
    const uint32_t idesc = swap ? idesc_bf16(128, nrows16, 0, 0) : idesc_bf16(m64 ? 64 : 128, nc, 0, 0);
    const uint32_t w0 = smem_u32(wbase);
    for (int kb0 = 0; kb0 < nkb; kb0 += ckb, p.advance(nst)) {
        const int st = (int)p.st;
        mbar_wait(checked ? &s.a_valid[st] : &s.a_full[st], p.ph);     // direct exchange: the epilogue threads have checked the stage
        tc_fence_after();
        if (stamps && lane0) stamps[kb0 == 0 ? 11 : 12] = clock64();
        const int n = min(ckb, nkb - kb0);
        // Descriptors are built once per stage and advanced by plain adds (K step of 16 bf16 = 32 B = +2 in
        // the descriptor's address field): every instruction between two tcgen05.mma costs tensor-pipe idle time.
        uint64_t da = desc_kmajor(smem_u32(s.ring + st * ckb * tile * (1 + split)));
        const int kbr = kb0 + rot >= nkb ? kb0 + rot - nkb : kb0 + rot;
        uint64_t db = desc_kmajor(w0 + (uint32_t)kbr * (uint32_t)nc * 128u);
        const uint64_t da_step = (uint64_t)(tile >> 4), db_step = (uint64_t)((nc * 128) >> 4);
        // One divergent region per stage: the elected lane issues the stage's MMAs back to back and commits.
        if (elect_one()) {
            if (split && !swap) {
                // Weights of a K block are resident as [hi rows ; lo rows] (2*nc rows): one MMA of N = 2*nc gives
                // A_hi*W_hi in accumulator columns [0, nc) and A_hi*W_lo in [nc, 2nc), a second of N = nc adds A_lo*W_hi
                // to [0, nc); the epilogue sums the two column halves.  Two instructions and one read of the A_hi tile
                // less per K step than three separate products (the MMAs are paced by their shared-memory reads).
                const uint32_t idesc2 = idesc_bf16(m64 ? 64 : 128, 2 * nc, 0, 0);
                db = desc_kmajor(w0 + (uint32_t)kbr * (uint32_t)(2 * nc) * 128u);
                const uint64_t db2_step = (uint64_t)((2 * nc * 128) >> 4);
                for (int j = 0; j < n; ++j, da += da_step, db += db2_step) {
                    if (lite) { mma_bf16(tmem, da, db, idesc2, (uint32_t)((kb0 + j) != 0)); continue; }   // timing experiment: an eighth of the MMAs
#pragma unroll
                    for (int k = 0; k < 8; k += 2) mma_bf16(tmem, da + k, db + k, idesc2, (uint32_t)((kb0 + j + k) != 0));
#pragma unroll
                    for (int k = 0; k < 8; k += 2) mma_bf16(tmem, da + a_lo + k, db + k, idesc, 1u);
                }
            } else if (split) {
                for (int j = 0; j < n; ++j, da += da_step, db += db_step) {
#pragma unroll
                    for (int k = 0; k < 8; k += 2) mma_bf16(tmem, db + k, da + k, idesc, (uint32_t)((kb0 + j + k) != 0));
#pragma unroll
                    for (int k = 0; k < 8; k += 2) mma_bf16(tmem, db + k, da + a_lo + k, idesc, 1u);
#pragma unroll
                    for (int k = 0; k < 8; k += 2) mma_bf16(tmem, db + w_lo + k, da + k, idesc, 1u);
                }
            } else if (!swap) {
                for (int j = 0; j < n; ++j, da += da_step, db += db_step) {
                    mma_bf16(tmem, da, db, idesc, (uint32_t)((kb0 + j) != 0));
                    if (lite) continue;
                    mma_bf16(tmem, da + 2, db + 2, idesc, 1u);
                    mma_bf16(tmem, da + 4, db + 4, idesc, 1u);
                    mma_bf16(tmem, da + 6, db + 6, idesc, 1u);
                }
            } else {
                for (int j = 0; j < n; ++j, da += da_step, db += db_step) {
                    mma_bf16(tmem, db, da, idesc, (uint32_t)((kb0 + j) != 0));
                    mma_bf16(tmem, db + 2, da + 2, idesc, 1u);
                    mma_bf16(tmem, db + 4, da + 4, idesc, 1u);
                    mma_bf16(tmem, db + 6, da + 6, idesc, 1u);
                }
            }
            mma_commit(&s.a_empty[st]);
            if (kb0 + ckb >= nkb) mma_commit(s.z_full);      // last stage: the accumulator is complete
        }
        __syncwarp();
    }
    ++p.z_i;
}

// Cluster exchange: the phase's whole streamed operand sits in one buffer `abuf` ([hi: nkb tiles][lo: nkb tiles], written by
// the slices of the cluster, see cx_push) whose barrier `full` completes when every slice has delivered its rows.
// Batch rows on M only.  Executed by all lanes of the MMA warp like rt_mma.
__device__ __forceinline__ void rt_mma_cx(const RtSmem& s, RtPipe& p, uint32_t tmem, const uint8_t* wbase, int nc, int nkb,
                                          const uint8_t* abuf, int tile, uint64_t* full, uint32_t parity, uint32_t expect_bytes,
                                          int m64, int split, long long* stamps = nullptr) {
    const bool lane0 = (threadIdx.x & 31) == 0;
    if (lane0) mbar_expect_tx(full, expect_bytes);  // the buffer's one arrival; the bytes come from the slices (or the launch's first TMA load)
    mbar_wait(s.z_empty, (p.z_i & 1) ^ 1);          // epilogue has drained the previous accumulator
    if (stamps && lane0) stamps[10] = clock64();
    mbar_wait_cluster(full, parity);                // acquire.cluster: the peers' rows have landed
    if (stamps && lane0) stamps[11] = clock64();
    fence_proxy_async();                            // (rows written by st.async -> read by the tensor core)
    tc_fence_after();
    if (stamps && lane0) stamps[12] = clock64();
    const uint32_t idesc = idesc_bf16(m64 ? 64 : 128, nc, 0, 0);
    const uint64_t a_lo = (uint64_t)(((uint32_t)nkb * (uint32_t)tile) >> 4);
    uint64_t da = desc_kmajor(smem_u32(abuf));
    const uint64_t da_step = (uint64_t)(tile >> 4);
    if (elect_one()) {
        if (split) {                                 // see rt_mma: [hi rows ; lo rows] per K block of the weights
            const uint32_t idesc2 = idesc_bf16(m64 ? 64 : 128, 2 * nc, 0, 0);
            uint64_t db = desc_kmajor(smem_u32(wbase));
            const uint64_t db_step = (uint64_t)((2 * nc * 128) >> 4);
            for (int j = 0; j < nkb; ++j, da += da_step, db += db_step) {
#pragma unroll
                for (int k = 0; k < 8; k += 2) mma_bf16(tmem, da + k, db + k, idesc2, (uint32_t)((j + k) != 0));
#pragma unroll
                for (int k = 0; k < 8; k += 2) mma_bf16(tmem, da + a_lo + k, db + k, idesc, 1u);
            }
        } else {
            uint64_t db = desc_kmajor(smem_u32(wbase));
            const uint64_t db_step = (uint64_t)((nc * 128) >> 4);
            for (int j = 0; j < nkb; ++j, da += da_step, db += db_step) {
                mma_bf16(tmem, da, db, idesc, (uint32_t)(j != 0));
                mma_bf16(tmem, da + 2, db + 2, idesc, 1u);
                mma_bf16(tmem, da + 4, db + 4, idesc, 1u);
                mma_bf16(tmem, da + 6, db + 6, idesc, 1u);
            }
        }
        mma_commit(s.z_full);
    }
    __syncwarp();
    if (stamps && lane0) stamps[3] = clock64();
    ++p.z_i;
}

// Grid barrier, split: epilogue threads publish, the TMA producer waits before touching
// rows written by other CTAs (generic-proxy writes -> async-proxy reads need the proxy fence).
__device__ __forceinline__ void rt_publish(unsigned* counter, int epi_tid, int synced = 0) {
    if (!synced) epi_bar_sync();                     // every epilogue thread's stores happen-before ...
    if (epi_tid == 0)                                // ... this release, which is cumulative over them
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(counter) : "memory");
}
// (Measured alternative: every epilogue warp releasing its own rows - no CTA barrier before the release - is slower,
//  config 3: 3.11 -> 3.26 ms per step: eight times as many atomics on the tile's counter serialise in L2.)
// Cluster form of the same hand-shake (the cluster = all slices of this batch tile): every slice
// arrives on every slice's step barrier; the waiting producer sleeps on its own shared-memory barrier
// instead of polling L2.  The data itself still travels through L2 (generic stores -> TMA loads), so
// the arriving threads fence at gpu scope first.
__device__ __forceinline__ void rt_publish_cluster(uint64_t* step_bar, int epi_tid, int csize) {
    epi_bar_sync();
    if (epi_tid < csize) {
        asm volatile("fence.acq_rel.gpu;" ::: "memory");
        mbar_arrive_remote(step_bar, (uint32_t)epi_tid);
    }
}
__device__ __forceinline__ void rt_await_cluster(uint64_t* step_bar, unsigned phases_done) {
    mbar_wait_cluster(step_bar, (phases_done - 1u) & 1u);
    asm volatile("fence.proxy.async;" ::: "memory");
}
__device__ __forceinline__ void rt_await(const unsigned* counter, unsigned target, int acq_poll) {
    unsigned v;
    if (acq_poll) {
        // acquire polls: the load that sees the last slice's release synchronises with it (the reds form its release
        // sequence); no separate fence afterwards, only the hand-over to the async proxy (TMA)
        do {
            asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(counter) : "memory");
        } while (v < target);
    } else {
        do {   // relaxed polls (one L2 round trip each) ...
            asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(counter) : "memory");
        } while (v < target);
        asm volatile("fence.acq_rel.gpu;" ::: "memory");  // ... then one acquire
    }
    asm volatile("fence.proxy.async;" ::: "memory");
}

// Persistent form: wait for the chunk flag (see RecTcArgs::pchunks).  One thread per CTA; bounded (~4 s), false on expiry.
__device__ __forceinline__ bool rt_wait_chunk(const unsigned* flag, unsigned target, unsigned* err) {
    unsigned v, spins = 0;
    unsigned long long t0 = 0;
    for (;;) {
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
        if (v >= target) return true;
        __nanosleep(128);
        if ((++spins & 255u) == 0u) {
            unsigned long long now;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
            if (t0 == 0) t0 = now;
            else if (now - t0 > 4000000000ull) { if (err) *err = 1u; return false; }
        }
    }
}
__device__ __forceinline__ int rt_chunk_t0(int S, int c, int C) { return (int)((long long)S * c / C); }

// Accumulator -> shared Zs[r*ldz + c] for batch rows r < nr and gate columns c < nc (ldz odd).
// A warp can only read its own TMEM lane quarter; quarters holding nothing useful skip the load.
// pair != 0: the accumulator holds two column halves [0, nc) and [nc, 2nc) whose sum is the result (split operands).
__device__ __forceinline__ void zs_dump(uint32_t tmem, int q, int half, int lane, int nr, int nc, float* __restrict__ Zs,
                                        int ldz, int swap, int m64, int pair = 0) {
    float v[32];
    float w[32];
    constexpr int NH = RT_EPI_WARPS / 4;             // warps per lane quarter
    if (m64 && !swap) {
        // M=64 accumulator: batch row 16*q + lane lives in lane (32*q + lane), lane < 16
        if (16 * q < nr) {
            const int r = 16 * q + lane;
            for (int c0 = 32 * half; c0 < nc; c0 += 32 * NH) {
                tmem_ld32(tmem + ((uint32_t)(32 * q) << 16) + (uint32_t)c0, v);
                if (pair) {
                    tmem_ld32(tmem + ((uint32_t)(32 * q) << 16) + (uint32_t)(nc + c0), w);
#pragma unroll
                    for (int i = 0; i < 32; ++i) v[i] += w[i];
                }
                if (lane < 16 && r < nr) {
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        if (c0 + i < nc) Zs[r * ldz + c0 + i] = v[i];
                }
            }
        }
    } else if (!swap) {
        if (32 * q < nr) {                           // lanes = batch rows
            const int r = 32 * q + lane;
            for (int c0 = 32 * half; c0 < nc; c0 += 32 * NH) {
                tmem_ld32(tmem + ((uint32_t)(32 * q) << 16) + (uint32_t)c0, v);
                if (pair) {
                    tmem_ld32(tmem + ((uint32_t)(32 * q) << 16) + (uint32_t)(nc + c0), w);
#pragma unroll
                    for (int i = 0; i < 32; ++i) v[i] += w[i];
                }
                if (r < nr) {
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        if (c0 + i < nc) Zs[r * ldz + c0 + i] = v[i];
                }
            }
        }
    } else {
        if (32 * q < nc) {                           // lanes = gate columns, TMEM columns = batch rows
            const int c = 32 * q + lane;
            for (int r0 = 32 * half; r0 < nr; r0 += 32 * NH) {
                tmem_ld32(tmem + ((uint32_t)(32 * q) << 16) + (uint32_t)r0, v);
                if (c < nc) {
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        if (r0 + i < nr) Zs[(r0 + i) * ldz + c] = v[i];
                }
            }
        }
    }
}


// Element -> (row, 4 consecutive units) mapping shared by all epilogues: a thread owns QUADS of adjacent
// hidden units of one batch row, so every global access is a 16-byte vector (8 bytes for bf16) and a warp's
// accesses cover whole contiguous segments.  Offsets are 32-bit and relative to per-pass base pointers
// (r < 128): r*H + jj into [B,H] tensors, r*gw + jj into the gate stash, r*zw + jj into dz.
struct EpiIdx {
    bool ok;
    int r, jj;       // batch row inside the pass, first unit of the quad inside the slice (multiple of 4)
    int zi;          // r*ldz + (jcol0 + jj)*gstride: Zs index of the quad's first accumulator column
};
template <int VEC>
__device__ __forceinline__ EpiIdx epi_index(int idx, int nquads, int hq, int ldz, int gstride, int jcol0 = 0) {
    EpiIdx e;
    e.ok = idx < nquads;
    e.r = e.ok ? idx / hq : 0;
    e.jj = e.ok ? VEC * (idx - e.r * hq) : 0;
    e.zi = e.r * ldz + (jcol0 + e.jj) * gstride;
    return e;
}
// VEC = 4: 16-byte vectors (8 bytes for bf16); VEC = 1: scalars (the K-split kernels, whose partial sums are read
// from distributed shared memory unit by unit with consecutive lanes on consecutive units)
template <int VEC>
__device__ __forceinline__ void ldv(const float* __restrict__ p, float* dst) {
    if (VEC == 4) { const float4 v = *reinterpret_cast<const float4*>(p); dst[0] = v.x; dst[1] = v.y; dst[2] = v.z; dst[3] = v.w; }
    else dst[0] = p[0];
}
template <int VEC>
__device__ __forceinline__ void stv(float* __restrict__ p, const float* src) {
    if (VEC == 4) *reinterpret_cast<float4*>(p) = make_float4(src[0], src[1], src[2], src[3]);
    else p[0] = src[0];
}
// adjacent bf16 (one 8-byte store for four), plus the lo twins when lo_ofs != 0 (see store_bf16_split).  These are the rows
// other CTAs read while this kernel runs: relaxed gpu-scope stores (strong operations, written through to L2).
template <int VEC>
__device__ __forceinline__ void stv_bf16_split(__nv_bfloat16* __restrict__ p, int64_t lo_ofs, const float* v) {
    if (VEC == 1) {
        const __nv_bfloat16 hi = __float2bfloat16_rn(v[0]);
        st_relaxed_gpu_b16(p, __bfloat16_as_ushort(hi));
        if (lo_ofs) st_relaxed_gpu_b16(p + lo_ofs, __bfloat16_as_ushort(__float2bfloat16_rn(v[0] - __bfloat162float(hi))));
        return;
    }
    const __nv_bfloat162 h0 = __floats2bfloat162_rn(v[0], v[VEC > 1 ? 1 : 0]), h1 = __floats2bfloat162_rn(v[VEC > 2 ? 2 : 0], v[VEC > 3 ? 3 : 0]);
    st_relaxed_gpu_v2(p, *reinterpret_cast<const uint32_t*>(&h0), *reinterpret_cast<const uint32_t*>(&h1));
    if (lo_ofs) {
        const float2 f0 = __bfloat1622float2(h0), f1 = __bfloat1622float2(h1);
        const __nv_bfloat162 l0 = __floats2bfloat162_rn(v[0] - f0.x, v[VEC > 1 ? 1 : 0] - f0.y),
                             l1 = __floats2bfloat162_rn(v[VEC > 2 ? 2 : 0] - f1.x, v[VEC > 3 ? 3 : 0] - f1.y);
        st_relaxed_gpu_v2(p + lo_ofs, *reinterpret_cast<const uint32_t*>(&l0), *reinterpret_cast<const uint32_t*>(&l1));
    }
}

// Accumulator values of a batch of quads: NV consecutive columns per unit, `gstride` columns from unit to unit,
// summed over the K parts (cluster ranks, always in rank order).  All remote loads are issued before any is
// consumed, so their distributed-shared-memory latencies overlap instead of adding up.
template <int NV, int UB, int VEC>
__device__ __forceinline__ void rt_zgather(float (&zv)[UB][4], const EpiIdx (&e)[UB / VEC], int gstride, const float* __restrict__ Zs,
                                           const uint32_t* zpeer, int Q, int kq) {
    if (Q == 1) {
#pragma unroll
        for (int u = 0; u < UB; ++u)
#pragma unroll
            for (int k = 0; k < NV; ++k) zv[u][k] = e[u / VEC].ok ? Zs[e[u / VEC].zi + (u % VEC) * gstride + k] : 0.f;
        return;
    }
    float part[UB][NV][4];
#pragma unroll
    for (int u = 0; u < UB; ++u)
#pragma unroll
        for (int k = 0; k < NV; ++k)
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                float v = 0.f;
                const int zi = e[u / VEC].zi + (u % VEC) * gstride + k;
                if (r < Q && e[u / VEC].ok) v = (r == kq) ? Zs[zi] : ld_dsmem_f32(zpeer[r] + 4u * (uint32_t)zi);
                part[u][k][r] = v;
            }
#pragma unroll
    for (int u = 0; u < UB; ++u)
#pragma unroll
        for (int k = 0; k < NV; ++k) zv[u][k] = ((part[u][k][0] + part[u][k][1]) + part[u][k][2]) + part[u][k][3];
}

// Wait for the accumulator, move it to shared memory, release TMEM to the MMA thread.
#define RT_TAKE_ACC(nr_, nc_)                                                     \
    do {                                                                          \
        if (Q > 1) mbar_wait_cluster(s.zs_free, (p.z_i & 1) ^ 1);  /* peers have read the previous partials */ \
        mbar_wait(s.z_full, p.z_i & 1);                                           \
        tc_fence_after();                                                         \
        zs_dump(tmem, q, ehalf, lane, nr_, nc_, Zs, ldz, a.swap, a.m64, a.split && !a.swap);            \
        tc_fence_before();                                                        \
        mbar_arrive(s.z_empty);                                                   \
        epi_bar_sync();                                                           \
        if (Q > 1) {   /* partial sums of all K parts are in place once every peer has arrived */ \
            if (epi_tid < Q) mbar_arrive_remote(s.zs_ready, (uint32_t)epi_tid);   \
            mbar_wait_cluster(s.zs_ready, p.z_i & 1);                             \
        }                                                                         \
        ++p.z_i;                                                                  \
    } while (0)

// All epilogue threads are done with the peers' partial sums of this pass.
#define RT_RELEASE_PARTIALS()                                                     \
    do {                                                                          \
        if (Q > 1 && epi_tid < Q) mbar_arrive_remote(s.zs_free, (uint32_t)epi_tid); \
    } while (0)

// ------------------------------------------------------------------------------------------
// forward.  Packed weights (pack_rec_weights_kernel): row (slice*GW*HC + lc) of Wr is the column
// of W_h this CTA's local column lc contracts with, K contiguous.  lc = jj*G1 + g for the first
// contraction; GRU candidate rows follow at lc = 2*HC + jj.
// ------------------------------------------------------------------------------------------
template <int MODEL, bool FAST, int UB, int VEC>
__global__ void __launch_bounds__(RT_THREADS, 1) rec_fwd_tc_kernel(const __grid_constant__ CUtensorMap tmH,
                                                                   const __grid_constant__ CUtensorMap tmRS,
                                                                   const __grid_constant__ CUtensorMap tmW0,
                                                                   const __grid_constant__ CUtensorMap tmW1,
                                                                   const __grid_constant__ CUtensorMap tmHl,    // lo twins (split operands)
                                                                   const __grid_constant__ CUtensorMap tmRSl,
                                                                   const __grid_constant__ CUtensorMap tmW0l,
                                                                   const __grid_constant__ CUtensorMap tmW1l,
                                                                   RecTcArgs a) {
    extern __shared__ uint8_t smem_raw[];
    constexpr int G1 = MODEL == CRNN_MODEL_LSTM ? 4 : (MODEL == CRNN_MODEL_GRU ? 2 : 1);
    constexpr int GW = MODEL == CRNN_MODEL_GRU ? 3 : G1;
    const int B = a.B, H = a.H, HC = a.HC, PR = a.pr;
    const int Q = a.ksplit;                           // K parts per unit group (= cluster size when > 1)
    const int HCG = HC * Q;                           // units whose gate columns this CTA contracts
    const int NCA = G1 * HCG;
    const int nkb = ceil_div(H, 64) / Q;              // K blocks of this CTA's part (host: H % (64 Q) == 0 when Q > 1)
    const int nsp = 1 + a.split;                      // operand copies resident / streamed
    const RtSmem s = rt_carve(smem_raw, (size_t)nsp * nkb * NCA * 128, MODEL == CRNN_MODEL_GRU ? (size_t)nsp * nkb * HCG * 128 : 0, a.zs_floats, a.ring_bytes);
    float* Zs = s.zs;
    // life-cycle stamps of CTA 0 (CRNN_TC_DEBUG): entry, set-up done, weights resident, first contraction issued, loop end, exit
    long long* life = (a.dbg && blockIdx.x == 0) ? a.dbg + 4 * 192 : nullptr;
    if (life && threadIdx.x == 0) { life[0] = clock64(); unsigned long long g; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g)); life[8] = (long long)g; }
    const int ldz = NCA + 1;                          // (GRU candidate pass uses the first HCG columns)
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;   // warp-uniform for ptxas
    const int slice = blockIdx.x % a.nslices, tile = blockIdx.x / a.nslices;
    const int grp = slice / Q, kq = slice % Q;        // unit group, K part (== rank in the cluster)
    const int j0 = grp * HCG + kq * HC, hcv = min(HC, H - j0);   // the units this CTA finishes in its epilogue
    const int b0 = tile * a.BT, bt = min(a.BT, B - b0);
    const int64_t gw = (int64_t)a.GS * H, BH = (int64_t)B * H;
    uint32_t tcols = 32;
    while ((int)tcols < (a.swap ? 128 : (a.split ? 2 * NCA + 32 : NCA))) tcols <<= 1;   // split: two column halves (zs_dump reads 32 columns from nc + c0)

    // zero the ring once: rows past the TMA box are never written and must not hold NaNs
    for (int i = threadIdx.x; i < a.ring_bytes / 16; i += RT_THREADS)
        reinterpret_cast<uint4*>(s.ring)[i] = make_uint4(0, 0, 0, 0);
    fence_proxy_async();
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmH); tma_prefetch_desc(&tmW0);
        // (cluster exchange: a_valid[b] is the barrier of operand buffer b: one arrival - the MMA warp's expect_tx - plus the bytes)
        for (int i = 0; i < RT_MAX_STAGES; ++i) { mbar_init(&s.a_full[i], 1); mbar_init(&s.a_empty[i], 1); mbar_init(&s.a_valid[i], a.xchg ? 1u : (uint32_t)RT_EPI_WARPS); }
        mbar_init(s.z_full, 1); mbar_init(s.z_empty, RT_EPI_THREADS); mbar_init(s.w_bar, 1); mbar_init(s.step_bar, (uint32_t)a.nslices); mbar_init(s.zs_ready, (uint32_t)a.ksplit); mbar_init(s.zs_free, (uint32_t)a.ksplit);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(s.tmem_slot, tcols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *s.tmem_slot;
    if (a.csize > 1) cluster_sync_all();             // every CTA's barriers exist (and its ring is zeroed) before any remote arrive / store
    // cluster exchange: two operand buffers in the ring - lstm / rnn: h_{t-1} in buffer t % 2; gru: h_{t-1} in 0, r*h in 1
    const int cx_tile = rt_tile_bytes(a.box_rows);
    const uint32_t cx_buf = (uint32_t)(nsp * nkb * cx_tile);
    if (life && threadIdx.x == 0) life[1] = clock64();
    int dbg_slot = -1;
    if (a.dbg) dbg_slot = blockIdx.x == 0 ? 0 : blockIdx.x == 1 ? 1 : blockIdx.x == gridDim.x / 2 ? 2 : blockIdx.x == gridDim.x - 1 ? 3 : -1;
    long long* dbgs = dbg_slot >= 0 ? s.dbg : nullptr;

    RtPipe p{0, 0, 0, 0};
    unsigned phases_done = 0;
    const int nphase = MODEL == CRNN_MODEL_GRU ? 2 : 1;
    const int rot = (a.rotate && nkb % a.ckb == 0) ? (grp % (nkb / a.ckb)) * a.ckb : 0;

    if (warp == 0) {
        if (lane == 0) {
            // resident weights: once
            uint32_t wbytes = (uint32_t)nkb * NCA * 128u + (MODEL == CRNN_MODEL_GRU ? (uint32_t)nkb * HCG * 128u : 0u);
            mbar_expect_tx(s.w_bar, wbytes * (uint32_t)nsp);
            for (int half = 0; half < nsp; ++half)
                for (int kb = 0; kb < nkb; ++kb) {
                    // (not swapped: [hi rows ; lo rows] per K block, see rt_mma; swapped: all hi blocks, then all lo blocks)
                    const int wblk = (nsp == 2 && !a.swap) ? kb * 2 + half : half * nkb + kb;
                    tma_load_2d(s.wsa + (size_t)wblk * NCA * 128, half ? &tmW0l : &tmW0, s.w_bar, (kq * nkb + kb) * 64, grp * GW * HCG);
                    if (MODEL == CRNN_MODEL_GRU)
                        tma_load_2d(s.wsb + (size_t)wblk * HCG * 128, half ? &tmW1l : &tmW1, s.w_bar, (kq * nkb + kb) * 64, grp * GW * HCG + NCA);
                }
        }
        if (a.xchg) {
            // Only the launch's first contraction reads rows from global memory (an earlier launch or the host left them
            // there); every later one finds its rows in shared memory, delivered by the slices of the cluster.
            if (lane == 0) {
                const int bidx = MODEL == CRNN_MODEL_GRU ? 0 : (a.t_begin & 1);
                uint8_t* dst = s.ring + (size_t)bidx * cx_buf;      // (the MMA warp expects these bytes on the buffer's barrier)
                for (int kb = 0; kb < nkb; ++kb) {
                    tma_load_2d(dst + (size_t)kb * cx_tile, &tmH, &s.a_valid[bidx], kb * 64, a.t_begin * B + b0);
                    if (a.split) tma_load_2d(dst + (size_t)(nkb + kb) * cx_tile, &tmHl, &s.a_valid[bidx], kb * 64, a.t_begin * B + b0);
                }
            }
        } else
        for (int t = a.t_begin; t < a.t_end; ++t) {
            for (int ph = 0; ph < nphase; ++ph) {
                const int dbg_phase = (int)phases_done;
                if (lane == 0) {
                    RT_STAMP(0);
                    if (!a.gather && !a.sub && phases_done > 0) { if (a.cbar) rt_await_cluster(s.step_bar, phases_done); else rt_await(a.barrier + tile, phases_done * a.nslices, a.acq_poll); }
                    RT_STAMP(1);
                }
                for (int ps = 0; ps < a.npass; ++ps) {
                    if (a.gather && phases_done > 0) {    // (the launch's first phase reads rows an earlier launch or the host wrote)
                        const bool second = MODEL == CRNN_MODEL_GRU && ph == 1;
                        rt_probe(second ? a.rsb : a.hxb, second ? a.rsb_lo : a.hxb_lo, (int64_t)t * B + b0 + ps * PR + min(PR, bt - ps * PR) - 1, H, 0, HC, H,
                                 a.nslices, lane, a.err);
                    }
                    if (lane == 0) {
                        // sub-chains: pass ps of this phase only needs pass ps of the previous phase from the other slices
                        if (a.sub && !a.gather && phases_done > 0) rt_await(a.barrier + tile * a.npass + ps, phases_done * a.nslices, a.acq_poll);
                        rt_produce(s, p, ph == 0 ? &tmH : &tmRS, t * B + b0 + ps * PR, nkb, a.box_rows, a.ckb, a.nst, a.use3d, kq * nkb,
                                   a.split ? (ph == 0 ? &tmHl : &tmRSl) : nullptr, rot,
                                   (!a.gather && dbgs && dbg_phase >= 20 && dbg_phase < 28) ? dbgs + (dbg_phase - 20) * 24 : nullptr,
                                   (a.expt & 64) ? a.hxb + ((int64_t)t * B + b0 + ps * PR) * H : nullptr,
                                   (a.expt & 64) ? a.hxb + a.hxb_lo + ((int64_t)t * B + b0 + ps * PR) * H : nullptr);
                    }
                }
                if (lane == 0) RT_STAMP(2);
                ++phases_done;
            }
        }
    } else if (warp == 1) {
        {   // the whole warp runs the issue loop (see mma_bf16_elect)
            const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem, 0);
            mbar_wait(s.w_bar, 0);
            if (life && lane == 0) life[2] = clock64();
            if (a.xchg) {
                uint32_t uses[2] = {0u, 0u};
                for (int t = a.t_begin; t < a.t_end; ++t)
                    for (int ph = 0; ph < nphase; ++ph) {
                        if (life && lane == 0 && t == a.t_begin + 1 && ph == 0) life[3] = clock64();
                        const int bidx = MODEL == CRNN_MODEL_GRU ? ph : (t & 1);
                        const int dbg_phase = (t - a.t_begin) * nphase + ph;
                        // bytes of a complete buffer: the launch's first one is TMA boxes, the others bt rows x H units from the slices
                        const uint32_t nbytes = (t == a.t_begin && ph == 0) ? (uint32_t)a.box_rows * 128u * (uint32_t)(nkb * nsp)
                                                                            : (uint32_t)(nsp * bt) * (uint32_t)H * 2u;
                        rt_mma_cx(s, p, tmem_u, ph == 0 ? s.wsa : s.wsb, ph == 0 ? NCA : HCG, nkb, s.ring + (size_t)bidx * cx_buf, cx_tile,
                                  &s.a_valid[bidx], uses[bidx] & 1u, nbytes, a.m64, a.split,
                                  (dbgs && dbg_phase >= 20 && dbg_phase < 28) ? dbgs + (dbg_phase - 20) * 24 : nullptr);
                        ++uses[bidx];
                    }
            } else
            for (int t = a.t_begin; t < a.t_end; ++t)
                for (int ph = 0; ph < nphase; ++ph) {
                    const int dbg_phase = (t - a.t_begin) * nphase + ph;
                    for (int ps = 0; ps < a.npass; ++ps)
                        rt_mma(s, p, tmem_u, ph == 0 ? s.wsa : s.wsb, ph == 0 ? NCA : HCG, nkb, a.swap, (min(PR, bt - ps * PR) + 15) & ~15, a.box_rows, a.ckb, a.nst, a.m64,
                               (dbgs && dbg_phase >= 20 && dbg_phase < 28) ? dbgs + (dbg_phase - 20) * 24 : nullptr, a.split, rot, a.gather, a.expt & 32);
                    if (lane == 0) RT_STAMP(3);
                }
        }
    } else {
        const int q = warp & 3;                       // TMEM lane quarter this warp may read
        const int ehalf = (warp - 2) >> 2;            // which of the quarter's warps this is
        const int epi_tid = (warp - 2) * 32 + lane;
        uint32_t zpeer[4] = {0, 0, 0, 0};            // the peers' accumulator staging areas (distributed shared memory)
        if (Q > 1) for (int r = 0; r < Q; ++r) zpeer[r] = dsmem_addr(Zs, (uint32_t)r);
        uint32_t pbase[RT_CX_MAX];                    // cluster exchange: every slice's ring
#pragma unroll
        for (int r = 0; r < RT_CX_MAX; ++r) pbase[r] = (a.xchg && r < a.csize) ? dsmem_addr(s.ring, (uint32_t)r) : 0u;
        uint32_t pbar[RT_CX_MAX];                     // ... and its operand-buffer barriers
#pragma unroll
        for (int r = 0; r < RT_CX_MAX; ++r) pbar[r] = (a.xchg && r < a.csize) ? dsmem_addr(s.a_valid, (uint32_t)r) : 0u;
        if (a.gather) {
            // the slots this CTA writes in the NEXT launch of the layer's chain (visible there by the kernel boundary)
            for (int t = a.fill_t0; t < a.fill_t1; ++t) {
                rt_fill_sentinel(a.hxb, a.hxb_lo, (int64_t)(t + 1) * B + b0, bt, H, j0, hcv, epi_tid);
                if (MODEL == CRNN_MODEL_GRU) rt_fill_sentinel(a.rsb, a.rsb_lo, (int64_t)t * B + b0, bt, H, j0, hcv, epi_tid);
            }
        }
        const RtGatherGeom gg{smem_u32(s.ring), smem_u32(s.a_full), smem_u32(s.a_valid), nsp, rt_tile_bytes(a.box_rows), a.g_ckb_log2, a.g_rp_log2, a.nst, a.expt, a.err};
        uint32_t gpos = 0;                            // ring position of the gather (stage | phase bit << 16)
        int pck = 0;                                  // persistent form: chunk in progress
        bool pdead = false;
        for (int t = a.t_begin; t < a.t_end; ++t) {
            if (a.pchunks && t == rt_chunk_t0(a.S, pck, a.pchunks)) {   // first step of a chunk: its hoisted x-projection must be in place
                if (a.in_flag && epi_tid == 0 && !pdead) pdead = !rt_wait_chunk(a.in_flag, (unsigned)pck + 1u, a.err);
                epi_bar_sync();
            }
            const int64_t tB = (int64_t)t * B;
            float* gt = a.gates;
            const float* hprev = a.hx + (int64_t)t * BH;
            float* hnext = a.hx + (int64_t)(t + 1) * BH;
            __nv_bfloat16* hnb = a.hxb + (int64_t)(t + 1) * BH;
            for (int ph = 0; ph < nphase; ++ph) {
                // The batch in flight.  Declared here so that the LAST batch of the phase can keep its results in
                // registers across the step barrier: only the bf16 rows other CTAs stream next (hxb / rsb) are
                // stored before the release, everything this CTA alone reads later (fp32 h, c, the gate stash for
                // BPTT) is stored after it - the release waits for every store issued before it.
                EpiIdx e[UB / VEC];
                float x0[UB], x1[UB], x2[UB], x3[UB], ax[UB], hv[UB];
                bool pending = false;
                const int hq = hcv / VEC;                                        // vectors per row of the slice
                for (int ps = 0; ps < a.npass; ++ps) {
                    const int nr = min(PR, bt - ps * PR);
                    const int nquads = nr * hq;
                    const int b0row = b0 + ps * PR;
                    if (a.gather) {   // this pass's rows: h_{t-1} (slot t of hxb) or r*h of this step, as the other slices finish them
                        const int gph = (t - a.t_begin) * nphase + ph;
                        long long* gstamp = (dbgs && gph >= 20 && gph < 28) ? dbgs + (gph - 20) * 24 : nullptr;
                        const bool second = MODEL == CRNN_MODEL_GRU && ph == 1;
                        gpos = rt_validate(gg, gpos, second ? a.rsb : a.hxb, second ? a.rsb_lo : a.hxb_lo, tB + b0row, nr, H, 0, H, nkb, epi_tid, lane, gstamp);
                    }
                    const int gstride = ph == 0 ? G1 : 1;
                    const int64_t ob = (int64_t)b0row * H + j0;                  // [B,H] tensors
                    float* gp = gt + (tB + b0row) * gw + j0;                     // gate stash rows of this pass
                    const float* hp_ = hprev + ob;
                    float* hn_ = hnext + ob;
                    __nv_bfloat16* hb_ = hnb + ob;
                    const float* cp_ = MODEL == CRNN_MODEL_LSTM ? a.cx + (int64_t)t * BH + ob : nullptr;
                    float* cn_ = MODEL == CRNN_MODEL_LSTM ? a.cx + (int64_t)(t + 1) * BH + ob : nullptr;
                    __nv_bfloat16* rs_ = MODEL == CRNN_MODEL_GRU ? a.rsb + (tB + b0row) * H + j0 : nullptr;
                    for (int base = 0; base < nquads; base += RT_EPI_THREADS * (UB / VEC)) {
                        const bool defer = a.defer && (a.sub || ps == a.npass - 1) && base + RT_EPI_THREADS * (UB / VEC) >= nquads;
                        // ---- issue every global load of this batch before touching the accumulator
#pragma unroll
                        for (int u = 0; u < UB; ++u) x0[u] = x1[u] = x2[u] = x3[u] = ax[u] = hv[u] = 0.f;
#pragma unroll
                        for (int qd = 0; qd < UB / VEC; ++qd) {
                            e[qd] = epi_index<VEC>(base + qd * RT_EPI_THREADS + epi_tid, nquads, hq, ldz, gstride, kq * HC);
                            if (e[qd].ok) {
                                const float* g = gp + e[qd].r * (int)gw + e[qd].jj;
                                const int o = e[qd].r * H + e[qd].jj;
                                if (MODEL == CRNN_MODEL_LSTM) {
                                    ldv<VEC>(g, x0 + VEC * qd); ldv<VEC>(g + H, x1 + VEC * qd); ldv<VEC>(g + 2 * H, x2 + VEC * qd); ldv<VEC>(g + 3 * H, x3 + VEC * qd);
                                    ldv<VEC>(cp_ + o, ax + VEC * qd);
                                } else if (MODEL == CRNN_MODEL_RNN) {
                                    ldv<VEC>(g, x0 + VEC * qd);
                                } else if (ph == 0) {
                                    ldv<VEC>(g, x0 + VEC * qd); ldv<VEC>(g + H, x1 + VEC * qd); ldv<VEC>(hp_ + o, ax + VEC * qd);
                                } else {
                                    ldv<VEC>(g + 2 * H, x0 + VEC * qd); ldv<VEC>(g + H, x1 + VEC * qd); ldv<VEC>(hp_ + o, ax + VEC * qd);
                                }
                            }
                        }
                        const int dbg_phase = (t - a.t_begin) * nphase + ph;
                        if (epi_tid == 0 && base == 0) RT_STAMP(4);
                        if (base == 0) {
                            if (dbgs) { mbar_wait(s.z_full, p.z_i & 1); if (epi_tid == 0) RT_STAMP(14); if (epi_tid == RT_EPI_THREADS - 1) RT_STAMP(15); }
                            RT_TAKE_ACC(nr, ph == 0 ? NCA : HCG);
                        }
                        if (epi_tid == 0 && base == 0) RT_STAMP(5);
                        float zv[UB][4];
                        if (ph == 0) rt_zgather<G1, UB, VEC>(zv, e, gstride, Zs, zpeer, Q, kq);
                        else         rt_zgather<1, UB, VEC>(zv, e, gstride, Zs, zpeer, Q, kq);
                        if (dbgs && epi_tid == 0 && base == 0) {
                            asm volatile("" ::"f"(zv[0][0]), "f"(zv[UB - 1][0]));
                            RT_STAMP(16);                                          // accumulator values in registers
                            asm volatile("" ::"f"(x0[0]), "f"(x0[UB - 1]), "f"(x1[0]), "f"(x1[UB - 1]));
                            RT_STAMP(17);                                          // gate-stash loads have landed
                            asm volatile("" ::"f"(ax[0]), "f"(ax[UB - 1]));
                            RT_STAMP(18);                                          // previous state has landed
                        }
                        // ---- cell update; results replace the inputs in x0..x3 / ax / hv.  No branches in the math
                        //      (inactive elements compute on zeros), so the UB dependency chains interleave.
#pragma unroll
                        for (int u = 0; u < UB; ++u) {
                            if (a.expt & 2) { hv[u] = zv[u][0] + x0[u] + x1[u] + ax[u]; continue; }
                            if (MODEL == CRNN_MODEL_LSTM) {
                                x0[u] = fast_sigmoid(zv[u][0] + x0[u]);                   // i
                                x1[u] = fast_tanh(zv[u][1] + x1[u]);                      // j
                                x2[u] = fast_sigmoid(zv[u][2] + x2[u] + 1.0f);            // f (forget_bias)
                                x3[u] = fast_sigmoid(zv[u][3] + x3[u]);                   // o
                                ax[u] = fmaf(x2[u], ax[u], x0[u] * x1[u]);                // c'
                                hv[u] = x3[u] * fast_tanh(ax[u]);
                            } else if (MODEL == CRNN_MODEL_RNN) {
                                hv[u] = fast_tanh(zv[u][0] + x0[u]);
                            } else if (ph == 0) {
                                x0[u] = fast_sigmoid(zv[u][0] + x0[u]);                   // r
                                x1[u] = fast_sigmoid(zv[u][1] + x1[u]);                   // u
                                hv[u] = x0[u] * ax[u];                                    // reset BEFORE the matmul
                            } else {
                                x0[u] = fast_tanh(zv[u][0] + x0[u]);                      // candidate
                                hv[u] = fmaf(x1[u], ax[u] - x0[u], x0[u]);                // u*h + (1-u)*cand
                            }
                        }
                        if (dbgs && epi_tid == 0 && base == 0) { asm volatile("" ::"f"(hv[0]), "f"(hv[UB - 1])); RT_STAMP(21); }   // cell math done
                        // the rows other CTAs stream in the next phase
                        const bool to_rs = MODEL == CRNN_MODEL_GRU && ph == 0;
                        // (cluster exchange: h of the launch's last step is read from global memory by the next launch)
                        const bool cx_send = a.xchg && (to_rs || t + 1 < a.t_end);
                        const int cx_b = to_rs ? 1 : (MODEL == CRNN_MODEL_GRU ? 0 : ((t + 1) & 1));
                        const uint32_t cx_dst = (uint32_t)cx_b * cx_buf;
#pragma unroll
                        for (int qd = 0; qd < UB / VEC; ++qd) {
                            if (!e[qd].ok || ((a.expt & 1) && !a.xchg)) continue;
                            const int o = e[qd].r * H + e[qd].jj;
                            if (a.xchg) {
                                if (cx_send && VEC == 4) cx_push(pbase, pbar, a.csize, cx_b, cx_dst, a.split ? (uint32_t)(nkb * cx_tile) : 0u, cx_tile, e[qd].r, j0 + e[qd].jj, hv + VEC * qd);
                            } else if (to_rs) stv_bf16_split<VEC>(rs_ + o, a.rsb_lo, hv + VEC * qd);
                            else              stv_bf16_split<VEC>(hb_ + o, a.hxb_lo, hv + VEC * qd);
                        }
                        pending = defer;
                        if (!defer) {
#pragma unroll
                            for (int qd = 0; qd < UB / VEC; ++qd) {
                                if (!e[qd].ok) continue;
                                float* g = gp + e[qd].r * (int)gw + e[qd].jj;
                                const int o = e[qd].r * H + e[qd].jj;
                                if (a.xchg) {          // the global copies (GEMM operands, the next launch): nobody in this launch waits for them
                                    if (to_rs) stv_bf16_split<VEC>(rs_ + o, a.rsb_lo, hv + VEC * qd);
                                    else       stv_bf16_split<VEC>(hb_ + o, a.hxb_lo, hv + VEC * qd);
                                }
                                if (MODEL == CRNN_MODEL_LSTM) {
                                    stv<VEC>(hn_ + o, hv + VEC * qd);
                                    stv<VEC>(cn_ + o, ax + VEC * qd);
                                    stv<VEC>(g, x0 + VEC * qd); stv<VEC>(g + H, x1 + VEC * qd); stv<VEC>(g + 2 * H, x2 + VEC * qd); stv<VEC>(g + 3 * H, x3 + VEC * qd);
                                } else if (MODEL == CRNN_MODEL_RNN) {
                                    stv<VEC>(hn_ + o, hv + VEC * qd);
                                } else if (ph == 0) {
                                    stv<VEC>(g, x0 + VEC * qd); stv<VEC>(g + H, x1 + VEC * qd); stv<VEC>(g + 3 * H, hv + VEC * qd);
                                } else {
                                    stv<VEC>(hn_ + o, hv + VEC * qd);
                                    stv<VEC>(g + 2 * H, x0 + VEC * qd);
                                }
                            }
                        }
                    }
                    { const int dbg_phase = (t - a.t_begin) * nphase + ph; if (epi_tid == 0) RT_STAMP(8); if (epi_tid == RT_EPI_THREADS - 1) RT_STAMP(9); }
                    epi_bar_sync();          // Zs is reused by the next pass / phase
                    RT_RELEASE_PARTIALS();
                    if (!a.sub && ps + 1 < a.npass) continue;
                    // ---- the pass (sub-chains) or the whole phase is done on this CTA: release it, then the deferred stores
                const int dbg_phase = (t - a.t_begin) * nphase + ph;
                if (epi_tid == 0) RT_STAMP(6);
                if (a.xchg) { /* the pushed rows count themselves on the receivers' barriers */ }
                else if (a.gather) { /* the stored rows are their own signal */ }
                else if (a.cbar) rt_publish_cluster(s.step_bar, epi_tid, a.csize);
                else rt_publish(a.sub ? a.barrier + tile * a.npass + ps : a.barrier + tile, epi_tid, a.sub);
                if (epi_tid == 0) RT_STAMP(7);
                if (pending && !(a.expt & 4)) {   // the stores nobody else waits for
                    const int b0row = b0 + ps * PR;                              // the deferred batch belongs to this (the last) pass
                    const int64_t ob = (int64_t)b0row * H + j0;
#pragma unroll
                    for (int qd = 0; qd < UB / VEC; ++qd) {
                        if (!e[qd].ok) continue;
                        float* g = gt + (tB + b0row) * gw + j0 + e[qd].r * (int)gw + e[qd].jj;
                        const int64_t o = ob + e[qd].r * H + e[qd].jj;
                        if (a.xchg) {
                            if (MODEL == CRNN_MODEL_GRU && ph == 0) stv_bf16_split<VEC>(a.rsb + tB * H + o, a.rsb_lo, hv + VEC * qd);
                            else                                    stv_bf16_split<VEC>(hnb + o, a.hxb_lo, hv + VEC * qd);
                        }
                        if (MODEL == CRNN_MODEL_LSTM) {
                            stv<VEC>(hnext + o, hv + VEC * qd);
                            stv<VEC>(a.cx + (int64_t)(t + 1) * BH + o, ax + VEC * qd);
                            stv<VEC>(g, x0 + VEC * qd); stv<VEC>(g + H, x1 + VEC * qd); stv<VEC>(g + 2 * H, x2 + VEC * qd); stv<VEC>(g + 3 * H, x3 + VEC * qd);
                        } else if (MODEL == CRNN_MODEL_RNN) {
                            stv<VEC>(hnext + o, hv + VEC * qd);
                        } else if (ph == 0) {
                            stv<VEC>(g, x0 + VEC * qd); stv<VEC>(g + H, x1 + VEC * qd); stv<VEC>(g + 3 * H, hv + VEC * qd);
                        } else {
                            stv<VEC>(hnext + o, hv + VEC * qd);
                            stv<VEC>(g + 2 * H, x0 + VEC * qd);
                        }
                    }
                }
                pending = false;
                if (epi_tid == 0) RT_STAMP(13);
                }   // passes
            }
            if (a.pchunks && t + 1 == rt_chunk_t0(a.S, pck + 1, a.pchunks)) {   // the chunk is complete on this CTA (all its stores issued)
                epi_bar_sync();
                if (epi_tid == 0 && a.out_flag) asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(a.out_flag + pck) : "memory");
                ++pck;
            }
        }
    }
    if (life && warp == 1 && lane == 0) life[4] = clock64();
    if (life && warp == 2 && lane == 0) life[7] = clock64();
    __syncthreads();
    if (life && threadIdx.x == 0) life[5] = clock64();
    if (dbgs) for (int i = threadIdx.x; i < 192; i += RT_THREADS) a.dbg[dbg_slot * 192 + i] = dbgs[i];
    if (a.csize > 1) cluster_sync_all();             // no CTA leaves while a peer may still signal its barriers
    if (life && threadIdx.x == 0) { life[6] = clock64(); unsigned long long g; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(g)); life[9] = (long long)g; }
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem, tcols);
    }
}

// ------------------------------------------------------------------------------------------
// backward.  dh_t = dy_t + dz_{t+1} * W_h^T: the B operand is rows H+j0.. of the bf16 copy of
// the cell kernel itself (K = gate columns, contiguous) - no packing needed.
// ------------------------------------------------------------------------------------------
template <int MODEL, bool FAST, int UB, int VEC>
__global__ void __launch_bounds__(RT_THREADS, 1) rec_bwd_tc_kernel(const __grid_constant__ CUtensorMap tmZ0,   // dzb, first contraction
                                                                   const __grid_constant__ CUtensorMap tmZ1,   // gru: dzb candidate columns
                                                                   const __grid_constant__ CUtensorMap tmW0,
                                                                   const __grid_constant__ CUtensorMap tmW1,
                                                                   const __grid_constant__ CUtensorMap tmZ0l,   // lo twins (split operands)
                                                                   const __grid_constant__ CUtensorMap tmZ1l,
                                                                   const __grid_constant__ CUtensorMap tmW0l,
                                                                   const __grid_constant__ CUtensorMap tmW1l,
                                                                   RecTcArgs a) {
    extern __shared__ uint8_t smem_raw[];
    constexpr int G1 = MODEL == CRNN_MODEL_LSTM ? 4 : (MODEL == CRNN_MODEL_GRU ? 2 : 1);
    constexpr int GZ = MODEL == CRNN_MODEL_LSTM ? 4 : (MODEL == CRNN_MODEL_GRU ? 3 : 1);
    const int B = a.B, H = a.H, HC = a.HC, PR = a.pr;
    const int Q = a.ksplit;
    const int HCG = HC * Q;
    const int nkb0 = ceil_div(G1 * H, 64) / Q;        // this CTA's part of the gate-column contraction
    const int nkb1 = ceil_div(H, 64) / Q;
    const int nsp = 1 + a.split;
    const RtSmem s = rt_carve(smem_raw, (size_t)nsp * nkb0 * HCG * 128, MODEL == CRNN_MODEL_GRU ? (size_t)nsp * nkb1 * HCG * 128 : 0, a.zs_floats, a.ring_bytes);
    float* Zs = s.zs;
    const int ldz = HCG + 1;
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;   // warp-uniform for ptxas
    const int slice = blockIdx.x % a.nslices, tile = blockIdx.x / a.nslices;
    const int grp = slice / Q, kq = slice % Q;
    const int jg0 = grp * HCG;                        // first unit of the group whose dh this CTA contracts
    const int j0 = jg0 + kq * HC, hcv = min(HC, H - j0);
    const int b0 = tile * a.BT, bt = min(a.BT, B - b0);
    const int64_t gw = (int64_t)a.GS * H, zw = (int64_t)GZ * H, BH = (int64_t)B * H;
    uint32_t tcols = 32;
    while ((int)tcols < (a.swap ? 128 : (a.split ? 2 * HCG + 32 : HCG))) tcols <<= 1;

    for (int i = threadIdx.x; i < a.ring_bytes / 16; i += RT_THREADS)
        reinterpret_cast<uint4*>(s.ring)[i] = make_uint4(0, 0, 0, 0);
    fence_proxy_async();
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmZ0); tma_prefetch_desc(&tmW0);
        // (cluster exchange: a_valid[0..1] are the barriers of the two dz_{t+1} buffers - the gru delivers them in two waves,
        //  du in the first phase of step t+1 and dr in its second -, a_valid[2] the gru's candidate-column buffer)
        for (int i = 0; i < RT_MAX_STAGES; ++i) { mbar_init(&s.a_full[i], 1); mbar_init(&s.a_empty[i], 1); mbar_init(&s.a_valid[i], a.xchg ? 1u : (uint32_t)RT_EPI_WARPS); }
        mbar_init(s.z_full, 1); mbar_init(s.z_empty, RT_EPI_THREADS); mbar_init(s.w_bar, 1); mbar_init(s.step_bar, (uint32_t)a.nslices); mbar_init(s.zs_ready, (uint32_t)a.ksplit); mbar_init(s.zs_free, (uint32_t)a.ksplit);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(s.tmem_slot, tcols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *s.tmem_slot;
    if (a.csize > 1) cluster_sync_all();             // every CTA's barriers exist (and its ring is zeroed) before any remote arrive / store
    // cluster exchange: dz_{t+1} (gate columns [0, G1*H)) in buffer t % 2 of two; gru: this step's candidate columns in a third
    const int cx_tile = rt_tile_bytes(a.box_rows);
    const uint32_t cx_buf0 = (uint32_t)(nsp * nkb0 * cx_tile), cx_buf1 = (uint32_t)(nsp * nkb1 * cx_tile);

    RtPipe p{0, 0, 0, 0};
    unsigned phases_done = 0;
    const bool rot_on = a.rotate != 0;
    const int rot0 = (rot_on && nkb0 % a.ckb == 0) ? (grp % (nkb0 / a.ckb)) * a.ckb : 0;
    const int rot1 = (rot_on && nkb1 % a.ckb == 0) ? (grp % (nkb1 / a.ckb)) * a.ckb : 0;

    if (warp == 0) {
        if (lane == 0) {
            uint32_t wbytes = (uint32_t)nkb0 * HCG * 128u + (MODEL == CRNN_MODEL_GRU ? (uint32_t)nkb1 * HCG * 128u : 0u);
            mbar_expect_tx(s.w_bar, wbytes * (uint32_t)nsp);
            for (int half = 0; half < nsp; ++half) {
                const bool il = nsp == 2 && !a.swap;         // [hi rows ; lo rows] per K block, see rt_mma
                for (int kb = 0; kb < nkb0; ++kb)
                    tma_load_2d(s.wsa + (size_t)(il ? kb * 2 + half : half * nkb0 + kb) * HCG * 128, half ? &tmW0l : &tmW0, s.w_bar, (kq * nkb0 + kb) * 64, H + jg0);
                if (MODEL == CRNN_MODEL_GRU)
                    for (int kb = 0; kb < nkb1; ++kb)
                        tma_load_2d(s.wsb + (size_t)(il ? kb * 2 + half : half * nkb1 + kb) * HCG * 128, half ? &tmW1l : &tmW1, s.w_bar, (kq * nkb1 + kb) * 64, H + jg0);
            }
        }
        if (a.xchg) {
            // only the launch's first contraction reads dz_{t+1} from global memory (the previous launch wrote it)
            const int t = a.t_end - 1;
            if (lane == 0 && t + 1 < a.S) {
                uint8_t* dst = s.ring + (size_t)(t & 1) * cx_buf0;      // (the MMA warp expects these bytes on the buffer's barrier)
                for (int kb = 0; kb < nkb0; ++kb) {
                    tma_load_2d(dst + (size_t)kb * cx_tile, &tmZ0, &s.a_valid[t & 1], kb * 64, (t + 1) * B + b0);
                    if (a.split) tma_load_2d(dst + (size_t)(nkb0 + kb) * cx_tile, &tmZ0l, &s.a_valid[t & 1], kb * 64, (t + 1) * B + b0);
                }
            }
        } else
        for (int t = a.t_end - 1; t >= a.t_begin; --t) {
            // contraction 0 reads dz_{t+1} (all CTAs, previous step; at a chunk boundary: previous launch)
            if (t + 1 < a.S) {
                if (lane == 0 && !a.gather && !a.sub && phases_done > 0) { if (a.cbar) rt_await_cluster(s.step_bar, phases_done); else rt_await(a.barrier + tile, phases_done * a.nslices, a.acq_poll); }
                for (int ps = 0; ps < a.npass; ++ps) {
                    if (a.gather && phases_done > 0)
                        rt_probe(a.dzb, a.dzb_lo, (int64_t)(t + 1) * B + b0 + ps * PR + min(PR, bt - ps * PR) - 1, (int)zw, 0, HC, H, a.nslices, lane, a.err);
                    if (lane == 0) {
                        if (a.sub && !a.gather && phases_done > 0) rt_await(a.barrier + tile * a.npass + ps, phases_done * a.nslices, a.acq_poll);
                        rt_produce(s, p, &tmZ0, (t + 1) * B + b0 + ps * PR, nkb0, a.box_rows, a.ckb, a.nst, a.use3d, kq * nkb0,
                                   a.split ? &tmZ0l : nullptr, rot0, nullptr,
                                   (a.expt & 64) ? a.dzb + ((int64_t)t * B + b0 + ps * PR) * zw : nullptr,
                                   (a.expt & 64) ? a.dzb + a.dzb_lo + ((int64_t)t * B + b0 + ps * PR) * zw : nullptr);
                    }
                }
            }
            ++phases_done;
            if (MODEL == CRNN_MODEL_GRU) {
                if (lane == 0 && !a.gather && !a.sub) { if (a.cbar) rt_await_cluster(s.step_bar, phases_done); else rt_await(a.barrier + tile, phases_done * a.nslices, a.acq_poll); }
                for (int ps = 0; ps < a.npass; ++ps) {
                    if (a.gather) rt_probe(a.dzb + 2 * H, a.dzb_lo, (int64_t)t * B + b0 + ps * PR + min(PR, bt - ps * PR) - 1, (int)zw, 0, HC, H, a.nslices, lane, a.err);
                    if (lane == 0) {
                        if (a.sub && !a.gather) rt_await(a.barrier + tile * a.npass + ps, phases_done * a.nslices, a.acq_poll);
                        rt_produce(s, p, &tmZ1, t * B + b0 + ps * PR, nkb1, a.box_rows, a.ckb, a.nst, a.use3d, kq * nkb1,
                                   a.split ? &tmZ1l : nullptr, rot1, nullptr,
                                   (a.expt & 64) ? a.dzb + ((int64_t)t * B + b0 + ps * PR) * zw : nullptr,
                                   (a.expt & 64) ? a.dzb + a.dzb_lo + ((int64_t)t * B + b0 + ps * PR) * zw : nullptr);
                    }
                }
                ++phases_done;
            }
        }
    } else if (warp == 1) {
        {   // the whole warp runs the issue loop (see mma_bf16_elect)
            const uint32_t tmem_u = __shfl_sync(0xffffffffu, tmem, 0);
            mbar_wait(s.w_bar, 0);
            if (a.xchg) {
                uint32_t uses[3] = {0u, 0u, 0u};
                for (int t = a.t_end - 1; t >= a.t_begin; --t) {
                    if (t + 1 < a.S) {
                        const uint32_t nbytes = t == a.t_end - 1 ? (uint32_t)a.box_rows * 128u * (uint32_t)(nkb0 * nsp)      // TMA boxes
                                                                 : (uint32_t)(nsp * bt) * (uint32_t)(G1 * H) * 2u;           // bt rows x G1*H gate columns
                        rt_mma_cx(s, p, tmem_u, s.wsa, HCG, nkb0, s.ring + (size_t)(t & 1) * cx_buf0, cx_tile, &s.a_valid[t & 1], uses[t & 1] & 1u, nbytes, a.m64, a.split);
                        ++uses[t & 1];
                    }
                    if (MODEL == CRNN_MODEL_GRU) {
                        rt_mma_cx(s, p, tmem_u, s.wsb, HCG, nkb1, s.ring + 2 * (size_t)cx_buf0, cx_tile, &s.a_valid[2], uses[2] & 1u,
                                  (uint32_t)(nsp * bt) * (uint32_t)H * 2u, a.m64, a.split);
                        ++uses[2];
                    }
                }
            } else
            for (int t = a.t_end - 1; t >= a.t_begin; --t) {
                if (t + 1 < a.S)
                    for (int ps = 0; ps < a.npass; ++ps) rt_mma(s, p, tmem_u, s.wsa, HCG, nkb0, a.swap, (min(PR, bt - ps * PR) + 15) & ~15, a.box_rows, a.ckb, a.nst, a.m64, nullptr, a.split, rot0, a.gather, a.expt & 32);
                if (MODEL == CRNN_MODEL_GRU)
                    for (int ps = 0; ps < a.npass; ++ps) rt_mma(s, p, tmem_u, s.wsb, HCG, nkb1, a.swap, (min(PR, bt - ps * PR) + 15) & ~15, a.box_rows, a.ckb, a.nst, a.m64, nullptr, a.split, rot1, a.gather, a.expt & 32);
            }
        }
    } else {
        const int q = warp & 3;                       // TMEM lane quarter this warp may read
        const int ehalf = (warp - 2) >> 2;            // which of the quarter's warps this is
        const int epi_tid = (warp - 2) * 32 + lane;
        uint32_t zpeer[4] = {0, 0, 0, 0};
        if (Q > 1) for (int r = 0; r < Q; ++r) zpeer[r] = dsmem_addr(Zs, (uint32_t)r);
        uint32_t pbase[RT_CX_MAX];                    // cluster exchange: every slice's ring
#pragma unroll
        for (int r = 0; r < RT_CX_MAX; ++r) pbase[r] = (a.xchg && r < a.csize) ? dsmem_addr(s.ring, (uint32_t)r) : 0u;
        uint32_t pbar[RT_CX_MAX];                     // ... and its operand-buffer barriers
#pragma unroll
        for (int r = 0; r < RT_CX_MAX; ++r) pbar[r] = (a.xchg && r < a.csize) ? dsmem_addr(s.a_valid, (uint32_t)r) : 0u;
        const uint32_t cx_lo0 = a.split ? (uint32_t)(nkb0 * cx_tile) : 0u, cx_lo1 = a.split ? (uint32_t)(nkb1 * cx_tile) : 0u;
        if (a.gather) {
            // the dz slots this CTA writes in the NEXT (earlier) launch of the layer's chain
            for (int t = a.fill_t0; t < a.fill_t1; ++t)
                for (int g = 0; g < GZ; ++g) rt_fill_sentinel(a.dzb, a.dzb_lo, (int64_t)t * B + b0, bt, (int)zw, g * H + j0, hcv, epi_tid);
        }
        const RtGatherGeom gg{smem_u32(s.ring), smem_u32(s.a_full), smem_u32(s.a_valid), nsp, rt_tile_bytes(a.box_rows), a.g_ckb_log2, a.g_rp_log2, a.nst, a.expt, a.err};
        uint32_t gpos = 0;
        // zero the CTA-private carries (they persist in global memory across the launches of one sequence)
        if (MODEL != CRNN_MODEL_RNN && a.t_end == a.S) {
            for (int idx = epi_tid; idx < bt * hcv; idx += RT_EPI_THREADS) {
                const int r = idx / hcv, jj = idx - r * hcv;
                a.carry0[(int64_t)(b0 + r) * H + j0 + jj] = 0.f;
            }
        }
        int pck = 0;                                  // persistent form: chunks done (they are processed last to first)
        bool pdead = false;
        for (int t = a.t_end - 1; t >= a.t_begin; --t) {
            if (a.pchunks && t == rt_chunk_t0(a.S, a.pchunks - pck, a.pchunks) - 1) {   // first processed step of a chunk: its dy must be in place
                if (a.in_flag && epi_tid == 0 && !pdead) pdead = !rt_wait_chunk(a.in_flag, (unsigned)pck + 1u, a.err);
                epi_bar_sync();
            }
            const bool has_next = t + 1 < a.S;
            const int64_t tB = (int64_t)t * B;
            // Same split as the forward epilogue: only the bf16 dz rows other CTAs stream next are stored before the
            // step barrier's release; fp32 dz (read by the weight-gradient GEMMs' bias sums after the kernel) and the
            // carries (read back by the same thread) of the phase's last batch go out after it.
            EpiIdx e[UB / VEC];
            float v0[UB], v1[UB], v2[UB], v3[UB], v4[UB];
            bool pending = false;
            const int hq = hcv / VEC;
            for (int ps = 0; ps < a.npass; ++ps) {
                const int nr = min(PR, bt - ps * PR);
                const int nquads = nr * hq;
                const int b0row = b0 + ps * PR;
                if (a.gather && has_next)   // dz_{t+1}, gate columns [0, G1*H), as the other slices finish them
                    gpos = rt_validate(gg, gpos, a.dzb, a.dzb_lo, (int64_t)(t + 1) * B + b0row, nr, (int)zw, 0, G1 * H, nkb0, epi_tid, lane, nullptr);
                const int64_t ob = (int64_t)b0row * H + j0;
                const float* gp = a.gates + (tB + b0row) * gw + j0;
                float* dzp = a.dz + (tB + b0row) * zw + j0;
                __nv_bfloat16* dzbp = a.dzb + (tB + b0row) * zw + j0;
                for (int base = 0; base < nquads; base += RT_EPI_THREADS * (UB / VEC)) {
                    const bool defer = a.defer && (a.sub || ps == a.npass - 1) && base + RT_EPI_THREADS * (UB / VEC) >= nquads;
                    float dyv[UB], c0v[UB], c1v[UB], cr[UB];
#pragma unroll
                    for (int u = 0; u < UB; ++u) dyv[u] = v0[u] = v1[u] = v2[u] = v3[u] = v4[u] = c0v[u] = c1v[u] = cr[u] = 0.f;
#pragma unroll
                    for (int qd = 0; qd < UB / VEC; ++qd) {
                        e[qd] = epi_index<VEC>(base + qd * RT_EPI_THREADS + epi_tid, nquads, hq, ldz, 1, kq * HC);
                        if (e[qd].ok) {
                            const float* g = gp + e[qd].r * (int)gw + e[qd].jj;
                            const int64_t o = ob + e[qd].r * H + e[qd].jj;
                            ldv<VEC>(a.dy + (int64_t)t * BH + o, dyv + VEC * qd);
                            if (MODEL == CRNN_MODEL_LSTM) {
                                ldv<VEC>(g, v0 + VEC * qd); ldv<VEC>(g + H, v1 + VEC * qd); ldv<VEC>(g + 2 * H, v2 + VEC * qd); ldv<VEC>(g + 3 * H, v3 + VEC * qd);
                                ldv<VEC>(a.cx + (int64_t)t * BH + o, c0v + VEC * qd);
                                ldv<VEC>(a.cx + (int64_t)(t + 1) * BH + o, c1v + VEC * qd);
                                ldv<VEC>(a.carry0 + o, cr + VEC * qd);
                            } else if (MODEL == CRNN_MODEL_RNN) {
                                ldv<VEC>(a.hx + (int64_t)(t + 1) * BH + o, v0 + VEC * qd);
                            } else {
                                ldv<VEC>(g + H, v1 + VEC * qd); ldv<VEC>(g + 2 * H, v2 + VEC * qd);
                                ldv<VEC>(a.hx + (int64_t)t * BH + o, c0v + VEC * qd);
                                ldv<VEC>(a.carry0 + o, cr + VEC * qd);
                            }
                        }
                    }
                    if (base == 0 && has_next) RT_TAKE_ACC(nr, HCG);
                    float zv[UB][4];
                    if (has_next) rt_zgather<1, UB, VEC>(zv, e, 1, Zs, zpeer, Q, kq);
                    // results: lstm v0..v3 = di,dj,df,dO, v4 = carry ; rnn v0 = d ; gru v1 = du, v2 = dc, v4 = dh
                    // (branch-free math: the UB chains interleave)
#pragma unroll
                    for (int u = 0; u < UB; ++u) {
                        float dht = dyv[u] + (has_next ? zv[u][0] : 0.f);
                        if (MODEL == CRNN_MODEL_LSTM) {
                            const float si = v0[u], tj = v1[u], sf = v2[u], so = v3[u];
                            const float tcn = fast_tanh(c1v[u]);
                            const float dct = cr[u] + dht * so * (1.0f - tcn * tcn);
                            v0[u] = dct * tj * si * (1.0f - si);
                            v1[u] = dct * si * (1.0f - tj * tj);
                            v2[u] = dct * c0v[u] * sf * (1.0f - sf);
                            v3[u] = dht * tcn * so * (1.0f - so);
                            v4[u] = dct * sf;
                        } else if (MODEL == CRNN_MODEL_RNN) {
                            v0[u] = dht * (1.0f - v0[u] * v0[u]);
                        } else {
                            dht += cr[u];
                            const float u_ = v1[u], cand = v2[u], hp = c0v[u];
                            v2[u] = dht * (1.0f - u_) * (1.0f - cand * cand);     // d candidate pre-activation
                            v1[u] = dht * (hp - cand) * u_ * (1.0f - u_);         // d update-gate pre-activation
                            v4[u] = dht;
                        }
                    }
                    // (cluster exchange: dz of the launch's first step t_begin is read from global memory by the next launch)
                    const bool cx_next = a.xchg && t > a.t_begin;
                    const uint32_t cx_r = (uint32_t)((t - 1) & 1) * cx_buf0;      // dz_t is the operand of step t-1
#pragma unroll
                    for (int qd = 0; qd < UB / VEC; ++qd) {
                        if (!e[qd].ok) continue;
                        __nv_bfloat16* dzb = dzbp + e[qd].r * (int)zw + e[qd].jj;
                        const int kc = j0 + e[qd].jj;
                        if (a.xchg && VEC == 4) {
                            if (MODEL == CRNN_MODEL_LSTM) {
                                if (cx_next) {
                                    cx_push(pbase, pbar, a.csize, (t - 1) & 1, cx_r, cx_lo0, cx_tile, e[qd].r, kc, v0 + VEC * qd);
                                    cx_push(pbase, pbar, a.csize, (t - 1) & 1, cx_r, cx_lo0, cx_tile, e[qd].r, H + kc, v1 + VEC * qd);
                                    cx_push(pbase, pbar, a.csize, (t - 1) & 1, cx_r, cx_lo0, cx_tile, e[qd].r, 2 * H + kc, v2 + VEC * qd);
                                    cx_push(pbase, pbar, a.csize, (t - 1) & 1, cx_r, cx_lo0, cx_tile, e[qd].r, 3 * H + kc, v3 + VEC * qd);
                                }
                            } else if (MODEL == CRNN_MODEL_RNN) {
                                if (cx_next) cx_push(pbase, pbar, a.csize, (t - 1) & 1, cx_r, cx_lo0, cx_tile, e[qd].r, kc, v0 + VEC * qd);
                            } else {
                                cx_push(pbase, pbar, a.csize, 2, 2u * cx_buf0, cx_lo1, cx_tile, e[qd].r, kc, v2 + VEC * qd);     // candidate: this step's second contraction
                                if (cx_next) cx_push(pbase, pbar, a.csize, (t - 1) & 1, cx_r, cx_lo0, cx_tile, e[qd].r, H + kc, v1 + VEC * qd);
                            }
                        } else if (MODEL == CRNN_MODEL_LSTM) {
                            stv_bf16_split<VEC>(dzb, a.dzb_lo, v0 + VEC * qd); stv_bf16_split<VEC>(dzb + H, a.dzb_lo, v1 + VEC * qd);
                            stv_bf16_split<VEC>(dzb + 2 * H, a.dzb_lo, v2 + VEC * qd); stv_bf16_split<VEC>(dzb + 3 * H, a.dzb_lo, v3 + VEC * qd);
                        } else if (MODEL == CRNN_MODEL_RNN) {
                            stv_bf16_split<VEC>(dzb, a.dzb_lo, v0 + VEC * qd);
                        } else {
                            stv_bf16_split<VEC>(dzb + 2 * H, a.dzb_lo, v2 + VEC * qd); stv_bf16_split<VEC>(dzb + H, a.dzb_lo, v1 + VEC * qd);
                        }
                    }
                    pending = defer;
                    if (!defer) {
#pragma unroll
                        for (int qd = 0; qd < UB / VEC; ++qd) {
                            if (!e[qd].ok) continue;
                            float* dz = dzp + e[qd].r * (int)zw + e[qd].jj;
                            const int64_t o = ob + e[qd].r * H + e[qd].jj;
                            if (a.xchg) {              // the global bf16 copies: GEMM operands and the next launch
                                __nv_bfloat16* dzb = dzbp + e[qd].r * (int)zw + e[qd].jj;
                                if (MODEL == CRNN_MODEL_LSTM) {
                                    stv_bf16_split<VEC>(dzb, a.dzb_lo, v0 + VEC * qd); stv_bf16_split<VEC>(dzb + H, a.dzb_lo, v1 + VEC * qd);
                                    stv_bf16_split<VEC>(dzb + 2 * H, a.dzb_lo, v2 + VEC * qd); stv_bf16_split<VEC>(dzb + 3 * H, a.dzb_lo, v3 + VEC * qd);
                                } else if (MODEL == CRNN_MODEL_RNN) {
                                    stv_bf16_split<VEC>(dzb, a.dzb_lo, v0 + VEC * qd);
                                } else {
                                    stv_bf16_split<VEC>(dzb + 2 * H, a.dzb_lo, v2 + VEC * qd); stv_bf16_split<VEC>(dzb + H, a.dzb_lo, v1 + VEC * qd);
                                }
                            }
                            if (MODEL == CRNN_MODEL_LSTM) {
                                stv<VEC>(a.carry0 + o, v4 + VEC * qd);
                                stv<VEC>(dz, v0 + VEC * qd); stv<VEC>(dz + H, v1 + VEC * qd); stv<VEC>(dz + 2 * H, v2 + VEC * qd); stv<VEC>(dz + 3 * H, v3 + VEC * qd);
                            } else if (MODEL == CRNN_MODEL_RNN) {
                                stv<VEC>(dz, v0 + VEC * qd);
                            } else {
                                stv<VEC>(dz + 2 * H, v2 + VEC * qd); stv<VEC>(dz + H, v1 + VEC * qd);
                                stv<VEC>(a.carry1 + o, v4 + VEC * qd);
                            }
                        }
                    }
                }
                epi_bar_sync();
                if (has_next) RT_RELEASE_PARTIALS();
                if (!a.sub && ps + 1 < a.npass) continue;
                // ---- the pass (sub-chains) or the whole phase is done on this CTA: release it, then the deferred stores
            if (a.xchg) { /* the pushed rows count themselves on the receivers' barriers */ }
            else if (a.gather) { /* the stored rows are their own signal */ }
            else if (a.cbar) rt_publish_cluster(s.step_bar, epi_tid, a.csize);
            else rt_publish(a.sub ? a.barrier + tile * a.npass + ps : a.barrier + tile, epi_tid, a.sub);
            if (pending) {
                const int b0row = b0 + ps * PR;
                const int64_t ob = (int64_t)b0row * H + j0;
                float* dzp = a.dz + (tB + b0row) * zw + j0;
#pragma unroll
                for (int qd = 0; qd < UB / VEC; ++qd) {
                    if (!e[qd].ok) continue;
                    float* dz = dzp + e[qd].r * (int)zw + e[qd].jj;
                    const int64_t o = ob + e[qd].r * H + e[qd].jj;
                    if (a.xchg) {
                        __nv_bfloat16* dzb = a.dzb + (tB + b0row) * zw + j0 + e[qd].r * (int)zw + e[qd].jj;
                        if (MODEL == CRNN_MODEL_LSTM) {
                            stv_bf16_split<VEC>(dzb, a.dzb_lo, v0 + VEC * qd); stv_bf16_split<VEC>(dzb + H, a.dzb_lo, v1 + VEC * qd);
                            stv_bf16_split<VEC>(dzb + 2 * H, a.dzb_lo, v2 + VEC * qd); stv_bf16_split<VEC>(dzb + 3 * H, a.dzb_lo, v3 + VEC * qd);
                        } else if (MODEL == CRNN_MODEL_RNN) {
                            stv_bf16_split<VEC>(dzb, a.dzb_lo, v0 + VEC * qd);
                        } else {
                            stv_bf16_split<VEC>(dzb + 2 * H, a.dzb_lo, v2 + VEC * qd); stv_bf16_split<VEC>(dzb + H, a.dzb_lo, v1 + VEC * qd);
                        }
                    }
                    if (MODEL == CRNN_MODEL_LSTM) {
                        stv<VEC>(a.carry0 + o, v4 + VEC * qd);
                        stv<VEC>(dz, v0 + VEC * qd); stv<VEC>(dz + H, v1 + VEC * qd); stv<VEC>(dz + 2 * H, v2 + VEC * qd); stv<VEC>(dz + 3 * H, v3 + VEC * qd);
                    } else if (MODEL == CRNN_MODEL_RNN) {
                        stv<VEC>(dz, v0 + VEC * qd);
                    } else {
                        stv<VEC>(dz + 2 * H, v2 + VEC * qd); stv<VEC>(dz + H, v1 + VEC * qd);
                        stv<VEC>(a.carry1 + o, v4 + VEC * qd);
                    }
                }
            }
            pending = false;
            }   // passes
            if (MODEL == CRNN_MODEL_GRU) {
                pending = false;
                for (int ps = 0; ps < a.npass; ++ps) {
                    const int nr = min(PR, bt - ps * PR);
                    const int nquads = nr * hq;
                    const int b0row = b0 + ps * PR;
                    if (a.gather)   // this step's candidate columns [2H, 3H) of dz
                        gpos = rt_validate(gg, gpos, a.dzb + 2 * H, a.dzb_lo, tB + b0row, nr, (int)zw, 0, H, nkb1, epi_tid, lane, nullptr);
                    const int64_t ob = (int64_t)b0row * H + j0;
                    const float* gp = a.gates + (tB + b0row) * gw + j0;
                    for (int base = 0; base < nquads; base += RT_EPI_THREADS * (UB / VEC)) {
                        const bool defer = a.defer && (a.sub || ps == a.npass - 1) && base + RT_EPI_THREADS * (UB / VEC) >= nquads;
                        float hp[UB], c1v[UB];
#pragma unroll
                        for (int u = 0; u < UB; ++u) v0[u] = v1[u] = v4[u] = hp[u] = c1v[u] = 0.f;
#pragma unroll
                        for (int qd = 0; qd < UB / VEC; ++qd) {
                            e[qd] = epi_index<VEC>(base + qd * RT_EPI_THREADS + epi_tid, nquads, hq, ldz, 1, kq * HC);
                            if (e[qd].ok) {
                                const float* g = gp + e[qd].r * (int)gw + e[qd].jj;
                                const int64_t o = ob + e[qd].r * H + e[qd].jj;
                                ldv<VEC>(g, v0 + VEC * qd); ldv<VEC>(g + H, v1 + VEC * qd);
                                ldv<VEC>(a.hx + (int64_t)t * BH + o, hp + VEC * qd);
                                ldv<VEC>(a.carry1 + o, c1v + VEC * qd);
                            }
                        }
                        if (base == 0) RT_TAKE_ACC(nr, HCG);
                        float zv[UB][4];
                        rt_zgather<1, UB, VEC>(zv, e, 1, Zs, zpeer, Q, kq);
                        // results: v0 = dr, v4 = carry into step t-1
#pragma unroll
                        for (int u = 0; u < UB; ++u) {
                            const float drs = zv[u][0];
                            const float r_ = v0[u], u_ = v1[u];
                            v0[u] = drs * hp[u] * r_ * (1.0f - r_);
                            v4[u] = fmaf(c1v[u], u_, drs * r_);
                        }
#pragma unroll
                        for (int qd = 0; qd < UB / VEC; ++qd) {
                            if (!e[qd].ok) continue;
                            if (a.xchg && VEC == 4) {
                                if (t > a.t_begin) cx_push(pbase, pbar, a.csize, (t - 1) & 1, (uint32_t)((t - 1) & 1) * cx_buf0, cx_lo0, cx_tile, e[qd].r, j0 + e[qd].jj, v0 + VEC * qd);
                            } else stv_bf16_split<VEC>(a.dzb + (tB + b0row) * zw + j0 + e[qd].r * (int)zw + e[qd].jj, a.dzb_lo, v0 + VEC * qd);
                        }
                        pending = defer;
                        if (!defer) {
#pragma unroll
                            for (int qd = 0; qd < UB / VEC; ++qd) {
                                if (!e[qd].ok) continue;
                                if (a.xchg) stv_bf16_split<VEC>(a.dzb + (tB + b0row) * zw + j0 + e[qd].r * (int)zw + e[qd].jj, a.dzb_lo, v0 + VEC * qd);
                                stv<VEC>(a.dz + (tB + b0row) * zw + j0 + e[qd].r * (int)zw + e[qd].jj, v0 + VEC * qd);
                                stv<VEC>(a.carry0 + ob + e[qd].r * H + e[qd].jj, v4 + VEC * qd);
                            }
                        }
                    }
                    epi_bar_sync();
                    RT_RELEASE_PARTIALS();
                    if (!a.sub && ps + 1 < a.npass) continue;
                if (a.xchg) { }
                else if (a.gather) { }
                else if (a.cbar) rt_publish_cluster(s.step_bar, epi_tid, a.csize);
                else rt_publish(a.sub ? a.barrier + tile * a.npass + ps : a.barrier + tile, epi_tid, a.sub);
                if (pending) {
                    const int b0row = b0 + ps * PR;
                    const int64_t ob = (int64_t)b0row * H + j0;
#pragma unroll
                    for (int qd = 0; qd < UB / VEC; ++qd) {
                        if (!e[qd].ok) continue;
                        if (a.xchg) stv_bf16_split<VEC>(a.dzb + (tB + b0row) * zw + j0 + e[qd].r * (int)zw + e[qd].jj, a.dzb_lo, v0 + VEC * qd);
                        stv<VEC>(a.dz + (tB + b0row) * zw + j0 + e[qd].r * (int)zw + e[qd].jj, v0 + VEC * qd);
                        stv<VEC>(a.carry0 + ob + e[qd].r * H + e[qd].jj, v4 + VEC * qd);
                    }
                }
                pending = false;
                }   // passes
            }
            if (a.pchunks && t == rt_chunk_t0(a.S, a.pchunks - 1 - pck, a.pchunks)) {   // the chunk is complete on this CTA
                epi_bar_sync();
                if (epi_tid == 0 && a.out_flag) asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(a.out_flag + pck) : "memory");
                ++pck;
            }
        }
    }
    __syncthreads();
    if (a.csize > 1) cluster_sync_all();             // no CTA leaves while a peer may still signal its barriers
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem, tcols);
    }
}

#undef RT_TAKE_ACC
#undef RT_RELEASE_PARTIALS
#undef fast_tanh
#undef fast_sigmoid

// Pack the recurrent halves of the cell kernels for the forward kernel (bf16, K-major rows).
//   out[(slice*GW*HC + lc) * Kp + k]:  first G1*HC rows lc = jj*G1 + g -> W[(H+k), g*H + j0+jj];
//   GRU: rows 2*HC + jj -> Wc[(H+k), j0+jj].  Rows of units past H and k >= H are zero.
__global__ void __launch_bounds__(256) pack_rec_weights_kernel(const float* __restrict__ W, const float* __restrict__ Wc,
                                                               int H, int HC, int nslices, int G1, int GW, int Kp,
                                                               __nv_bfloat16* __restrict__ out, int64_t lo_ofs = 0) {
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
        store_bf16_split(out + i, lo_ofs, v);
    }
}

}  // namespace tc
}  // namespace crnn
