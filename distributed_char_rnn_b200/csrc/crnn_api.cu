// crnn_api.cu — the C-ABI of include/crnn.h: handle, flat layout, workspace plan and the
// launch sequence of one training step.  Compiled for sm_100a only.
#include "../../include/crnn.h"
#include "common.cuh"
#include "gemm_simt.cuh"
#include "recurrence_simt.cuh"
#include "elementwise.cuh"
#include "gemm_tc.cuh"
#include "recurrence_tc.cuh"
#include "proj_xent_tc.cuh"

#include <stdarg.h>
#include <stdlib.h>
#include <string>
#include <vector>

namespace crnn {

thread_local char g_err[512] = "";
long long g_launches = 0;

inline int set_err(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}
inline int fail_cuda(cudaError_t e, const char* what) {
    return set_err(CRNN_ERR_CUDA, "CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
}

struct TensorInfo {
    std::string name;
    int64_t offset, rows, cols;
};

struct RecPlan {
    int HC, BT, nslices, ntiles, TXa, TXb;
    size_t smem;
};

struct WsPlan {
    size_t scratch;                 // barrier words + double accumulators + norm partials
    size_t x0;
    std::vector<size_t> hx, cx, gates;
    size_t dz, dya, dyb, carry0, carry1, logits;
    // CRNN_PREC_BF16 only: bf16 MMA operands
    std::vector<size_t> xin, xinb;   // dropout: masked copy of each layer's input (fp32 / bf16)
    size_t topd, topdb;              // dropout: masked top-layer output as the projection sees it
    size_t pbf, wsmb, x0b, dlb, gx0;
    std::vector<size_t> wr, hxb, rsb;
    std::vector<size_t> dzv, dzbv, gout, carry0v, carry1v;   // per layer: layers run concurrently (wavefront)
    std::vector<size_t> csum;        // per layer (+ one for softmax_b): row-block partials of the bias-gradient column sums
    size_t es_part, es_sq;           // ordered embedding scatter: per-segment tables [nseg][V][H], partial sums of squares
    int es_nseg, es_tps;             // segments / tokens per segment (0: the table does not fit shared memory -> atomics)
    // sorted form (elementwise.cuh: es_rank_kernel ...): token order by id, per-piece partial rows
    int es_sorted, es_nblk, es_pmax, es_piece_rows;
    size_t es_lrank, es_order, es_hist, es_off, es_poff, es_piece, es_np;
    size_t total;
};

struct RecPlanTc {
    int HC, BT, nslices, ntiles, npass, box_rows, zs_fwd, zs_bwd, swap_fwd, swap_bwd;
    int q_fwd, q_bwd;       // K parts per unit group (cluster size) of the forward / backward kernel
    int ring_fwd, ring_bwd; // TMA ring bytes
    int cx_fwd, cx_bwd;     // cluster exchange (per direction): all slices of a tile in one cluster, rows delivered through distributed shared memory
    int pr, sub;            // rows per pass; 1: the passes of a tile are independent chains (RecTcArgs::sub)
    int gather;             // 1: direct exchange of the streamed rows (recurrence_tc.cuh: rt_gather) instead of TMA loads behind a step barrier
    int g_rp_log2, g_ckb_log2;
    size_t smem_fwd, smem_bwd;
};

}  // namespace crnn

using namespace crnn;

// Launch-time switches, read from the environment ONCE per handle (crnn_create): a step launches ~100 kernels.
struct LaunchEnv {
    int m64, dbg, dbg_layer, one_stage, ckb, rotate, min_stages, use3d, ub8, defer, acq_poll, expt, vec4, cbar, hoist0, fuse_proj;
};

struct crnn_handle {
    crnn_config cfg;
    LaunchEnv env;
    int G;            // gate blocks in the cell kernel(s): lstm 4, gru 3 (2+1), rnn 1
    int GS;           // gate-stash width / H
    int GZ;           // pre-activation-gradient width / H
    int nstate;       // tensors per layer in the RNN state
    std::vector<TensorInfo> tensors;
    int64_t P;
    int64_t emb_off;
    RecPlan fwd, bwd;
    RecPlanTc tcp;
    // wavefront over layers (tensor path): wf_W layer streams, each sequence cut into wf_C chunks
    int wf_W, wf_C;
    int wf_Cb;                           // chunks of the BPTT wavefront (CRNN_WAVEFRONT_CHUNKS_BWD; default: as forward)
    std::vector<cudaStream_t> wf_streams;
    cudaStream_t aux_stream = nullptr;   // small id-only work that runs beside the step (the scatter's token sort)
    int persist = 0;                     // wavefront with ONE recurrence launch per layer and direction (chunk flags + stream memory operations)
    std::vector<cudaEvent_t> wf_events;
    // dropout (training calls only): keep probabilities, seed, and the count of training calls so far
    float ikp, okp;
    bool drop_on;
    bool tc;              // tensor-core path (CRNN_PREC_BF16 or CRNN_PREC_BF16X3)
    int split;            // 1: CRNN_PREC_BF16X3, every bf16 operand array is followed by its lo twin
    long long fb_calls;
    long long* dbg_ptr;   // scratch tail for CRNN_TC_DEBUG stamps (set per call)
    unsigned* err_word;   // host-mapped word a kernel sets when one of its bounded waits expires
    cudaEvent_t bucket_event;   // caller's event, recorded when cells >= bucket_layer have their final gradients (nullptr: off)
    int bucket_layer;
    int dbg_layer;        // the layer whose forward launches record stamps
    int Vp;           // vocab size padded to 8 (bf16 row pitch of the projection operands)
    int fast_act;
    WsPlan ws;
    int sm_count;
    int max_smem;
    // Adam beta powers, tracked like TF's fp32 beta1_power / beta2_power variables
    int64_t pow_step;
    float b1p, b2p;
    // optional per-phase timing (CUDA events on the launching stream)
    bool prof_on;
    std::vector<cudaEvent_t> prof_ev;
    std::vector<int> prof_id;
    int prof_n;
    double prof_ms[CRNN_NUM_PHASES];
    // per-launch timing of the recurrence kernels (events on the stream each launch goes to)
    std::vector<cudaEvent_t> kev;
    std::vector<int> kev_id;
    int kev_n;
    double kern_ms[2];
    long long kern_launches[2];
};

// Bracket one kernel launch with events on ITS stream (which: 0 = rec_fwd, 1 = rec_bwd).
static void prof_kernel_mark(crnn_handle* h, cudaStream_t s, int which) {
    if (!h->prof_on) return;
    if (h->kev_n >= (int)h->kev.size()) {
        cudaEvent_t e;
        if (cudaEventCreate(&e) != cudaSuccess) return;
        h->kev.push_back(e);
        h->kev_id.push_back(-1);
    }
    cudaEventRecord(h->kev[h->kev_n], s);
    h->kev_id[h->kev_n] = which;
    ++h->kev_n;
}

static const char* kPhaseNames[CRNN_NUM_PHASES] = {
    "embed_gather", "xproj_gemm", "rec_fwd", "logits_gemm", "softmax_xent", "proj_grads",
    "rec_bwd", "weight_grads", "input_grads", "embed_scatter", "clip_adam", "misc"};

// Close the running phase and open phase `id` (id < 0: just close).
static void prof_mark(crnn_handle* h, cudaStream_t st, int id) {
    if (!h->prof_on) return;
    if (h->prof_n >= (int)h->prof_ev.size()) {
        cudaEvent_t e;
        if (cudaEventCreate(&e) != cudaSuccess) return;
        h->prof_ev.push_back(e);
        h->prof_id.push_back(-1);
    }
    cudaEventRecord(h->prof_ev[h->prof_n], st);
    h->prof_id[h->prof_n] = id;
    ++h->prof_n;
}

static const float kBeta1 = 0.9f, kBeta2 = 0.999f, kEps = 1e-8f;

// scratch layout (bytes): [0,256) barrier words; [256,512) doubles {loss_sum, sq_sum};
// [1024, 1024 + 8*NORM_BLOCKS) norm partials.
static const size_t kScratchBytes = 32768 + 512 * 256;   // 512 barrier blocks after the first 32 KB
// one 256-byte block of barrier counters per recurrence launch (2 per layer), after the first 8 KB
static inline unsigned* ws_barrier(void* ws, int i) { return reinterpret_cast<unsigned*>((char*)ws + 32768 + (size_t)i * 256); }
// Persistent form of the wavefront (one recurrence launch per layer and direction): chunk flags of (direction, layer) in the
// barrier blocks 448..511 - word 0: "the batch GEMMs of chunks 0..v-1 are done" (written by a stream memory operation, read by
// the kernel), words 32..63: per chunk (in processing order), the number of CTAs that have finished it (read by
// cuStreamWaitValue32; at most 32 chunks).  Zeroed with the blocks.
static inline unsigned* ws_pflag(void* ws, int dir, int l) { return ws_barrier(ws, 448 + dir * 32 + l); }
typedef CUresult (*StreamValueFn)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
static StreamValueFn stream_value_fn(bool write) {
    static StreamValueFn fn[2] = {nullptr, nullptr};
    static bool tried = false;
    if (!tried) {
        tried = true;
        const char* names[2] = {"cuStreamWaitValue32", "cuStreamWriteValue32"};
        for (int i = 0; i < 2; ++i) {
            void* p = nullptr;
            cudaDriverEntryPointQueryResult q;
            if (cudaGetDriverEntryPoint(names[i], &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess) fn[i] = (StreamValueFn)p;
        }
    }
    return fn[write ? 1 : 0];
}
static int wf_wait_value(cudaStream_t s, unsigned* addr, unsigned v) {       // stream s continues once *addr >= v
    const CUresult r = stream_value_fn(false)((CUstream)s, (CUdeviceptr)(uintptr_t)addr, v, CU_STREAM_WAIT_VALUE_GEQ);
    return r == CUDA_SUCCESS ? CRNN_OK : set_err(CRNN_ERR_CUDA, "cuStreamWaitValue32 failed (%d)", (int)r);
}
static int wf_write_value(cudaStream_t s, unsigned* addr, unsigned v) {      // *addr = v once the stream's earlier work is done
    const CUresult r = stream_value_fn(true)((CUstream)s, (CUdeviceptr)(uintptr_t)addr, v, CU_STREAM_WRITE_VALUE_DEFAULT);
    return r == CUDA_SUCCESS ? CRNN_OK : set_err(CRNN_ERR_CUDA, "cuStreamWriteValue32 failed (%d)", (int)r);
}
static inline double* ws_loss(void* ws) { return reinterpret_cast<double*>((char*)ws + 256); }
static inline double* ws_sq(void* ws) { return reinterpret_cast<double*>((char*)ws + 264); }
static inline double* ws_partials(void* ws) { return reinterpret_cast<double*>((char*)ws + 1024); }

static int next_pow2(int v) { int p = 1; while (p < v) p <<= 1; return p; }

static int env_int(const char* name, int dflt) {
    const char* s = getenv(name);
    return (s && *s) ? atoi(s) : dflt;
}

// Pick the (hidden slice, batch tile) decomposition of the persistent recurrence kernels.
static int plan_recurrence(const crnn_handle* h, bool backward, RecPlan* out) {
    const int H = h->cfg.rnn_size, B = h->cfg.batch_size, model = h->cfg.model;
    const int G = h->G;
    const int gfirst = model == CRNN_MODEL_LSTM ? 4 : (model == CRNN_MODEL_GRU ? 2 : 1);
    const int max_ctas = h->sm_count;                    // 1 CTA per SM (shared-memory bound)
    // columns handled by one thread row: fwd NC = gfirst*HC <= 64, bwd NC = HC <= 64
    int hc_tx = backward ? 64 : 64 / gfirst;
    // resident weight slice: G*H*HC floats (fwd and bwd alike) within ~128 KB
    int hc_smem = (int)((128 * 1024 / 4) / ((int64_t)G * ((H + 3) & ~3)));
    int HC = std::min(std::min(hc_tx, hc_smem), H);
    if (HC >= 4) { int p2 = 4; while (p2 * 2 <= HC) p2 *= 2; HC = p2; }   // power of two: no idle thread columns
    if (HC < 1) return set_err(CRNN_ERR_UNSUP, "rnn_size %d too large for the fp32 recurrence kernel", H);
    const double work = (double)B * G * H * (double)H;   // FMAs per step
    int target = (int)std::min<double>(max_ctas, std::max(8.0, work / 65536.0));
    int nslices, ntiles;
    for (;;) {
        nslices = ceil_div(H, HC);
        ntiles = std::min(std::max(1, max_ctas / nslices), ceil_div(B, 8));
        if (nslices * ntiles >= target / 2 || HC <= 4) break;
        HC = std::max(4, HC / 2);
    }
    if (nslices > max_ctas) return set_err(CRNN_ERR_UNSUP, "rnn_size %d needs %d slices > %d SMs", H, nslices, max_ctas);
    ntiles = std::min(ntiles, std::max(1, target / nslices));
    int hc_env = env_int(backward ? "CRNN_REC_BWD_HC" : "CRNN_REC_FWD_HC", 0);
    int nt_env = env_int(backward ? "CRNN_REC_BWD_TILES" : "CRNN_REC_FWD_TILES", 0);
    if (hc_env > 0) { HC = std::min(hc_env, std::min(std::min(hc_tx, hc_smem), H)); nslices = ceil_div(H, HC); ntiles = std::min(ntiles, std::max(1, max_ctas / nslices)); }
    if (nt_env > 0) ntiles = std::min(nt_env, std::max(1, max_ctas / nslices));
    int BT = ceil_div(B, ntiles);
    ntiles = ceil_div(B, BT);
    out->HC = HC; out->BT = BT; out->nslices = nslices; out->ntiles = ntiles;
    const int nc = backward ? HC : gfirst * HC;
    out->TXa = std::max(4, next_pow2(nc));
    out->TXb = std::max(4, next_pow2(HC));
    const size_t K4 = (H + 3) & ~3;
    size_t wfl;
    if (!backward) wfl = K4 * (size_t)G * HC;
    else wfl = (size_t)(((int64_t)gfirst * H + 3) & ~3) * HC + (model == CRNN_MODEL_GRU ? K4 * HC : 0);
    out->smem = (wfl + 2 * REC_RB * REC_LDA + (size_t)REC_RB * (nc + 1)) * sizeof(float);
    if ((int)out->smem > h->max_smem)
        return set_err(CRNN_ERR_UNSUP, "recurrence kernel needs %zu B shared memory > %d", out->smem, h->max_smem);
    return CRNN_OK;
}

// Decomposition of the tensor-core recurrence kernels: HC hidden units (multiple of 16) per slice,
// BT batch rows per tile.  The resident weight slice (2*GW*H*HC bytes) plus the 64 KB A ring must
// fit in shared memory; one CTA per SM, all co-resident.
static int plan_recurrence_tc(const crnn_handle* h, int max_ctas, RecPlanTc* out) {
    const int H = h->cfg.rnn_size, B = h->cfg.batch_size, model = h->cfg.model;
    if (H % 16 != 0) return set_err(CRNN_ERR_UNSUP, "the tensor-core precisions need rnn_size %% 16 == 0 (got %d)", H);
    const int g1 = model == CRNN_MODEL_LSTM ? 4 : (model == CRNN_MODEL_GRU ? 2 : 1);
    const int nkb = ceil_div(H, 64);
    const int hc_env = env_int("CRNN_TC_HC", 0), nt_env = env_int("CRNN_TC_TILES", 0);
    const int dbg = h->env.dbg != 0;                         // the stamp area of tools/dbg_rec_timeline.py
    auto fill = [&](int HC, RecPlanTc* o) -> bool {
        const int nslices = ceil_div(H, HC);
        if (nslices > max_ctas) return false;
        int ntiles = std::max(1, std::min(max_ctas / nslices, ceil_div(B, 16)));
        if (nt_env > 0) ntiles = std::max(1, std::min(nt_env, max_ctas / nslices));
        if (ntiles > 64) ntiles = 64;                        // 64 barrier words per launch
        const int BT = ceil_div(B, ntiles);
        o->HC = HC; o->BT = BT; o->nslices = nslices; o->ntiles = ceil_div(B, BT);
        o->npass = ceil_div(BT, 128);
        o->box_rows = o->npass > 1 ? 128 : std::min(128, (BT + 7) & ~7);
        int rows = std::min(128, BT);
        // MMA orientation: cost ~ N with M fixed at 128, so put the smaller of (gate columns, rows) on N
        const int rows16 = (rows + 15) & ~15;
        const int sw_env = env_int("CRNN_TC_SWAP", -1);
        // (measured on B200, config 3: the batch-rows-on-M form is 3-10 % faster end to end even though its
        //  MMAs are longer - the transposed accumulator read-out costs more than the MMA saves - so the
        //  swapped form is opt-in.)
        o->swap_fwd = o->swap_bwd = 0;
        (void)rows16;
        if (sw_env == 1) { o->swap_fwd = g1 * HC <= 128; o->swap_bwd = HC <= 128; }
        // Split K over Q CTAs of a cluster when the per-CTA MMAs would otherwise be many and narrow
        // (large H: the issue rate of small-N MMAs, not the tensor pipe, bounds the step).  Same resident
        // bytes per CTA; N grows Q-fold, the K extent and the streamed rows shrink Q-fold; the partial sums
        // are exchanged through distributed shared memory, whose staging area grows with N.
        const size_t wbytes = ((size_t)nkb * g1 * HC * 128 + (model == CRNN_MODEL_GRU ? (size_t)nkb * HC * 128 : 0)) * (size_t)(1 + h->split);
        int ring_q[2] = {tc::RT_RING_BYTES, tc::RT_RING_BYTES};          // base ring per direction (a K split may trade ring for staging)
        auto pick_q = [&](int ncols_per_unit, int env, int* ring) -> int {
            for (int Qc = 4; Qc >= 2; Qc >>= 1) {
                if (env > 0 && Qc != env) continue;
                // only long K, keep N <= 128
                if (env == 0 && (nkb < 16 || ncols_per_unit * HC * Qc > 128)) continue;
                // measured (B200): Q = 4 pays (config 4 BPTT 9.9 -> 5.9 ms); Q = 2 pays only when one column per
                // unit is exchanged (config 5: 26.3 -> 23.9 ms per direction) and loses for the LSTM forward
                // pass, which gathers four (config 4: 4.6 -> 6.0 ms)
                if (env == 0 && Qc == 2 && ncols_per_unit > 1) continue;
                if (H % (64 * Qc) != 0 || H % (HC * Qc) != 0 || nslices % Qc != 0) continue;
                if (ncols_per_unit * HC * Qc > (h->split ? 128 : 256)) continue;   // split: the MMA's N is twice the columns
                // (the partial sums' staging area grows with Q; taking ring space for it does not pay: config 5 in plain bf16 with
                //  Q = 4 and a 32 KB ring 71.4 ms per step against 66.0 with Q = 2 and 64 KB - CRNN_TC_RING_MIN_KB for experiments)
                int rb = tc::RT_RING_BYTES;
                const int rb_min = std::max(16, env_int("CRNN_TC_RING_MIN_KB", 64)) * 1024;
                while (rb > rb_min && (int)tc::rt_smem_bytes(wbytes, 0, rows * (ncols_per_unit * HC * Qc + 1), rb, dbg) > h->max_smem) rb -= 16 * 1024;
                if ((int)tc::rt_smem_bytes(wbytes, 0, rows * (ncols_per_unit * HC * Qc + 1), rb, dbg) > h->max_smem) continue;
                if (tc::rt_num_stages(o->box_rows, 1, h->split, rb) < 2) continue;      // the ring must still hold two stages
                *ring = rb;
                return Qc;
            }
            return 1;
        };
        const int q_env = env_int("CRNN_TC_KSPLIT", -1);       // -1 auto, 0/1 off, 2/4 forced
        o->q_fwd = q_env == 0 || q_env == 1 ? 1 : pick_q(g1, std::max(q_env, 0), &ring_q[0]);
        o->q_bwd = q_env == 0 || q_env == 1 ? 1 : pick_q(1, std::max(q_env, 0), &ring_q[1]);
        // Sub-chains (CRNN_TC_SUB=n): a tile's rows are cut into n passes that are chains of their own (own step barrier
        // word, released pass by pass).  The CTA keeps ONE resident weight slice for all of them and works on pass k+1
        // while pass k waits for the other slices: the exchange latency hides behind useful work of the same CTA.
        o->pr = 128; o->sub = 0;
        {
            // Measured on B200, config 3 (ms per step): split precision (tiles of 128 rows -> 2 x 64) 4.65 -> 3.98, four passes
            // 5.27; plain bf16 (tiles of 64 rows -> 2 x 32, half-empty M = 64 MMAs and twice as many of them) 3.02 -> 3.26.
            // At the same slice width (HC = 32) two passes also help plain bf16 a little: 3.06 -> 2.94.  Default: two passes for
            // tiles of at least 64 rows.
            const int nsub = env_int("CRNN_TC_SUB", BT >= 64 ? 2 : 0);
            if (nsub > 1 && o->q_fwd == 1 && o->q_bwd == 1 && !o->swap_fwd && !o->swap_bwd && BT >= 16 * nsub) {
                const int pr = (ceil_div(BT, nsub) + 7) & ~7;
                const int np = ceil_div(BT, pr);
                if (pr <= 128 && np > 1 && o->ntiles * np <= 64) { o->pr = pr; o->sub = 1; o->npass = np; o->box_rows = pr; rows = pr; }
            }
        }
        o->zs_fwd = rows * (g1 * HC * o->q_fwd + 1);
        o->zs_bwd = rows * (HC * o->q_bwd + 1);
        o->ring_fwd = ring_q[0]; o->ring_bwd = ring_q[1];
        o->smem_fwd = tc::rt_smem_bytes(wbytes, 0, o->zs_fwd, o->ring_fwd, dbg);
        o->smem_bwd = tc::rt_smem_bytes(wbytes, 0, o->zs_bwd, o->ring_bwd, dbg);
        if ((int)std::max(o->smem_fwd, o->smem_bwd) > h->max_smem) return false;
        // A larger ring (CRNN_TC_RING_KB) where shared memory is left.  Plain bf16: measured no gain (config 3 at 64 / 96 /
        // 128 KB: 4.20 / 4.22 / 4.22 ms).  Split operands stream twice the bytes per K block, a 64 KB ring holds two
        // stages of them, and a third stage in flight pays once the MMAs no longer pace the stream (config 3: 4.84 ->
        // 4.62 ms at 96 KB), so the split ring takes what is left up to 112 KB.
        const int ring_cap = env_int("CRNN_TC_RING_KB", h->split ? 112 : 64) * 1024;
        auto grow = [&](size_t* smem, int* ring) {
            const int spare = ((h->max_smem - (int)*smem) / 1024) * 1024;
            const int add = std::max(0, std::min(spare, ring_cap - *ring));
            *ring += add; *smem += (size_t)add;
        };
        grow(&o->smem_fwd, &o->ring_fwd);
        grow(&o->smem_bwd, &o->ring_bwd);
        // One TMA instruction per phase when the phase's whole K extent of streamed rows fits: the loads of a phase
        // do not overlap each other in practice (measured: 2 x 32 KB complete 1.3 k and 3.0 k cycles after issue, one
        // 64 KB load after 1.6 k), so the ring is grown to hold the H extent as a single stage where that is possible.
        if (env_int("CRNN_TC_ONE_STAGE", 1) && o->npass == 1) {
            const int tile = tc::rt_tile_bytes(o->box_rows);
            auto one = [&](size_t* smem, int* ring, int q) {
                const int want = ((nkb / q) * tile * (1 + h->split) + tc::RT_MROWS_BYTES - tile + 1023) & ~1023;
                if (want > *ring && want <= 96 * 1024 && (int)*smem + (want - *ring) <= h->max_smem) { *smem += (size_t)(want - *ring); *ring = want; }
            };
            one(&o->smem_fwd, &o->ring_fwd, o->q_fwd);
            one(&o->smem_bwd, &o->ring_bwd, o->q_bwd);
        }
        // Direct exchange (CRNN_TC_GATHER=1, recurrence_tc.cuh: rt_validate / rt_probe): no step barrier; the streamed rows
        // carry their own readiness (sentinel-filled slots, landed stages checked in shared memory).  Single-pass tiles,
        // batch rows on M, no K split; stages of at most 16 KB per operand half and at most 4096 chunks.
        // Measured on B200 (config 3, bf16, ms per step): barrier protocol 3.04; direct exchange 3.83-4.05 - the check of
        // the landed rows by the epilogue threads costs ~4 k cycles per phase next to the other layer's kernel; with the
        // check switched off (wrong results, timing only) 3.02, i.e. removing MEMBAR + flag + acquire poll from the chain
        // buys ~1 k of ~10 k cycles per phase in the forward kernel and nothing over the whole step.  A kernel running
        // alone gains 18 % (6.9 k vs 8.4-9.3 k cycles per phase).  Off by default.
        o->gather = 0; o->g_rp_log2 = o->g_ckb_log2 = 0;
        if (env_int("CRNN_TC_GATHER", 0) && (o->npass == 1 || o->sub) && o->q_fwd == 1 && o->q_bwd == 1 && !o->swap_fwd && !o->swap_bwd) {
            const int tile = tc::rt_tile_bytes(o->box_rows), nsp = 1 + h->split;
            int rp = 5; while ((1 << rp) < rows) ++rp;
            int ck = 0;
            while ((2 << ck) <= nkb && nsp * (1 << rp) * (2 << ck) * 8 <= 4096 && (2 << ck) * tile <= 16384) ++ck;
            { const int ck_env = env_int("CRNN_TC_CKB", 0); if (ck_env > 0) { ck = 0; while ((2 << ck) <= ck_env && (2 << ck) <= nkb && nsp * (1 << rp) * (2 << ck) * 8 <= 4096) ++ck; } }
            if (nsp * (1 << rp) * (1 << ck) * 8 <= 4096) {
                o->gather = 1; o->g_rp_log2 = rp; o->g_ckb_log2 = ck;
                const int cap = env_int("CRNN_TC_RING_KB", 128) * 1024;
                auto grow2 = [&](size_t* smem, int* ring) {
                    const int spare = ((h->max_smem - (int)*smem) / 1024) * 1024;
                    const int add = std::max(0, std::min(spare, cap - *ring));
                    *ring += add; *smem += (size_t)add;
                };
                grow2(&o->smem_fwd, &o->ring_fwd);
                grow2(&o->smem_bwd, &o->ring_bwd);
            }
        }
        // Cluster exchange (recurrence_tc.cuh: cx_push / rt_mma_cx): when ALL slices of a tile fit a portable cluster, the rows
        // a slice finishes are written straight into every slice's operand buffers (distributed shared memory) and nothing of
        // the step hand-shake goes through L2.  Needs the phase's whole K extent per buffer: two buffers forward, two (gru:
        // two + one) backward.  (tools/ubench_dsmem.cu: pays for clusters of <= 8 with a few KB per slice; at config 3's
        // geometry - 16 slices x 4 KB - the exchange through L2 is faster, and its weights leave no room for the buffers.)
        o->cx_fwd = o->cx_bwd = 0;
        if (env_int("CRNN_TC_CX", 1) && (nslices == 2 || nslices == 4 || nslices == 8) && nslices <= std::min(8, h->env.cbar) &&
            o->npass == 1 && !o->sub && !o->gather && o->q_fwd == 1 && o->q_bwd == 1 && !o->swap_fwd && !o->swap_bwd) {
            const int tile = tc::rt_tile_bytes(o->box_rows), nsp = 1 + h->split;
            const int nkbg = ceil_div(g1 * H, 64);
            const int need_f = (2 * nsp * nkb * tile + tc::RT_MROWS_BYTES - tile + 1023) & ~1023;
            const int need_b = ((2 * nkbg + (model == CRNN_MODEL_GRU ? nkb : 0)) * nsp * tile + tc::RT_MROWS_BYTES - tile + 1023) & ~1023;
            // (each direction on its own: BPTT's buffers hold the G1*H gate columns and stop fitting first - LSTM 256 in
            //  split precision exchanges inside the cluster forward and through L2 backward)
            const int add_f = std::max(0, need_f - o->ring_fwd), add_b = std::max(0, need_b - o->ring_bwd);
            if ((int)o->smem_fwd + add_f <= h->max_smem) { o->cx_fwd = 1; o->ring_fwd += add_f; o->smem_fwd += (size_t)add_f; }
            if ((int)o->smem_bwd + add_b <= h->max_smem) { o->cx_bwd = 1; o->ring_bwd += add_b; o->smem_bwd += (size_t)add_b; }
        }
        return true;
    };
    // Largest slice that fits (fewest slices -> least L2 traffic for the h all-gather), shrunk
    // while that leaves SMs idle: every SM should hold a (slice, tile) pair.
    const int hc_max = std::min((h->split ? 128 : 256) / g1, H) & ~15;   // MMA N <= 256 (split operands: N = 2 * columns)
    int best = 0;
    RecPlanTc cand;
    // (slice widths that divide H first: a ragged last slice idles most of one CTA per tile - HC = 48 at H = 512 measured
    //  3.29 ms per step against 2.94 at HC = 32)
    for (int even = 1; even >= 0 && best == 0; --even)
        for (int HC = hc_max; HC >= 16; HC -= 16) {
            if (hc_env >= 16 && HC > hc_env) continue;
            if (even && hc_env < 16 && H % HC != 0) continue;
            if (!fill(HC, &cand)) continue;
            if (best == 0 || cand.nslices * cand.ntiles > out->nslices * out->ntiles) { best = HC; *out = cand; }
            if (hc_env >= 16 || cand.nslices * cand.ntiles >= (max_ctas * 3) / 4) break;
        }
    // A slice width at which the tile's slices fit one cluster (cluster exchange, at least forward) beats more, narrower
    // slices that must exchange through L2 (measured, split precision, 2 layers, B = 64, S = 50: LSTM 256 1.03 -> 0.97 ms,
    // GRU 256 1.47 -> 0.95, tanh-RNN 256 0.76 -> 0.53); among such widths the one with the most CTAs.
    if (best != 0 && hc_env < 16 && !out->cx_fwd) {
        RecPlanTc bx; int best_x = 0;
        for (int HC = hc_max; HC >= 16; HC -= 16) {
            if (H % HC != 0 || !fill(HC, &cand) || !cand.cx_fwd) continue;
            if (best_x == 0 || cand.nslices * cand.ntiles > bx.nslices * bx.ntiles) { best_x = HC; bx = cand; }
        }
        if (best_x != 0) { best = best_x; *out = bx; }
    }
    // LSTM slices of 8 units (32 gate columns, N = 32 / 64 MMAs) where 16 do not fit: hi + lo slices of H = 1024 (config 4 in
    // split precision) are 128 KB per CTA at 128 slices.  BPTT contracts one column per unit, so its MMA needs the K split
    // (N = 8 Q >= 16) or M = 64 tiles (N = 8 is legal there).
    if ((best == 0 || hc_env == 8) && model == CRNN_MODEL_LSTM && H % 8 == 0 && fill(8, &cand) &&
        (cand.q_bwd >= 2 || std::min(cand.pr, cand.BT) <= 64) && !cand.swap_fwd && !cand.swap_bwd) {
        best = 8; *out = cand;
    }
    if (best == 0) return set_err(CRNN_ERR_UNSUP, "rnn_size %d: recurrent weight slice does not fit in shared memory%s", H,
                                  h->split ? " with split operands (use CRNN_PREC_BF16 or CRNN_PREC_FP32)" : "");
    return CRNN_OK;
}

static void build_layout(crnn_handle* h) {
    const int H = h->cfg.rnn_size, V = h->cfg.vocab_size, L = h->cfg.num_layers;
    int64_t off = 0;
    auto add = [&](const std::string& name, int64_t r, int64_t c) {
        TensorInfo t{name, off, r, c};
        h->tensors.push_back(t);
        off = round_up64(off + r * c, 64);
    };
    add("rnnlm/softmax_w", H, V);
    add("rnnlm/softmax_b", 1, V);
    h->emb_off = off;
    add("embedding", V, H);
    for (int l = 0; l < L; ++l) {
        std::string p = "rnnlm/multi_rnn_cell/cell_" + std::to_string(l) + "/";
        if (h->cfg.model == CRNN_MODEL_LSTM) {
            add(p + "lstm_cell/kernel", 2 * H, 4 * H);
            add(p + "lstm_cell/bias", 1, 4 * H);
        } else if (h->cfg.model == CRNN_MODEL_GRU) {
            add(p + "gru_cell/gates/kernel", 2 * H, 2 * H);
            add(p + "gru_cell/gates/bias", 1, 2 * H);
            add(p + "gru_cell/candidate/kernel", 2 * H, H);
            add(p + "gru_cell/candidate/bias", 1, H);
        } else {
            add(p + "basic_rnn_cell/kernel", 2 * H, H);
            add(p + "basic_rnn_cell/bias", 1, H);
        }
    }
    h->P = off;
}

static void build_workspace(crnn_handle* h) {
    const int64_t H = h->cfg.rnn_size, V = h->cfg.vocab_size, L = h->cfg.num_layers;
    const int64_t S = h->cfg.seq_length, B = h->cfg.batch_size;
    size_t off = 0;
    auto take = [&](int64_t floats) { size_t o = off; off = (size_t)round_up64(off + floats * 4, 256); return o; };
    WsPlan& w = h->ws;
    w.scratch = off; off += kScratchBytes;
    w.x0 = take(S * B * H);
    for (int l = 0; l < L; ++l) {
        w.hx.push_back(take((S + 1) * B * H));
        w.cx.push_back(h->cfg.model == CRNN_MODEL_LSTM ? take((S + 1) * B * H) : 0);
        w.gates.push_back(take(S * B * h->GS * H));
    }
    const bool fp32_path = h->cfg.precision == CRNN_PREC_FP32;
    w.dz = fp32_path ? take((S + 1) * B * h->GZ * H) : 0;
    w.dya = fp32_path ? take(S * B * H) : 0;
    w.dyb = fp32_path ? take(S * B * H) : 0;
    w.carry0 = fp32_path ? take(B * H) : 0;
    w.carry1 = fp32_path ? take(B * H) : 0;
    w.logits = take(S * B * V);
    const int64_t nsp = 1 + h->split;                 // bf16 operand arrays: hi (+ lo twin right behind it)
    auto take2 = [&](int64_t halves) { return take((halves * nsp + 1) / 2); };
    w.topd = w.topdb = 0;
    if (h->drop_on) {
        const bool bf = h->tc;
        for (int l = 0; l < L; ++l) {
            w.xin.push_back(take(S * B * H));
            w.xinb.push_back(bf ? take2(S * B * H) : 0);
        }
        w.topd = take(S * B * H);
        w.topdb = bf ? take2(S * B * H) : 0;
    }
    if (h->tc) {
        const int64_t GW = h->G;
        w.pbf = take2(h->P);
        w.wsmb = take2(H * h->Vp);
        w.x0b = take2(S * B * H);
        for (int l = 0; l < L; ++l) {
            w.wr.push_back(take2((int64_t)h->tcp.nslices * GW * h->tcp.HC * H));
            w.hxb.push_back(take2((S + 1) * B * H));
            w.rsb.push_back(h->cfg.model == CRNN_MODEL_GRU ? take2(S * B * H) : 0);
        }
        for (int l = 0; l < L; ++l) {
            w.dzv.push_back(take(S * B * h->GZ * H));
            w.dzbv.push_back(take2((S + 1) * B * h->GZ * H));
            w.gout.push_back(take(S * B * H));
            w.carry0v.push_back(take(B * H));
            w.carry1v.push_back(take(B * H));
        }
        w.gx0 = take(S * B * H);
        w.dlb = take2(S * B * h->Vp);
    }
    // bias-gradient column sums, combined in row-block order (one region per layer: layers overlap; the last for softmax_b)
    // (the last region also holds the fused projection kernel's per-CTA column sums: one row per 128 rows)
    for (int l = 0; l <= L; ++l) w.csum.push_back(take(l < L ? colsum_scratch_floats(S * B, h->GZ * H) : ceil_div64(S * B, 128) * V));
    // ordered embedding scatter: token segments of >= 128 rows, at most 32 MB of per-segment tables
    w.es_nseg = 0; w.es_tps = 0; w.es_part = w.es_sq = 0;
    const bool es_sort_ok = env_int("CRNN_ES_SORT", 1) && H % 4 == 0 && V <= 8192 && S * B < (1LL << 31);
    if (!es_sort_ok && (int64_t)es_smem_bytes((int)V) + 1024 <= h->max_smem && H % 4 == 0) {
        // segments: about two CTAs per SM over all column chunks, at least 128 rows each, at most 32 MB of tables
        const int64_t N = S * B, tab = V * H * 4, nchunk = ceil_div64(H, ES_COLS);
        int64_t nseg = std::max<int64_t>(1, std::min<int64_t>(std::min<int64_t>(N / 128, (32LL << 20) / tab), std::max<int64_t>(1, 2 * h->sm_count / nchunk)));
        w.es_tps = (int)ceil_div64(N, nseg);
        w.es_nseg = (int)ceil_div64(N, w.es_tps);
        w.es_part = take((int64_t)w.es_nseg * V * H);
        w.es_sq = take(2 * (int64_t)w.es_nseg * ceil_div64(H, ES_COLS));
    }
    // sorted form: the default when the vocabulary's counters fit one block's shared memory (CRNN_ES_SORT=0: table form)
    w.es_sorted = 0;
    if (es_sort_ok) {
        const int64_t N = S * B;
        w.es_sorted = 1;
        w.es_nblk = (int)ceil_div64(N, ES_SORT_BLOCK);
        // pieces of 64 tokens, or of 32 where 64 would leave the GPU short of CTAs (measured, config 3: 0.048 -> 0.041 ms with
        // 32; config 5: 0.23 -> 0.34 ms, its per-piece partial rows double)
        w.es_piece_rows = (N / 64 + V) * ceil_div64(H, 4 * ES_SUM_THREADS) < 4LL * h->sm_count ? 32 : 64;
        w.es_pmax = (int)(N / w.es_piece_rows + V + 1);
        w.es_lrank = take(N); w.es_order = take(N);
        w.es_hist = take((int64_t)w.es_nblk * V);
        w.es_off = take(V + 1); w.es_poff = take(V + 1);
        w.es_piece = take(2 * (int64_t)w.es_pmax); w.es_np = take(4);
        w.es_part = take((int64_t)w.es_pmax * H);
        w.es_sq = take(2 * (int64_t)w.es_pmax * ceil_div64(H, 4 * ES_SUM_THREADS));
    }
    w.total = off;
}

extern "C" {

const char* crnn_last_error(void) { return g_err; }
const char* crnn_version(void) { return "crnn-b200 0.1.0 (sm_100a)"; }

int crnn_create(const crnn_config* cfg, crnn_handle** out) {
    if (!cfg || !out) return set_err(CRNN_ERR_ARG, "null argument");
    *out = nullptr;
    if (cfg->model != CRNN_MODEL_RNN && cfg->model != CRNN_MODEL_GRU && cfg->model != CRNN_MODEL_LSTM)
        return set_err(CRNN_ERR_ARG, "model type not supported: %d", cfg->model);
    if (cfg->rnn_size < 1 || cfg->num_layers < 1 || cfg->seq_length < 1 || cfg->batch_size < 1 || cfg->vocab_size < 1)
        return set_err(CRNN_ERR_ARG, "rnn_size, num_layers, seq_length, batch_size and vocab_size must be >= 1");
    if (cfg->input_keep_prob < 0.f || cfg->input_keep_prob > 1.f || cfg->output_keep_prob < 0.f || cfg->output_keep_prob > 1.f)
        return set_err(CRNN_ERR_ARG, "keep probabilities must be in (0, 1] (0 = unset = 1.0)");
    if (cfg->precision != CRNN_PREC_FP32 && cfg->precision != CRNN_PREC_BF16 && cfg->precision != CRNN_PREC_BF16X3)
        return set_err(CRNN_ERR_ARG, "unknown precision %d", cfg->precision);
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return set_err(CRNN_ERR_CUDA, "no CUDA device: this library has no CPU path (%s)", cudaGetErrorString(e));
    }
    if (cfg->device < 0 || cfg->device >= ndev) return set_err(CRNN_ERR_ARG, "device %d out of range [0,%d)", cfg->device, ndev);
    cudaDeviceProp prop;
    CRNN_CUDA(cudaGetDeviceProperties(&prop, cfg->device));
    if (prop.major != 10)
        return set_err(CRNN_ERR_UNSUP, "device %d is sm_%d%d; this library is built for sm_100a only", cfg->device, prop.major, prop.minor);
    CRNN_CUDA(cudaSetDevice(cfg->device));
    crnn_handle* h = new crnn_handle();
    h->cfg = *cfg;
    h->G = cfg->model == CRNN_MODEL_LSTM ? 4 : (cfg->model == CRNN_MODEL_GRU ? 3 : 1);
    h->GS = cfg->model == CRNN_MODEL_RNN ? 1 : 4;     // lstm: i,j,f,o   gru: r,u,cand,r*h
    h->GZ = h->G;
    h->nstate = cfg->model == CRNN_MODEL_LSTM ? 2 : 1;
    h->tc = cfg->precision != CRNN_PREC_FP32;
    h->split = cfg->precision == CRNN_PREC_BF16X3 ? 1 : 0;
    h->sm_count = prop.multiProcessorCount;
    h->max_smem = (int)prop.sharedMemPerBlockOptin;
    h->pow_step = 0; h->b1p = 1.0f; h->b2p = 1.0f;
    h->prof_on = false; h->prof_n = 0; h->kev_n = 0;
    h->kern_ms[0] = h->kern_ms[1] = 0.0; h->kern_launches[0] = h->kern_launches[1] = 0;
    for (int i = 0; i < CRNN_NUM_PHASES; ++i) h->prof_ms[i] = 0.0;
    build_layout(h);
    int rc = plan_recurrence(h, false, &h->fwd);
    if (rc == CRNN_OK) rc = plan_recurrence(h, true, &h->bwd);
    h->Vp = (cfg->vocab_size + 7) & ~7;
    h->fast_act = h->split ? 0 : env_int("CRNN_FAST_ACT", 1);   // MUFU.TANH (5e-4) has no place in the fp32-grade mode
    h->ikp = (cfg->input_keep_prob > 0.f && cfg->input_keep_prob < 1.f) ? cfg->input_keep_prob : 1.0f;
    h->okp = (cfg->output_keep_prob > 0.f && cfg->output_keep_prob < 1.f) ? cfg->output_keep_prob : 1.0f;
    h->drop_on = h->ikp < 1.0f || h->okp < 1.0f;
    h->fb_calls = 0;
    h->err_word = nullptr;
    h->bucket_event = nullptr; h->bucket_layer = 0;
    h->env = LaunchEnv{env_int("CRNN_TC_M64", 1), env_int("CRNN_TC_DEBUG", 0), env_int("CRNN_TC_DEBUG_LAYER", 0), env_int("CRNN_TC_ONE_STAGE", 1),
                       env_int("CRNN_TC_CKB", 0), env_int("CRNN_TC_ROTATE", 0), env_int("CRNN_TC_MIN_STAGES", 2), env_int("CRNN_TC_3D", 1),
                       env_int("CRNN_TC_UB8", 1), env_int("CRNN_TC_DEFER", 1), env_int("CRNN_TC_ACQ_POLL", 1), env_int("CRNN_TC_EXPT", 0),
                       env_int("CRNN_TC_VEC4", -1), env_int("CRNN_TC_CBAR", 8), env_int("CRNN_WF_HOIST0", 0), env_int("CRNN_FUSE_PROJ", 1)};
    h->wf_W = 1; h->wf_C = 1; h->wf_Cb = 1;
    if (rc == CRNN_OK && h->tc) {
        // Layers are independent chains once their inputs exist: run up to W of them at once, each on
        // 1/W of the SMs, the sequence cut into C chunks so layer l+1 starts when layer l has produced
        // the first chunk.  W = 1 is the plain sequential schedule.
        // Measured on B200: overlapping two layers pays while a layer's recurrence is latency-bound
        // (config 1: 1.09 -> 0.85 ms/step, config 3: 6.03 -> 5.18 ms) and costs when it is bound by the
        // L2 all-gather of h (config 4: 19.3 -> 21.5 ms), so it is on for recurrent weights <= 4 MB/layer.
        const int64_t wbytes = 2LL * h->G * cfg->rnn_size * cfg->rnn_size;
        int W = std::min(cfg->num_layers, env_int("CRNN_WAVEFRONT", wbytes <= (4LL << 20) ? 2 : 1));
        if (W < 1) W = 1;
        for (; W >= 1; --W) {
            // W > 1: leave a few SMs to the batch GEMMs that run beside the persistent recurrence kernels
            const int reserve = W > 1 ? env_int("CRNN_WF_RESERVE", 20) : 0;     // measured best on config 3 (0: 4.96, 20: 4.81, 36: 5.47 ms)
            int ctas = (h->sm_count - reserve) / W;
            { const int cap = env_int("CRNN_TC_MAXCTAS", 0); if (cap > 0) ctas = std::min(ctas, cap); }   // experiments only
            rc = plan_recurrence_tc(h, ctas, &h->tcp);
            if (rc == CRNN_OK) break;
        }
        if (rc == CRNN_OK) {
            h->wf_W = W;
            int C = W > 1 ? std::max(1, std::min(16, cfg->seq_length / 10)) : 1;
            const int c_env = env_int("CRNN_WAVEFRONT_CHUNKS", 0);
            if (c_env > 0) C = std::min(std::min(c_env, cfg->seq_length), 32);
            if (W == 1 && c_env <= 0) C = 1;                 // (chunks without overlap only on request: measures the cost of a launch)
            h->wf_C = C;
            while (C > 1 && 2 * cfg->num_layers * C > 384) --C, h->wf_C = C;   // barrier blocks 0..383
            h->wf_Cb = h->wf_C;
            { const int cb = env_int("CRNN_WAVEFRONT_CHUNKS_BWD", 0);
              if (cb > 0 && W > 1 && cfg->num_layers * (h->wf_C + std::min(cb, cfg->seq_length)) <= 384) h->wf_Cb = std::min(std::min(cb, cfg->seq_length), 32); }
            for (int i = 0; i < 2 * W && W > 1; ++i) {      // [0,W): recurrence streams, [W,2W): GEMM streams
                cudaStream_t s;
                // The recurrence launches are the step's critical path and need a whole set of free SMs each (cooperative);
                // the chunk GEMMs only have to be done a chunk later.  Highest priority for the former, lowest for the
                // latter: an SM that frees up goes to a waiting recurrence CTA before another GEMM CTA (CRNN_WF_PRIO=0: off).
                int prio_lo = 0, prio_hi = 0;
                cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);
                const int prio = !env_int("CRNN_WF_PRIO", 1) ? 0 : (i < W ? prio_hi : prio_lo);
                if (cudaStreamCreateWithPriority(&s, cudaStreamNonBlocking, prio) != cudaSuccess) { rc = set_err(CRNN_ERR_CUDA, "cudaStreamCreate failed"); break; }
                h->wf_streams.push_back(s);
            }
        }
    }
    // One launch per layer and direction instead of one per chunk (each launch costs ~10 us: set-up, resident weights,
    // the gap to the next kernel): needs the driver's stream memory operations; not with the direct exchange, whose
    // launches pre-fill the next launch's slots.  CRNN_WF_PERSIST=0: a launch per chunk, ordered by events.
    // Measured (ms per step, split precision): config 1 0.60 -> 0.56 with both directions in this form; config 3 4.21 ->
    // 4.51 with both - in the forward pass its two resident kernels leave the chunk GEMMs 20 SMs all the time, where a launch
    // per chunk lets them use the SMs between launches - but 4.21 -> 4.14 (bf16: 3.04 -> 2.97) with BPTT alone.  Default:
    // both directions while all recurrence CTAs of the wavefront take at most half of the SMs, else BPTT only.
    h->persist = 0;
    if (rc == CRNN_OK && h->tc && h->wf_W > 1 && h->wf_C > 1 && !h->tcp.gather && stream_value_fn(false) && stream_value_fn(true)) {
        const bool small = h->wf_W * h->tcp.nslices * h->tcp.ntiles * 2 <= h->sm_count;
        const int pe = env_int("CRNN_WF_PERSIST", small ? 1 : 2);          // 1: both directions, 2: BPTT only, 3: forward only
        h->persist = pe == 1 ? 3 : (pe == 2 ? 2 : (pe == 3 ? 1 : (pe == 4 ? 6 : 0)));   // bit 0: forward, bit 1: BPTT, bit 2: forward, layer 0 only
        // Chunks cost no launch in this form, so BPTT of the large wavefronts is cut finer (less skew between the layers):
        // config 3, ms per step at 10 / 20 / 25 / 32 chunks: 4.16 / 4.13 / 4.09 / 4.15 (config 1 prefers its 5).
        if ((h->persist & 2) && !small && env_int("CRNN_WAVEFRONT_CHUNKS_BWD", 0) <= 0) {
            int cb = std::max(h->wf_C, std::min(32, cfg->seq_length / 4));
            while (cb > h->wf_C && cfg->num_layers * (h->wf_C + cb) > 384) --cb;
            h->wf_Cb = cb;
        }
    }
    if (rc == CRNN_OK && h->tc) {
        if (cudaHostAlloc((void**)&h->err_word, 64, cudaHostAllocMapped | cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); rc = set_err(CRNN_ERR_CUDA, "cudaHostAlloc failed"); }
        else *h->err_word = 0u;
    }
    if (rc == CRNN_OK && h->tc && cudaStreamCreateWithFlags(&h->aux_stream, cudaStreamNonBlocking) != cudaSuccess) { h->aux_stream = nullptr; rc = set_err(CRNN_ERR_CUDA, "cudaStreamCreate failed"); }
    if (rc != CRNN_OK) { for (cudaStream_t s_ : h->wf_streams) cudaStreamDestroy(s_); if (h->aux_stream) cudaStreamDestroy(h->aux_stream); if (h->err_word) cudaFreeHost(h->err_word); delete h; return rc; }
    build_workspace(h);
    if (env_int("CRNN_VERBOSE", 0) && h->tc)
        fprintf(stderr, "[crnn] plan: HC %d BT %d slices %d tiles %d passes %d x %d rows%s, K split fwd/bwd %d/%d, ring %d/%d KB, smem %zu/%zu, "
                "wavefront %d layers x %d chunks%s, workspace %.1f MB\n", h->tcp.HC, h->tcp.BT, h->tcp.nslices, h->tcp.ntiles, h->tcp.npass, h->tcp.pr,
                h->tcp.sub ? " (sub-chains)" : (h->tcp.cx_fwd && h->tcp.cx_bwd ? " (cluster exchange)" : h->tcp.cx_fwd ? " (cluster exchange forward)" : h->tcp.cx_bwd ? " (cluster exchange backward)" : ""), h->tcp.q_fwd, h->tcp.q_bwd, h->tcp.ring_fwd / 1024, h->tcp.ring_bwd / 1024, h->tcp.smem_fwd,
                h->tcp.smem_bwd, h->wf_W, h->wf_C, h->persist == 3 ? " (one launch per layer and direction)" : h->persist == 2 ? " (BPTT: one launch per layer)" : h->persist == 1 ? " (forward: one launch per layer)" : "", h->ws.total / 1e6);
    *out = h;
    return CRNN_OK;
}

void crnn_destroy(crnn_handle* h) {
    if (!h) return;
    for (cudaEvent_t e : h->prof_ev) cudaEventDestroy(e);
    for (cudaEvent_t e : h->kev) cudaEventDestroy(e);
    for (cudaEvent_t e : h->wf_events) if (e) cudaEventDestroy(e);
    for (cudaStream_t s : h->wf_streams) cudaStreamDestroy(s);
    if (h->aux_stream) cudaStreamDestroy(h->aux_stream);
    if (h->err_word) cudaFreeHost(h->err_word);
    delete h;
}

long long crnn_launch_count(void) { return g_launches; }

int crnn_profile_enable(crnn_handle* h, int32_t on) {
    if (!h) return set_err(CRNN_ERR_ARG, "null handle");
    h->prof_on = on != 0;
    h->prof_n = 0;
    h->kev_n = 0;
    return CRNN_OK;
}
int crnn_profile_read_kernels(crnn_handle* h, double* ms_out, int64_t* launches_out, int32_t reset) {
    if (!h || !ms_out || !launches_out) return set_err(CRNN_ERR_ARG, "null argument");
    for (int i = 0; i + 1 < h->kev_n; i += 2) {          // (begin, end) pairs
        CRNN_CUDA(cudaEventSynchronize(h->kev[i + 1]));
        float ms = 0.f;
        CRNN_CUDA(cudaEventElapsedTime(&ms, h->kev[i], h->kev[i + 1]));
        const int id = h->kev_id[i];
        if (id == 0 || id == 1) { h->kern_ms[id] += ms; ++h->kern_launches[id]; }
    }
    h->kev_n = 0;
    for (int i = 0; i < 2; ++i) { ms_out[i] = h->kern_ms[i]; launches_out[i] = h->kern_launches[i]; }
    if (reset) { h->kern_ms[0] = h->kern_ms[1] = 0.0; h->kern_launches[0] = h->kern_launches[1] = 0; }
    return CRNN_OK;
}
const char* crnn_profile_phase_name(int32_t i) { return (i >= 0 && i < CRNN_NUM_PHASES) ? kPhaseNames[i] : ""; }
int crnn_profile_read(crnn_handle* h, double* ms_out, int32_t reset) {
    if (!h || !ms_out) return set_err(CRNN_ERR_ARG, "null argument");
    if (h->prof_n > 0) {
        CRNN_CUDA(cudaEventSynchronize(h->prof_ev[h->prof_n - 1]));
        for (int i = 0; i + 1 < h->prof_n; ++i) {
            int id = h->prof_id[i];
            if (id < 0 || id >= CRNN_NUM_PHASES) continue;
            float ms = 0.f;
            CRNN_CUDA(cudaEventElapsedTime(&ms, h->prof_ev[i], h->prof_ev[i + 1]));
            h->prof_ms[id] += ms;
        }
        h->prof_n = 0;
    }
    for (int i = 0; i < CRNN_NUM_PHASES; ++i) ms_out[i] = h->prof_ms[i];
    if (reset) for (int i = 0; i < CRNN_NUM_PHASES; ++i) h->prof_ms[i] = 0.0;
    return CRNN_OK;
}

int64_t crnn_flat_size(const crnn_handle* h) { return h ? h->P : -1; }
int32_t crnn_num_tensors(const crnn_handle* h) { return h ? (int32_t)h->tensors.size() : -1; }
int crnn_param_layout(const crnn_handle* h, int32_t i, char* name, int32_t cap, int64_t* off, int64_t* rows, int64_t* cols) {
    if (!h || i < 0 || i >= (int32_t)h->tensors.size()) return set_err(CRNN_ERR_ARG, "tensor index out of range");
    const TensorInfo& t = h->tensors[i];
    if (name && cap > 0) { strncpy(name, t.name.c_str(), cap - 1); name[cap - 1] = 0; }
    if (off) *off = t.offset;
    if (rows) *rows = t.rows;
    if (cols) *cols = t.cols;
    return CRNN_OK;
}
int64_t crnn_state_size(const crnn_handle* h) {
    return h ? (int64_t)h->cfg.num_layers * h->nstate * h->cfg.batch_size * h->cfg.rnn_size : -1;
}
int64_t crnn_workspace_bytes(const crnn_handle* h) { return h ? (int64_t)h->ws.total : -1; }

}  // extern "C"

// ---- launch helpers ------------------------------------------------------------------

template <int MODEL>
static int launch_rec(const RecPlan& p, bool backward, RecArgs& a, cudaStream_t st) {
    void (*fn)(RecArgs) = backward ? rec_bwd_kernel<MODEL> : rec_fwd_kernel<MODEL>;
    CRNN_CUDA(cudaFuncSetAttribute((const void*)fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem));
    a.HC = p.HC; a.BT = p.BT; a.nslices = p.nslices; a.ntiles = p.ntiles; a.TXa = p.TXa; a.TXb = p.TXb;
    void* args[] = {&a};
    // cooperative launch: all CTAs co-resident, or the launch fails instead of deadlocking
    CRNN_CUDA(cudaLaunchCooperativeKernel((const void*)fn, dim3(p.nslices * p.ntiles), dim3(REC_THREADS), args, p.smem, st));
    ++g_launches;
    return CRNN_OK;
}

static int launch_rec_model(const crnn_handle* h, bool backward, RecArgs& a, cudaStream_t st) {
    const RecPlan& p = backward ? h->bwd : h->fwd;
    switch (h->cfg.model) {
        case CRNN_MODEL_LSTM: return launch_rec<CRNN_MODEL_LSTM>(p, backward, a, st);
        case CRNN_MODEL_GRU:  return launch_rec<CRNN_MODEL_GRU>(p, backward, a, st);
        default:              return launch_rec<CRNN_MODEL_RNN>(p, backward, a, st);
    }
}

struct LayerParams {
    const float *W, *b, *Wc, *bc;   // gru: W/b = gates, Wc/bc = candidate
    int64_t oW, ob, oWc, obc;       // offsets into the flat buffers
};

static LayerParams layer_params(const crnn_handle* h, const float* params, int l) {
    LayerParams lp{};
    const int per = h->cfg.model == CRNN_MODEL_GRU ? 4 : 2;
    const TensorInfo* t = &h->tensors[3 + l * per];
    lp.oW = t[0].offset; lp.ob = t[1].offset;
    lp.W = params + lp.oW; lp.b = params + lp.ob;
    if (per == 4) { lp.oWc = t[2].offset; lp.obc = t[3].offset; lp.Wc = params + lp.oWc; lp.bc = params + lp.obc; }
    return lp;
}

static inline float* wsf(void* ws, size_t off) { return reinterpret_cast<float*>((char*)ws + off); }

static int grid_for(int64_t n_threads_needed, int sm_count) {
    int64_t blocks = ceil_div64(n_threads_needed, 256);
    return (int)std::max<int64_t>(1, std::min<int64_t>(blocks, (int64_t)sm_count * 8));
}

static inline double* ws_loss_part(void* ws) { return reinterpret_cast<double*>((char*)ws + 8192); }   // one slot per softmax block (<= 8 * SMs)

// Token order by id for the sorted embedding scatter: depends on the ids only, so it runs early in the step (beside the
// forward pass where a second stream exists).
static int scatter_sort(crnn_handle* h, const int32_t* x, int64_t ld, void* ws, cudaStream_t st) {
    if (!h->ws.es_sorted) return CRNN_OK;
    const crnn_config& c = h->cfg;
    const int V = c.vocab_size, B = c.batch_size;
    const int64_t N = (int64_t)c.seq_length * B;
    int* hist = reinterpret_cast<int*>((char*)ws + h->ws.es_hist);
    int* lrank = reinterpret_cast<int*>((char*)ws + h->ws.es_lrank);
    int* order = reinterpret_cast<int*>((char*)ws + h->ws.es_order);
    int* off = reinterpret_cast<int*>((char*)ws + h->ws.es_off);
    int* poff = reinterpret_cast<int*>((char*)ws + h->ws.es_poff);
    int2* piece = reinterpret_cast<int2*>((char*)ws + h->ws.es_piece);
    int* np = reinterpret_cast<int*>((char*)ws + h->ws.es_np);
    es_rank_kernel<<<h->ws.es_nblk, ES_SORT_BLOCK, V * sizeof(int), st>>>(x, ld, B, N, V, hist, lrank);
    CRNN_LAUNCH_CHECK("es_rank_kernel");
    es_offsets_kernel<<<1, 1024, 2 * V * sizeof(int), st>>>(h->ws.es_nblk, V, h->ws.es_piece_rows, hist, off, poff, piece, np);
    CRNN_LAUNCH_CHECK("es_offsets_kernel");
    es_place_kernel<<<grid_for(N, h->sm_count), 256, 0, st>>>(x, ld, B, N, V, hist, off, lrank, order);
    CRNN_LAUNCH_CHECK("es_place_kernel");
    return CRNN_OK;
}

// Embedding gradient (scatter-add of the slices) + sum of squares of the slices + the two trailing scalars.
static int scatter_and_finalize(crnn_handle* h, const int32_t* x, int64_t ld, const float* dX, float* grads, void* ws,
                                int nloss, cudaStream_t st) {
    const crnn_config& c = h->cfg;
    const int H = c.rnn_size, V = c.vocab_size, S = c.seq_length, B = c.batch_size;
    const int64_t SB = (int64_t)S * B;
    if (nloss >= 0) prof_mark(h, st, CRNN_PHASE_EMBED_SCATTER);      // (nloss < 0: on a side stream, outside the phase profile)
    if (h->ws.es_sorted) {                            // (the token order was built by scatter_sort, early in the step)
        const int nchunk = ceil_div(H, 4 * ES_SUM_THREADS);
        const int* order = reinterpret_cast<const int*>((char*)ws + h->ws.es_order);
        const int2* piece = reinterpret_cast<const int2*>((char*)ws + h->ws.es_piece);
        const int* np = reinterpret_cast<const int*>((char*)ws + h->ws.es_np);
        const int* poff = reinterpret_cast<const int*>((char*)ws + h->ws.es_poff);
        float* part = wsf(ws, h->ws.es_part);
        double* sqp = reinterpret_cast<double*>((char*)ws + h->ws.es_sq);
        if (h->ws.es_piece_rows == 32) es_sum_kernel<32><<<dim3(nchunk, h->ws.es_pmax), ES_SUM_THREADS, 0, st>>>(H, dX, order, piece, np, part, sqp);
        else                           es_sum_kernel<64><<<dim3(nchunk, h->ws.es_pmax), ES_SUM_THREADS, 0, st>>>(H, dX, order, piece, np, part, sqp);
        CRNN_LAUNCH_CHECK("es_sum_kernel");
        es_finish_kernel<<<grid_for((int64_t)V * H, h->sm_count), 256, 0, st>>>(grads + h->emb_off, V, H, poff, part, sqp, np, nchunk, ws_sq(ws));
        CRNN_LAUNCH_CHECK("es_finish_kernel");
    } else if (h->ws.es_nseg > 0) {
        const int nchunk = ceil_div(H, ES_COLS);
        const int smem = (int)es_smem_bytes(V);
        static bool attr_done[64] = {};
        if (!attr_done[c.device & 63]) {
            CRNN_CUDA(cudaFuncSetAttribute(embed_scatter_part_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, h->max_smem - 1024));
            attr_done[c.device & 63] = true;
        }
        float* part = wsf(ws, h->ws.es_part);
        double* sqp = reinterpret_cast<double*>((char*)ws + h->ws.es_sq);
        embed_scatter_part_kernel<<<dim3(nchunk, h->ws.es_nseg), ES_COLS, smem, st>>>(V, H, x, ld, B, S, dX, h->ws.es_tps, part, sqp);
        CRNN_LAUNCH_CHECK("embed_scatter_part_kernel");
        embed_scatter_reduce_kernel<<<grid_for((int64_t)V * H, h->sm_count), 256, 0, st>>>(grads + h->emb_off, (int64_t)V * H, h->ws.es_nseg, part, sqp,
                                                                                           h->ws.es_nseg * nchunk, ws_sq(ws));
        CRNN_LAUNCH_CHECK("embed_scatter_reduce_kernel");
    } else {
        embed_scatter_kernel<<<grid_for(SB * H, h->sm_count), 256, 0, st>>>(grads + h->emb_off, V, H, x, ld, B, S, dX, ws_sq(ws));
        CRNN_LAUNCH_CHECK("embed_scatter_kernel");
    }
    if (nloss < 0) return CRNN_OK;                    // (scatter only: the caller finalises after joining its streams)
    finalize_scalars_kernel<<<1, 256, 0, st>>>(ws_sq(ws), ws_loss(ws), (double)SB, grads + h->P, ws_loss_part(ws), nloss);
    CRNN_LAUNCH_CHECK("finalize_scalars_kernel");
    return CRNN_OK;
}

// ---- dropout sites ---------------------------------------------------------------------------
// site l < L: what cell l reads = (embedding mask | output mask of cell l-1) x input mask of cell l;
// site L: the top cell's output as the projection reads it.
static DropArgs drop_args(const crnn_handle* h, long long call, int site, int64_t idx0) {
    const int L = h->cfg.num_layers;
    DropArgs d;
    d.seed = (uint64_t)(uint32_t)h->cfg.dropout_seed * 0x100000001B3ull + (uint64_t)call * 0x9E3779B97F4A7C15ull + 0x5851F42D4C957F2Dull;
    d.idx0 = idx0;
    if (site >= L) { d.stage_a = 2 + 2 * (L - 1); d.keep_a = h->okp; d.stage_b = 0; d.keep_b = 1.0f; }
    else if (site == 0) { d.stage_a = 0; d.keep_a = h->okp; d.stage_b = 1; d.keep_b = h->ikp; }
    else { d.stage_a = 2 + 2 * (site - 1); d.keep_a = h->okp; d.stage_b = 1 + 2 * site; d.keep_b = h->ikp; }
    return d;
}
// dst(/dst_b) = src * masks of `site` for elements [idx0, idx0+n) of the [S,B,H] tensor.
static int drop_apply(const crnn_handle* h, long long call, int site, const float* src, float* dst, __nv_bfloat16* dst_b,
                      int64_t idx0, int64_t n, cudaStream_t st, int64_t lo_ofs = 0) {
    if (n <= 0) return CRNN_OK;
    const int64_t blocks = std::max<int64_t>(1, std::min<int64_t>(ceil_div64(n, 256), (int64_t)h->sm_count * 8));
    dropout_apply_kernel<<<(int)blocks, 256, 0, st>>>(src, dst, dst_b, n, drop_args(h, call, site, idx0), lo_ofs);
    CRNN_LAUNCH_CHECK("dropout_apply_kernel");
    return CRNN_OK;
}

// Forward pass over all layers; leaves hx/cx/gates stashes in the workspace.
static int run_forward(crnn_handle* h, const int32_t* x, int64_t ldx, const float* state_in, float* state_out,
                       const float* params, void* ws, cudaStream_t st, long long drop_call = -1) {
    const crnn_config& c = h->cfg;
    const bool drop = h->drop_on && drop_call >= 0;       // training calls only (model.py:31, training=True)
    const int H = c.rnn_size, V = c.vocab_size, L = c.num_layers, S = c.seq_length, B = c.batch_size;
    const int64_t BH = (int64_t)B * H, SB = (int64_t)S * B;
    const int gw = h->GS * H;
    float* x0 = wsf(ws, h->ws.x0);
    const float* emb = params + h->emb_off;
    prof_mark(h, st, CRNN_PHASE_EMBED_GATHER);
    embed_gather_kernel<<<grid_for(SB * H / 4 + 1, h->sm_count), 256, 0, st>>>(emb, V, H, x, ldx, B, S, x0, nullptr);
    CRNN_LAUNCH_CHECK("embed_gather_kernel");
    prof_mark(h, st, CRNN_PHASE_MISC);
    CRNN_CUDA(cudaMemsetAsync(ws_barrier(ws, 0), 0, 512 * 256, st));
    for (int l = 0; l < L; ++l) {
        LayerParams lp = layer_params(h, params, l);
        float* hx = wsf(ws, h->ws.hx[l]);
        float* cx = c.model == CRNN_MODEL_LSTM ? wsf(ws, h->ws.cx[l]) : nullptr;
        float* gates = wsf(ws, h->ws.gates[l]);
        const float* in = l == 0 ? x0 : wsf(ws, h->ws.hx[l - 1]) + BH;
        if (drop) {
            int rcd = drop_apply(h, drop_call, l, in, wsf(ws, h->ws.xin[l]), nullptr, 0, SB * H, st);
            if (rcd != CRNN_OK) return rcd;
            in = wsf(ws, h->ws.xin[l]);
        }
        const float* sin = state_in + (int64_t)l * h->nstate * BH;
        if (c.model == CRNN_MODEL_LSTM) {
            CRNN_CUDA(cudaMemcpyAsync(cx, sin, BH * 4, cudaMemcpyDeviceToDevice, st));
            CRNN_CUDA(cudaMemcpyAsync(hx, sin + BH, BH * 4, cudaMemcpyDeviceToDevice, st));
        } else {
            CRNN_CUDA(cudaMemcpyAsync(hx, sin, BH * 4, cudaMemcpyDeviceToDevice, st));
        }
        // hoisted x half of [x,h]*W (+ bias) for all S steps at once
        prof_mark(h, st, CRNN_PHASE_XPROJ_GEMM);
        int rc;
        if (c.model == CRNN_MODEL_LSTM) {
            rc = gemm_f32(0, 0, (int)SB, 4 * H, H, in, H, lp.W, 4 * H, gates, gw, lp.b, 0, st, h->sm_count);
        } else if (c.model == CRNN_MODEL_GRU) {
            rc = gemm_f32(0, 0, (int)SB, 2 * H, H, in, H, lp.W, 2 * H, gates, gw, lp.b, 0, st, h->sm_count);
            if (rc == CRNN_OK) rc = gemm_f32(0, 0, (int)SB, H, H, in, H, lp.Wc, H, gates + 2 * H, gw, lp.bc, 0, st, h->sm_count);
        } else {
            rc = gemm_f32(0, 0, (int)SB, H, H, in, H, lp.W, H, gates, gw, lp.b, 0, st, h->sm_count);
        }
        if (rc != CRNN_OK) return rc;
        RecArgs a{};
        a.B = B; a.H = H; a.S = S; a.GS = h->GS;
        a.W = lp.W; a.Wc = lp.Wc; a.gates = gates; a.hx = hx; a.cx = cx;
        a.barrier = ws_barrier(ws, 2 * l);
        prof_mark(h, st, CRNN_PHASE_REC_FWD);
        rc = launch_rec_model(h, false, a, st);
        if (rc != CRNN_OK) return rc;
        prof_mark(h, st, CRNN_PHASE_MISC);
        float* sout = state_out + (int64_t)l * h->nstate * BH;
        if (c.model == CRNN_MODEL_LSTM) {
            CRNN_CUDA(cudaMemcpyAsync(sout, cx + (int64_t)S * BH, BH * 4, cudaMemcpyDeviceToDevice, st));
            CRNN_CUDA(cudaMemcpyAsync(sout + BH, hx + (int64_t)S * BH, BH * 4, cudaMemcpyDeviceToDevice, st));
        } else {
            CRNN_CUDA(cudaMemcpyAsync(sout, hx + (int64_t)S * BH, BH * 4, cudaMemcpyDeviceToDevice, st));
        }
    }
    // logits (time-major rows t*B+b) = top outputs * softmax_w + softmax_b   (model.py:76)
    const float* top = wsf(ws, h->ws.hx[L - 1]) + BH;
    if (drop) {
        int rcd = drop_apply(h, drop_call, L, top, wsf(ws, h->ws.topd), nullptr, 0, SB * H, st);
        if (rcd != CRNN_OK) return rcd;
        top = wsf(ws, h->ws.topd);
    }
    const float* Wsm = params + h->tensors[0].offset;
    const float* bsm = params + h->tensors[1].offset;
    prof_mark(h, st, CRNN_PHASE_LOGITS_GEMM);
    int rc = gemm_f32(0, 0, (int)SB, V, H, top, H, Wsm, V, wsf(ws, h->ws.logits), V, bsm, 0, st, h->sm_count);
    prof_mark(h, st, CRNN_PHASE_MISC);
    return rc;
}


// ======================================================================================
// CRNN_PREC_BF16: tensor-core path.  Same phase structure as the fp32 path; every
// contraction runs on tcgen05 with bf16 operands that the producing kernels wrote.
// ======================================================================================
typedef __nv_bfloat16 bf16_t;
static inline bf16_t* wsb(void* ws, size_t off) { return reinterpret_cast<bf16_t*>((char*)ws + off); }

template <int MODEL>
static int launch_rec_tc(crnn_handle* h, bool backward, tc::RecTcArgs& a, const bf16_t* pbf_W, const bf16_t* pbf_Wc,
                         const bf16_t* wr, cudaStream_t st) {
    const RecPlanTc& p = h->tcp;
    const int H = h->cfg.rnn_size, S = h->cfg.seq_length, B = h->cfg.batch_size;
    const int g1 = MODEL == CRNN_MODEL_LSTM ? 4 : (MODEL == CRNN_MODEL_GRU ? 2 : 1);
    const int GW = h->G, GZ = h->GZ;
    const int Q = backward ? p.q_bwd : p.q_fwd;
    const int HCG = p.HC * Q;                        // units per group (the columns one CTA contracts)
    a.ksplit = Q;
    a.HC = p.HC; a.BT = p.BT; a.nslices = p.nslices; a.ntiles = p.ntiles; a.npass = p.npass; a.box_rows = p.box_rows;
    a.pr = p.pr; a.sub = p.sub;
    a.fast_act = h->fast_act;
    a.zs_floats = backward ? p.zs_bwd : p.zs_fwd;
    a.swap = backward ? p.swap_bwd : p.swap_fwd;
    a.m64 = (!a.swap && std::min(p.pr, p.BT) <= 64 && h->env.m64) ? 1 : 0;
    a.dbg = (!backward && h->env.dbg && h->dbg_layer == h->env.dbg_layer) ? h->dbg_ptr : nullptr;
    CUtensorMap t0, t1, w0, w1, t0l, t1l, w0l, w1l;
    int rc;
    const int split = h->split;
    a.split = split;
    // lo twins sit right behind their hi arrays (build_workspace): element offsets
    const int64_t lo_pbf = split ? h->P : 0;
    const int64_t lo_wr = split ? (int64_t)p.nslices * GW * p.HC * H : 0;
    a.hxb_lo = split ? (long long)(S + 1) * B * H : 0;
    a.rsb_lo = split ? (long long)S * B * H : 0;
    a.dzb_lo = split ? (long long)(S + 1) * B * GZ * H : 0;
    // ring geometry is per launch (every contraction of the kernel shares the ring and its barriers):
    // sized for the shortest K extent, H
    a.ckb = tc::rt_chunk_kb(p.box_rows, ceil_div(H, 64), split);
    a.ring_bytes = backward ? p.ring_bwd : p.ring_fwd;
    {   // the whole H extent (of this CTA's K part) as one stage when the ring holds it
        const int nkb_h = ceil_div(H, 64) / Q, tile = tc::rt_tile_bytes(p.box_rows);
        if (h->env.one_stage && p.npass == 1 && nkb_h * tile * (1 + split) + tc::RT_MROWS_BYTES - tile <= a.ring_bytes) a.ckb = nkb_h;
    }
    { const int ck_env = h->env.ckb; if (ck_env > 0) a.ckb = std::min(ck_env, ceil_div(H, 64)); }
    a.rotate = h->env.rotate;
    a.ring_bytes = backward ? p.ring_bwd : p.ring_fwd;
    {   // at least two stages must fit in the ring
        const int tile = tc::rt_tile_bytes(p.box_rows);
        while (a.ckb > 1 && (a.ring_bytes - tc::RT_MROWS_BYTES + tile) / (tile * a.ckb * (1 + split)) < h->env.min_stages) --a.ckb;
    }
    a.gather = p.gather; a.g_rp_log2 = p.g_rp_log2; a.g_ckb_log2 = p.g_ckb_log2;
    a.err = h->err_word;
    if (a.gather) a.ckb = 1 << p.g_ckb_log2;
    a.nst = tc::rt_num_stages(p.box_rows, a.ckb, split, a.ring_bytes);
    a.use3d = (H % 64 == 0 && h->env.use3d && !(backward ? p.cx_bwd : p.cx_fwd)) ? 1 : 0;      // (cluster exchange: the map only serves the launch's first load, K block by K block)
    auto amap = [&](CUtensorMap* m, const bf16_t* base, int64_t rows, int64_t cols, int64_t ld) -> int {
        if (a.use3d) return tc::make_tmap_bf16_kblocks(m, base, rows, cols, ld, p.box_rows, a.ckb);
        return tc::make_tmap_bf16(m, base, rows, cols, ld, p.box_rows);
    };
    if (!backward) {
        rc = amap(&t0, a.hxb, (int64_t)(S + 1) * B, H, H);
        if (rc == CRNN_OK) rc = MODEL == CRNN_MODEL_GRU ? amap(&t1, a.rsb, (int64_t)S * B, H, H) : (t1 = t0, CRNN_OK);
        if (rc == CRNN_OK) rc = tc::make_tmap_bf16(&w0, wr, (int64_t)p.nslices * GW * p.HC, H, H, g1 * HCG);
        if (rc == CRNN_OK) rc = MODEL == CRNN_MODEL_GRU ? tc::make_tmap_bf16(&w1, wr, (int64_t)p.nslices * GW * p.HC, H, H, HCG) : (w1 = w0, CRNN_OK);
        t0l = t0; t1l = t1; w0l = w0; w1l = w1;
        if (rc == CRNN_OK && split) {
            rc = amap(&t0l, a.hxb + a.hxb_lo, (int64_t)(S + 1) * B, H, H);
            if (rc == CRNN_OK) rc = MODEL == CRNN_MODEL_GRU ? amap(&t1l, a.rsb + a.rsb_lo, (int64_t)S * B, H, H) : (t1l = t0l, CRNN_OK);
            if (rc == CRNN_OK) rc = tc::make_tmap_bf16(&w0l, wr + lo_wr, (int64_t)p.nslices * GW * p.HC, H, H, g1 * HCG);
            if (rc == CRNN_OK) rc = MODEL == CRNN_MODEL_GRU ? tc::make_tmap_bf16(&w1l, wr + lo_wr, (int64_t)p.nslices * GW * p.HC, H, H, HCG) : (w1l = w0l, CRNN_OK);
        }
    } else {
        const int64_t zw = (int64_t)GZ * H;
        rc = amap(&t0, a.dzb, (int64_t)(S + 1) * B, (int64_t)g1 * H, zw);
        if (rc == CRNN_OK) rc = MODEL == CRNN_MODEL_GRU ? amap(&t1, a.dzb + 2 * H, (int64_t)(S + 1) * B, H, zw) : (t1 = t0, CRNN_OK);
        if (rc == CRNN_OK) rc = tc::make_tmap_bf16(&w0, pbf_W, 2 * H, (int64_t)g1 * H, (int64_t)g1 * H, HCG);
        if (rc == CRNN_OK) rc = MODEL == CRNN_MODEL_GRU ? tc::make_tmap_bf16(&w1, pbf_Wc, 2 * H, H, H, HCG) : (w1 = w0, CRNN_OK);
        t0l = t0; t1l = t1; w0l = w0; w1l = w1;
        if (rc == CRNN_OK && split) {
            rc = amap(&t0l, a.dzb + a.dzb_lo, (int64_t)(S + 1) * B, (int64_t)g1 * H, zw);
            if (rc == CRNN_OK) rc = MODEL == CRNN_MODEL_GRU ? amap(&t1l, a.dzb + a.dzb_lo + 2 * H, (int64_t)(S + 1) * B, H, zw) : (t1l = t0l, CRNN_OK);
            if (rc == CRNN_OK) rc = tc::make_tmap_bf16(&w0l, pbf_W + lo_pbf, 2 * H, (int64_t)g1 * H, (int64_t)g1 * H, HCG);
            if (rc == CRNN_OK) rc = MODEL == CRNN_MODEL_GRU ? tc::make_tmap_bf16(&w1l, pbf_Wc + lo_pbf, 2 * H, H, H, HCG) : (w1l = w0l, CRNN_OK);
        }
    }
    if (rc != CRNN_OK) return rc;
    // elements per epilogue thread per batch: one batch per 128-row pass when 8 per thread cover it, so every global
    // load of the pass is in flight before the accumulator arrives
    const int per_pass = std::min(p.pr, p.BT) * p.HC;
    const bool ub8 = per_pass > tc::RT_EPI_THREADS * 4 && h->env.ub8;
    a.defer = h->env.defer;
    a.acq_poll = h->env.acq_poll;
    a.expt = h->env.expt;
    // K-split launches read their partial sums from distributed shared memory; there consecutive lanes on consecutive
    // units (scalar epilogue) beat quads (config 4 BPTT 5.4 vs 6.6 ms, split-precision config 3 BPTT 2.9 vs 3.6 ms)
    const bool vec4 = (backward ? p.cx_bwd : p.cx_fwd) || (h->env.vec4 >= 0 ? h->env.vec4 : (Q > 1 ? 0 : 1)) != 0;   // (the cluster exchange pushes quads)
    typedef void (*RecFn)(CUtensorMap, CUtensorMap, CUtensorMap, CUtensorMap, CUtensorMap, CUtensorMap, CUtensorMap, CUtensorMap, tc::RecTcArgs);
    RecFn fn;
#define CRNN_PICK(KERNEL, F)                                                                                   \
    (ub8 ? (vec4 ? (RecFn)tc::KERNEL<MODEL, F, 8, 4> : (RecFn)tc::KERNEL<MODEL, F, 8, 1>)                       \
         : (vec4 ? (RecFn)tc::KERNEL<MODEL, F, 4, 4> : (RecFn)tc::KERNEL<MODEL, F, 4, 1>))
    if (h->fast_act) fn = backward ? CRNN_PICK(rec_bwd_tc_kernel, true) : CRNN_PICK(rec_fwd_tc_kernel, true);
    else             fn = backward ? CRNN_PICK(rec_bwd_tc_kernel, false) : CRNN_PICK(rec_fwd_tc_kernel, false);
#undef CRNN_PICK
    const size_t smem = backward ? p.smem_bwd : p.smem_fwd;
    {   // the attribute is per function and device; set it again only when the size changes
        static std::unordered_map<uint64_t, int> smem_set;
        const uint64_t key = (uint64_t)(uintptr_t)fn ^ ((uint64_t)h->cfg.device << 56);
        auto it = smem_set.find(key);
        if (it == smem_set.end() || it->second != (int)smem) {
            CRNN_CUDA(cudaFuncSetAttribute((const void*)fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            smem_set[key] = (int)smem;
        }
    }
    int csize = 1;
    a.cbar = 0;
    // Cluster barrier: when ALL slices of a batch tile fit in one cluster, the steps are ordered by remote
    // mbarrier arrives inside the cluster instead of a global counter polled through L2, and nothing ties
    // different clusters together any more (no cooperative launch needed).  On by default for portable
    // cluster sizes (config 1: 4 slices, -2 %); 16-CTA clusters (CRNN_TC_CBAR=16) gain 4 % per kernel on
    // config 3 but eight of them with ~190 KB of shared memory per CTA are not co-scheduled, which costs
    // the layer wavefront far more than that.
    const int cbar_max = h->env.cbar;
    if (!a.gather && !a.sub && p.nslices <= cbar_max && (p.nslices == 2 || p.nslices == 4 || p.nslices == 8 || p.nslices == 16)) {
        csize = p.nslices;
        a.cbar = 1;
    }
    if (Q > 1) { csize = Q; a.cbar = 0; }     // the cluster is the K-split group
    a.xchg = 0;
    if ((backward ? p.cx_bwd : p.cx_fwd) && Q == 1) { a.xchg = 1; a.cbar = 0; csize = p.nslices; }   // cluster exchange: the rows themselves travel inside the cluster
    if (csize > 8) CRNN_CUDA(cudaFuncSetAttribute((const void*)fn, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    a.csize = csize;
    void* args[] = {&t0, &t1, &w0, &w1, &t0l, &t1l, &w0l, &w1l, &a};
    prof_kernel_mark(h, st, backward ? 1 : 0);
    if (csize == 1) {
        CRNN_CUDA(cudaLaunchCooperativeKernel((const void*)fn, dim3(p.nslices * p.ntiles), dim3(tc::RT_THREADS), args, smem, st));
    } else {
        cudaLaunchConfig_t cfgl{};
        cfgl.gridDim = dim3(p.nslices * p.ntiles);
        cfgl.blockDim = dim3(tc::RT_THREADS);
        cfgl.dynamicSmemBytes = smem;
        cfgl.stream = st;
        cudaLaunchAttribute attrs[2];
        attrs[0].id = cudaLaunchAttributeClusterDimension;
        attrs[0].val.clusterDim.x = (unsigned)csize; attrs[0].val.clusterDim.y = 1; attrs[0].val.clusterDim.z = 1;
        attrs[1].id = cudaLaunchAttributeCooperative;           // all clusters co-resident: tiles span several clusters
        attrs[1].val.cooperative = 1;
        cfgl.attrs = attrs; cfgl.numAttrs = (a.cbar || a.xchg) ? 1 : 2;     // cluster barrier / exchange: clusters are independent
        CRNN_CUDA(cudaLaunchKernelExC(&cfgl, (const void*)fn, args));
    }
    prof_kernel_mark(h, st, backward ? 1 : 0);
    ++g_launches;
    return CRNN_OK;
}

static int launch_rec_tc_model(crnn_handle* h, bool backward, tc::RecTcArgs& a, const bf16_t* W, const bf16_t* Wc,
                               const bf16_t* wr, cudaStream_t st) {
    switch (h->cfg.model) {
        case CRNN_MODEL_LSTM: return launch_rec_tc<CRNN_MODEL_LSTM>(h, backward, a, W, Wc, wr, st);
        case CRNN_MODEL_GRU:  return launch_rec_tc<CRNN_MODEL_GRU>(h, backward, a, W, Wc, wr, st);
        default:              return launch_rec_tc<CRNN_MODEL_RNN>(h, backward, a, W, Wc, wr, st);
    }
}

// bf16 shadows of the parameters: the flat buffer as is (pbf), softmax_w re-pitched to Vp,
// and the per-layer recurrent halves packed for the forward kernel.
static int refresh_bf16_params(crnn_handle* h, const float* params, void* ws, cudaStream_t st) {
    const crnn_config& c = h->cfg;
    const int H = c.rnn_size, V = c.vocab_size, L = c.num_layers;
    const int64_t sp = h->split;                       // lo twins behind the hi arrays
    tc::cast_bf16_kernel<<<grid_for(h->P, h->sm_count), 256, 0, st>>>(params, h->P, 1, (int)h->P, wsb(ws, h->ws.pbf), h->P, sp * h->P);
    CRNN_LAUNCH_CHECK("cast_bf16_kernel");
    tc::cast_bf16_kernel<<<grid_for((int64_t)H * V, h->sm_count), 256, 0, st>>>(params + h->tensors[0].offset, V, H, V, wsb(ws, h->ws.wsmb), h->Vp,
                                                                                 sp * H * h->Vp);
    CRNN_LAUNCH_CHECK("cast_bf16_kernel");
    const int g1 = c.model == CRNN_MODEL_LSTM ? 4 : (c.model == CRNN_MODEL_GRU ? 2 : 1);
    for (int l = 0; l < L; ++l) {
        const int per = c.model == CRNN_MODEL_GRU ? 4 : 2;
        const float* W = params + h->tensors[3 + l * per].offset;
        const float* Wc = per == 4 ? params + h->tensors[3 + l * per + 2].offset : nullptr;
        const int64_t n = (int64_t)h->tcp.nslices * h->G * h->tcp.HC * H;
        // one packed group per K-split cluster: q_fwd * HC units, nslices / q_fwd groups (same total size)
        tc::pack_rec_weights_kernel<<<grid_for(n, h->sm_count), 256, 0, st>>>(W, Wc, H, h->tcp.HC * h->tcp.q_fwd, h->tcp.nslices / h->tcp.q_fwd,
                                                                              g1, h->G, H, wsb(ws, h->ws.wr[l]), sp * n);
        CRNN_LAUNCH_CHECK("pack_rec_weights_kernel");
    }
    return CRNN_OK;
}

// ---- wavefront helpers ---------------------------------------------------------------------
// Stream of layer l: the caller's stream when layers run one after another, else one of the W
// internal streams.  Events are created on first use and reused by every later step.
static cudaStream_t wf_stream(crnn_handle* h, cudaStream_t st, int l) {
    return h->wf_W > 1 ? h->wf_streams[l % h->wf_W] : st;
}
static int wf_event(crnn_handle* h, int idx, cudaEvent_t* out) {
    if ((int)h->wf_events.size() <= idx) h->wf_events.resize(idx + 1, nullptr);
    if (!h->wf_events[idx]) CRNN_CUDA(cudaEventCreateWithFlags(&h->wf_events[idx], cudaEventDisableTiming));
    *out = h->wf_events[idx];
    return CRNN_OK;
}
// The batch GEMMs that feed / drain a layer's recurrence go to a second stream per layer, so a
// chunk's GEMM never sits between two recurrence launches of the same layer.
static cudaStream_t wf_gstream(crnn_handle* h, cudaStream_t st, int l) {
    return h->wf_W > 1 ? h->wf_streams[h->wf_W + l % h->wf_W] : st;
}
static int wf_record(crnn_handle* h, int idx, cudaStream_t s) {
    cudaEvent_t e;
    int rc = wf_event(h, idx, &e);
    if (rc != CRNN_OK) return rc;
    CRNN_CUDA(cudaEventRecord(e, s));
    return CRNN_OK;
}
static int wf_wait(crnn_handle* h, int idx, cudaStream_t s) {
    cudaEvent_t e;
    int rc = wf_event(h, idx, &e);
    if (rc != CRNN_OK) return rc;
    CRNN_CUDA(cudaStreamWaitEvent(s, e, 0));
    return CRNN_OK;
}
// event ids: family f in {0: fwd recurrence done, 1: fwd x-projection done, 2: bwd recurrence done,
// 3: bwd input gradients done}: f*L*C + l*C + c ; fork/join events after those
static inline int wf_id(const crnn_handle* h, int fam, int l, int c) {
    return (fam * h->cfg.num_layers + l) * h->wf_C + c;
}
static int wf_fork(crnn_handle* h, cudaStream_t st, int slot) {
    if (h->wf_W <= 1) return CRNN_OK;
    cudaEvent_t e;
    int rc = wf_event(h, 4 * h->cfg.num_layers * h->wf_C + slot, &e);
    if (rc != CRNN_OK) return rc;
    CRNN_CUDA(cudaEventRecord(e, st));
    for (cudaStream_t s : h->wf_streams) CRNN_CUDA(cudaStreamWaitEvent(s, e, 0));
    return CRNN_OK;
}
static int wf_join(crnn_handle* h, cudaStream_t st, int slot) {
    if (h->wf_W <= 1) return CRNN_OK;
    for (int i = 0; i < 2 * h->wf_W; ++i) {
        cudaEvent_t e;
        int rc = wf_event(h, 4 * h->cfg.num_layers * h->wf_C + slot + i, &e);
        if (rc != CRNN_OK) return rc;
        CRNN_CUDA(cudaEventRecord(e, h->wf_streams[i]));
        CRNN_CUDA(cudaStreamWaitEvent(st, e, 0));
    }
    return CRNN_OK;
}
static inline int wf_t0(const crnn_handle* h, int c) { return (int)((int64_t)h->cfg.seq_length * c / h->wf_C); }

// Element offsets of the lo twins (0 unless CRNN_PREC_BF16X3), and the GEMM call that passes them on.
struct LoOfs {
    int64_t act, hxb, pbf, wsm, dzb, dlb;
};
static LoOfs lo_offsets(const crnn_handle* h) {
    const crnn_config& c = h->cfg;
    const int64_t sp = h->split, S = c.seq_length, B = c.batch_size, H = c.rnn_size;
    LoOfs o;
    o.act = sp * S * B * H;
    o.hxb = sp * (S + 1) * B * H;
    o.pbf = sp * h->P;
    o.wsm = sp * H * h->Vp;
    o.dzb = sp * (S + 1) * B * h->GZ * H;
    o.dlb = sp * S * B * h->Vp;
    return o;
}
static int gemm_tc(const crnn_handle* h, int M, int N, int K, const bf16_t* A, int64_t lda, int a_mn, int64_t a_lo,
                   const bf16_t* B, int64_t ldb, int b_mn, int64_t b_lo, float* C, int64_t ldc, const float* bias, int accumulate,
                   cudaStream_t st) {
    return tc::gemm_bf16_tc(M, N, K, A, lda, a_mn, B, ldb, b_mn, C, ldc, bias, accumulate, st, h->sm_count,
                            h->split ? A + a_lo : nullptr, h->split ? B + b_lo : nullptr);
}

// y != nullptr (training): the projection is fused with softmax + cross-entropy and *fused_blocks receives the number of
// loss / column-sum partials it wrote (0: vocabulary too large for the fused kernel, fp32 logits are in the workspace).
static int run_forward_bf16(crnn_handle* h, const int32_t* x, int64_t ldx, const float* state_in, float* state_out,
                            const float* params, void* ws, cudaStream_t st, long long drop_call = -1,
                            const int32_t* y = nullptr, int* fused_blocks = nullptr) {
    const crnn_config& c = h->cfg;
    const bool drop = h->drop_on && drop_call >= 0;
    const int H = c.rnn_size, V = c.vocab_size, L = c.num_layers, S = c.seq_length, B = c.batch_size;
    const int64_t BH = (int64_t)B * H;
    const int64_t SB = (int64_t)S * B;
    const int gw = h->GS * H;
    const int sc = h->sm_count;
    const int C = h->wf_C;
    const bool wf = h->wf_W > 1;
    prof_mark(h, st, CRNN_PHASE_MISC);
    h->dbg_ptr = reinterpret_cast<long long*>((char*)ws + 24576);      // 8 KB before the barrier blocks
    int rc = refresh_bf16_params(h, params, ws, st);
    if (rc != CRNN_OK) return rc;
    const bf16_t* pbf = wsb(ws, h->ws.pbf);
    float* x0 = wsf(ws, h->ws.x0);
    bf16_t* x0b = wsb(ws, h->ws.x0b);
    prof_mark(h, st, CRNN_PHASE_EMBED_GATHER);
    const LoOfs lo = lo_offsets(h);
    embed_gather_kernel<<<grid_for(SB * H / 4 + 1, sc), 256, 0, st>>>(params + h->emb_off, V, H, x, ldx, B, S, x0, x0b, lo.act);
    CRNN_LAUNCH_CHECK("embed_gather_kernel");
    prof_mark(h, st, CRNN_PHASE_MISC);
    CRNN_CUDA(cudaMemsetAsync(ws_barrier(ws, 0), 0, 512 * 256, st));
    // In wavefront mode the "rec_fwd" phase is the whole fork..join span: the recurrences of all layers
    // plus their chunked x-projections, which run interleaved with them.
    // Layer 0's x-projection depends only on the embeddings: with CRNN_WF_HOIST0=1 it runs as ONE full-GPU GEMM here (and
    // layer 0's input-gradient GEMM as one after the BPTT join), so the SMs left beside the recurrence kernels serve only
    // the chunk GEMMs on the layers' critical path.  Measured on config 3: the recurrence spans shrink (2.14 -> 2.09 and
    // 2.65 -> 2.58 ms, split precision) but the two GEMMs are then serial, 5.03 -> 5.19 ms per step: off by default.
    const bool hoist0 = wf && h->env.hoist0;
    if (hoist0) {
        LayerParams lp = layer_params(h, params, 0);
        const bf16_t* inb0 = x0b;
        if (drop) {
            rc = drop_apply(h, drop_call, 0, x0, nullptr, wsb(ws, h->ws.xinb[0]), 0, SB * H, st, lo.act);
            if (rc != CRNN_OK) return rc;
            inb0 = wsb(ws, h->ws.xinb[0]);
        }
        prof_mark(h, st, CRNN_PHASE_XPROJ_GEMM);
        float* g0 = wsf(ws, h->ws.gates[0]);
        if (c.model == CRNN_MODEL_GRU) {
            rc = gemm_tc(h, (int)SB, 2 * H, H, inb0, H, 0, lo.act, pbf + lp.oW, 2 * H, 1, lo.pbf, g0, gw, lp.b, 0, st);
            if (rc == CRNN_OK) rc = gemm_tc(h, (int)SB, H, H, inb0, H, 0, lo.act, pbf + lp.oWc, H, 1, lo.pbf, g0 + 2 * H, gw, lp.bc, 0, st);
        } else {
            rc = gemm_tc(h, (int)SB, h->G * H, H, inb0, H, 0, lo.act, pbf + lp.oW, h->G * H, 1, lo.pbf, g0, gw, lp.b, 0, st);
        }
        if (rc != CRNN_OK) return rc;
        prof_mark(h, st, CRNN_PHASE_MISC);
    }
    if (wf) prof_mark(h, st, CRNN_PHASE_REC_FWD);
    rc = wf_fork(h, st, 0);
    if (rc != CRNN_OK) return rc;
    for (int l = 0; l < L; ++l) {
        h->dbg_layer = l;
        cudaStream_t s = wf_stream(h, st, l);
        LayerParams lp = layer_params(h, params, l);
        float* hx = wsf(ws, h->ws.hx[l]);
        float* cx = c.model == CRNN_MODEL_LSTM ? wsf(ws, h->ws.cx[l]) : nullptr;
        float* gates = wsf(ws, h->ws.gates[l]);
        bf16_t* hxb = wsb(ws, h->ws.hxb[l]);
        const bf16_t* inb = drop ? wsb(ws, h->ws.xinb[l]) : (l == 0 ? x0b : wsb(ws, h->ws.hxb[l - 1]) + BH);
        const int64_t in_lo = (drop || l == 0) ? lo.act : lo.hxb;
        const float* in32 = l == 0 ? x0 : wsf(ws, h->ws.hx[l - 1]) + BH;      // dropout masks the fp32 source
        const float* sin = state_in + (int64_t)l * h->nstate * BH;
        const float* hin = c.model == CRNN_MODEL_LSTM ? sin + BH : sin;
        if (c.model == CRNN_MODEL_LSTM) CRNN_CUDA(cudaMemcpyAsync(cx, sin, BH * 4, cudaMemcpyDeviceToDevice, s));
        CRNN_CUDA(cudaMemcpyAsync(hx, hin, BH * 4, cudaMemcpyDeviceToDevice, s));
        tc::cast_bf16_kernel<<<grid_for(BH, sc), 256, 0, s>>>(hin, H, B, H, hxb, H, lo.hxb);
        CRNN_LAUNCH_CHECK("cast_bf16_kernel");
        cudaStream_t gs = wf_gstream(h, st, l);
        if (h->tcp.gather) {
            // direct exchange: the slots of the FIRST chunk start as the sentinel (later chunks are filled by the launch before them)
            const int64_t n0 = (int64_t)wf_t0(h, 1) * BH;
            for (int half = 0; half <= h->split; ++half) {
                CRNN_CUDA(cudaMemsetAsync(hxb + BH + half * lo.hxb, 0xFF, n0 * sizeof(bf16_t), s));
                if (c.model == CRNN_MODEL_GRU) CRNN_CUDA(cudaMemsetAsync(wsb(ws, h->ws.rsb[l]) + half * lo.act, 0xFF, n0 * sizeof(bf16_t), s));
            }
        }
        // (persist & 4: forward pass, layer 0 only - the first layer's launches are the forward pass's critical chain)
        const bool persist = wf && ((h->persist & 1) || ((h->persist & 4) && l == 0));
        const bool persist_below = wf && l > 0 && ((h->persist & 1) || ((h->persist & 4) && l == 1));
        const unsigned nctas = (unsigned)(h->tcp.nslices * h->tcp.ntiles);
        for (int ck = 0; ck < C; ++ck) {
            const int t0 = wf_t0(h, ck), t1 = wf_t0(h, ck + 1);
            const int rows = (t1 - t0) * B;
            if (wf && l > 0) {                       // layer l-1 has produced this chunk's inputs
                rc = persist_below ? wf_wait_value(gs, ws_pflag(ws, 0, l - 1) + 32 + ck, nctas) : wf_wait(h, wf_id(h, 0, l - 1, ck), gs);
                if (rc != CRNN_OK) return rc;
            }
            const bool hoisted = hoist0 && l == 0;   // done above for the whole sequence
            if (drop && !hoisted) {                  // masked bf16 copy of this chunk's inputs
                rc = drop_apply(h, drop_call, l, in32 + (int64_t)t0 * BH, nullptr, wsb(ws, h->ws.xinb[l]) + (int64_t)t0 * BH,
                                (int64_t)t0 * BH, (int64_t)rows * H, gs, lo.act);
                if (rc != CRNN_OK) return rc;
            }
            if (!wf) prof_mark(h, st, CRNN_PHASE_XPROJ_GEMM);
            // x half of [x,h]*W for the chunk's steps: A = inputs (K-major), B = kernel rows 0..H stored [K,N]
            const bf16_t* ain = inb + (int64_t)t0 * BH;
            float* gch = gates + (int64_t)t0 * B * gw;
            if (hoisted) {
                rc = CRNN_OK;
            } else if (c.model == CRNN_MODEL_GRU) {
                rc = gemm_tc(h, rows, 2 * H, H, ain, H, 0, in_lo, pbf + lp.oW, 2 * H, 1, lo.pbf, gch, gw, lp.b, 0, gs);
                if (rc == CRNN_OK) rc = gemm_tc(h, rows, H, H, ain, H, 0, in_lo, pbf + lp.oWc, H, 1, lo.pbf, gch + 2 * H, gw, lp.bc, 0, gs);
            } else {
                const int GH = h->G * H;
                rc = gemm_tc(h, rows, GH, H, ain, H, 0, in_lo, pbf + lp.oW, GH, 1, lo.pbf, gch, gw, lp.b, 0, gs);
            }
            if (rc != CRNN_OK) return rc;
            if (persist) {                           // the kernel (launched below, once) reads the flag at the chunk's first step
                if (!hoisted) { rc = wf_write_value(gs, ws_pflag(ws, 0, l), (unsigned)(ck + 1)); if (rc != CRNN_OK) return rc; }
                if (ck == 0) {                       // the launch itself waits for the first chunk: a kernel that only spins would keep its SMs from the chunk GEMMs
                    rc = wf_record(h, wf_id(h, 1, l, 0), gs);
                    if (rc == CRNN_OK) rc = wf_wait(h, wf_id(h, 1, l, 0), s);
                    // ... and for ALL chunk GEMMs of the layer below (they have no dependencies and finish early): once this
                    // kernel is resident they would be left with the few spare SMs and stall the layer below (CRNN_WF_WAIT_BELOW=0: off; config 3 forward span 2.08 -> 1.81 ms)
                    if (rc == CRNN_OK && l > 0 && env_int("CRNN_WF_WAIT_BELOW", 1)) rc = wf_wait(h, wf_id(h, 1, l - 1, C - 1), s);
                    if (rc != CRNN_OK) return rc;
                }
                if (ck == C - 1) { rc = wf_record(h, wf_id(h, 1, l, C - 1), gs); if (rc != CRNN_OK) return rc; }
                if (ck + 1 < C) continue;
            } else if (wf) {
                rc = wf_record(h, wf_id(h, 1, l, ck), gs);
                if (rc == CRNN_OK) rc = wf_wait(h, wf_id(h, 1, l, ck), s);
                if (rc != CRNN_OK) return rc;
            }
            tc::RecTcArgs a{};
            a.B = B; a.H = H; a.S = S; a.GS = h->GS; a.t_begin = persist ? 0 : t0; a.t_end = persist ? S : t1;
            a.gates = gates; a.hx = hx; a.cx = cx; a.hxb = hxb;
            a.rsb = c.model == CRNN_MODEL_GRU ? wsb(ws, h->ws.rsb[l]) : nullptr;
            a.fill_t0 = t1; a.fill_t1 = ck + 1 < C ? wf_t0(h, ck + 2) : t1;
            a.barrier = ws_barrier(ws, persist ? l * C : l * C + ck);
            if (persist) { a.pchunks = C; a.in_flag = hoisted ? nullptr : ws_pflag(ws, 0, l); a.out_flag = ws_pflag(ws, 0, l) + 32; }
            if (!wf) prof_mark(h, st, CRNN_PHASE_REC_FWD);
            rc = launch_rec_tc_model(h, false, a, nullptr, nullptr, wsb(ws, h->ws.wr[l]), s);
            if (rc != CRNN_OK) return rc;
            if (!wf) prof_mark(h, st, CRNN_PHASE_MISC);
            if (wf) {
                rc = wf_record(h, wf_id(h, 0, l, ck), s);
                if (rc != CRNN_OK) return rc;
            }
        }
        float* sout = state_out + (int64_t)l * h->nstate * BH;
        if (c.model == CRNN_MODEL_LSTM) {
            CRNN_CUDA(cudaMemcpyAsync(sout, cx + (int64_t)S * BH, BH * 4, cudaMemcpyDeviceToDevice, s));
            CRNN_CUDA(cudaMemcpyAsync(sout + BH, hx + (int64_t)S * BH, BH * 4, cudaMemcpyDeviceToDevice, s));
        } else {
            CRNN_CUDA(cudaMemcpyAsync(sout, hx + (int64_t)S * BH, BH * 4, cudaMemcpyDeviceToDevice, s));
        }
    }
    rc = wf_join(h, st, 1);                      // slots 1 .. 2W
    if (rc != CRNN_OK) return rc;
    const bf16_t* topb = wsb(ws, h->ws.hxb[L - 1]) + BH;
    if (drop) {
        rc = drop_apply(h, drop_call, L, wsf(ws, h->ws.hx[L - 1]) + BH, nullptr, wsb(ws, h->ws.topdb), 0, SB * H, st, lo.act);
        if (rc != CRNN_OK) return rc;
        topb = wsb(ws, h->ws.topdb);
    }
    prof_mark(h, st, CRNN_PHASE_LOGITS_GEMM);
    const int64_t top_lo = drop ? lo.act : lo.hxb;
    if (fused_blocks) *fused_blocks = 0;
    rc = CRNN_ERR_UNSUP;
    if (h->env.fuse_proj && ceil_div64(SB, 128) <= 2048) {
        // fused projection + softmax + cross-entropy (training) / projection with fp32 logits out (forward only)
        int nblk = 0;
        rc = tc::proj_xent_tc((int)SB, V, h->Vp, H, B, S, topb, h->split ? topb + top_lo : nullptr, wsb(ws, h->ws.wsmb),
                              h->split ? wsb(ws, h->ws.wsmb) + lo.wsm : nullptr, params + h->tensors[1].offset, y, ldx,
                              y ? wsb(ws, h->ws.dlb) : nullptr, y ? lo.dlb : 0, y ? ws_loss_part(ws) : nullptr, y ? wsf(ws, h->ws.csum[L]) : nullptr,
                              y ? nullptr : wsf(ws, h->ws.logits), st, &nblk);
        if (rc == CRNN_OK && fused_blocks) *fused_blocks = nblk;
    }
    if (rc == CRNN_ERR_UNSUP)
        rc = gemm_tc(h, (int)SB, V, H, topb, H, 0, top_lo, wsb(ws, h->ws.wsmb), h->Vp, 1, lo.wsm, wsf(ws, h->ws.logits), V,
                     params + h->tensors[1].offset, 0, st);
    prof_mark(h, st, CRNN_PHASE_MISC);
    return rc;
}

static int forward_backward_bf16(crnn_handle* h, const int32_t* x, const int32_t* y, int64_t ld, const float* state_in,
                                 float* state_out, const float* params, float* grads, void* ws, cudaStream_t st, long long call) {
    const crnn_config& c = h->cfg;
    const bool drop = h->drop_on;
    const int H = c.rnn_size, V = c.vocab_size, L = c.num_layers, S = c.seq_length, B = c.batch_size;
    const int64_t BH = (int64_t)B * H, SB = (int64_t)S * B;
    const int gw = h->GS * H, zw = h->GZ * H, Vp = h->Vp;
    const int sc = h->sm_count;
    int C = h->wf_C;                             // chunks of the forward wavefront; the BPTT count from the fork below on
    const bool wf = h->wf_W > 1;
    prof_mark(h, st, CRNN_PHASE_MISC);
    CRNN_CUDA(cudaMemsetAsync(grads, 0, (h->P + CRNN_GRAD_TAIL) * sizeof(float), st));
    CRNN_CUDA(cudaMemsetAsync(ws, 0, kScratchBytes, st));
    int fused = 0;
    // the scatter's token order depends on the ids only: built on a side stream while the forward pass runs
    const int ev_sort = 4 * L * C + 128;
    int rc = CRNN_OK;
    if (h->ws.es_sorted) {
        rc = wf_record(h, ev_sort, st);
        if (rc == CRNN_OK) rc = wf_wait(h, ev_sort, h->aux_stream);
        if (rc == CRNN_OK) rc = scatter_sort(h, x, ld, ws, h->aux_stream);
        if (rc == CRNN_OK) rc = wf_record(h, ev_sort + 1, h->aux_stream);
        if (rc != CRNN_OK) return rc;
    }
    rc = run_forward_bf16(h, x, ld, state_in, state_out, params, ws, st, call, y, &fused);
    if (rc != CRNN_OK) return rc;
    const bf16_t* pbf = wsb(ws, h->ws.pbf);
    float* logits = wsf(ws, h->ws.logits);
    bf16_t* dlb = wsb(ws, h->ws.dlb);
    prof_mark(h, st, CRNN_PHASE_SOFTMAX_XENT);
    const LoOfs lo = lo_offsets(h);
    const int64_t top_lo = drop ? lo.act : lo.hxb;
    int nloss = fused;
    if (!fused) {
        nloss = grid_for(SB * 32, sc);
        softmax_xent_kernel<<<nloss, 256, 0, st>>>(logits, B, S, V, y, ld, ws_loss(ws), dlb, Vp, lo.dlb, ws_loss_part(ws));
        CRNN_LAUNCH_CHECK("softmax_xent_kernel");
    }
    const bf16_t* topb = drop ? wsb(ws, h->ws.topdb) : wsb(ws, h->ws.hxb[L - 1]) + BH;
    prof_mark(h, st, CRNN_PHASE_PROJ_GRADS);
    // d softmax_w [H,V] = out^T * dlogits ; d softmax_b = colsum ; d out = dlogits * softmax_w^T
    rc = gemm_tc(h, H, V, (int)SB, topb, H, 1, top_lo, dlb, Vp, 1, lo.dlb, grads + h->tensors[0].offset, V, nullptr, 1, st);
    if (rc == CRNN_OK && fused) {                    // the fused kernel left one row of column sums per CTA: add them in CTA order
        colsum_finish_kernel<<<ceil_div(V, 256), 256, 0, st>>>(wsf(ws, h->ws.csum[L]), fused, V, grads + h->tensors[1].offset);
        CRNN_LAUNCH_CHECK("colsum_finish_kernel");
    } else if (rc == CRNN_OK) rc = colsum_f32(logits, V, (int)SB, V, grads + h->tensors[1].offset, st, wsf(ws, h->ws.csum[L]));
    if (rc == CRNN_OK) rc = gemm_tc(h, (int)SB, H, V, dlb, Vp, 0, lo.dlb, wsb(ws, h->ws.wsmb), Vp, 0, lo.wsm, wsf(ws, h->ws.gout[L - 1]), H, nullptr, 0, st);
    if (rc == CRNN_OK && drop) rc = drop_apply(h, call, L, wsf(ws, h->ws.gout[L - 1]), wsf(ws, h->ws.gout[L - 1]), nullptr, 0, SB * H, st);
    if (rc != CRNN_OK) return rc;

    // BPTT, top layer first.  Wavefront: layer l works on chunk c (late steps first) as soon as layer
    // l+1 has turned that chunk's dz into input gradients; "rec_bwd" then times the whole fork..join
    // span (recurrences + interleaved input-gradient GEMMs), weight gradients included at its tail.
    const bool hoist0 = wf && h->env.hoist0;     // see run_forward_bf16
    // BPTT may use its own chunk count: the helpers (wf_t0, wf_id, fork / join slots) read h->wf_C, restored on every exit
    struct ChunkGuard { crnn_handle* h; int saved; ~ChunkGuard() { h->wf_C = saved; } } chunk_guard{h, h->wf_C};
    const int Cf = h->wf_C;                      // the forward count: its barrier blocks 0 .. L*Cf-1 are used up
    C = h->wf_Cb;
    h->wf_C = C;
    if (wf) prof_mark(h, st, CRNN_PHASE_REC_BWD);
    rc = wf_fork(h, st, 1 + 2 * h->wf_W);
    if (rc != CRNN_OK) return rc;
    for (int l = L - 1; l >= 0; --l) {
        cudaStream_t s = wf_stream(h, st, l);
        LayerParams lp = layer_params(h, params, l);
        float* hx = wsf(ws, h->ws.hx[l]);
        float* cx = c.model == CRNN_MODEL_LSTM ? wsf(ws, h->ws.cx[l]) : nullptr;
        float* gates = wsf(ws, h->ws.gates[l]);
        const bf16_t* hxb = wsb(ws, h->ws.hxb[l]);
        const bf16_t* inb = drop ? wsb(ws, h->ws.xinb[l]) : (l == 0 ? wsb(ws, h->ws.x0b) : wsb(ws, h->ws.hxb[l - 1]) + BH);
        const int64_t in_lo = (drop || l == 0) ? lo.act : lo.hxb;
        float* dz = wsf(ws, h->ws.dzv[l]);
        bf16_t* dzb = wsb(ws, h->ws.dzbv[l]);
        float* dnext = l > 0 ? wsf(ws, h->ws.gout[l - 1]) : wsf(ws, h->ws.gx0);
        cudaStream_t gs = wf_gstream(h, st, l);
        if (h->tcp.gather) {                         // sentinel over the LAST chunk's dz slots (see run_forward_bf16)
            const int64_t r0 = (int64_t)wf_t0(h, C - 1) * B * zw, n0 = (int64_t)(S - wf_t0(h, C - 1)) * B * zw;
            for (int half = 0; half <= h->split; ++half)
                CRNN_CUDA(cudaMemsetAsync(dzb + r0 + half * lo.dzb, 0xFF, n0 * sizeof(bf16_t), s));
        }
        const bool persist = wf && (h->persist & 2);
        const unsigned nctas = (unsigned)(h->tcp.nslices * h->tcp.ntiles);
        for (int ck = C - 1; ck >= 0; --ck) {
            const int t0 = wf_t0(h, ck), t1 = wf_t0(h, ck + 1);
            const int rows = (t1 - t0) * B;
            if (wf && (!persist || ck == C - 1) && l < L - 1) {   // layer l+1 has produced this chunk's output gradients
                rc = wf_wait(h, wf_id(h, 3, l + 1, ck), s);         // (persistent form: the launch waits for the first chunk only)
                if (rc != CRNN_OK) return rc;
            }
            if (!persist || ck == C - 1) {           // persistent form: ONE launch for the whole sequence, before the chunk GEMMs that follow it
                tc::RecTcArgs a{};
                a.B = B; a.H = H; a.S = S; a.GS = h->GS; a.t_begin = persist ? 0 : t0; a.t_end = persist ? S : t1;
                a.gates = gates; a.hx = hx; a.cx = cx;
                a.dy = wsf(ws, h->ws.gout[l]); a.dz = dz; a.dzb = dzb;
                a.carry0 = wsf(ws, h->ws.carry0v[l]); a.carry1 = wsf(ws, h->ws.carry1v[l]);
                a.fill_t0 = ck > 0 ? wf_t0(h, ck - 1) : t0; a.fill_t1 = t0;
                a.barrier = ws_barrier(ws, persist ? L * Cf + l * C : L * Cf + l * C + ck);
                if (persist) {   // dy of chunk k (processing order) is there when the layer above's GEMM stream has written k + 1
                    a.pchunks = C; a.in_flag = l < L - 1 ? ws_pflag(ws, 1, l) : nullptr; a.out_flag = ws_pflag(ws, 1, l) + 32;
                }
                if (!wf) prof_mark(h, st, CRNN_PHASE_REC_BWD);
                rc = launch_rec_tc_model(h, true, a, pbf + lp.oW, c.model == CRNN_MODEL_GRU ? pbf + lp.oWc : nullptr, nullptr, s);
                if (rc != CRNN_OK) return rc;
            }
            if (persist) {                           // the chunk's dz rows: every CTA of the kernel has finished the chunk
                rc = wf_wait_value(gs, ws_pflag(ws, 1, l) + 32 + (C - 1 - ck), nctas);      // (counters are indexed in processing order)
                if (rc != CRNN_OK) return rc;
            } else if (wf) {
                rc = wf_record(h, wf_id(h, 2, l, ck), s);
                if (rc == CRNN_OK) rc = wf_wait(h, wf_id(h, 2, l, ck), gs);
                if (rc != CRNN_OK) return rc;
            }
            // input gradients of the chunk: dIn = dz * Wx^T (A = dz K-major, B = kernel rows 0..H as [N,K])
            if (!wf) prof_mark(h, st, CRNN_PHASE_INPUT_GRADS);
            const bf16_t* dzc = dzb + (int64_t)t0 * B * zw;
            float* dn = dnext + (int64_t)t0 * BH;
            if (hoist0 && l == 0) {
                continue;                            // nothing waits for the embedding-side gradients: one GEMM after the join
            } else if (c.model == CRNN_MODEL_GRU) {
                rc = gemm_tc(h, rows, H, 2 * H, dzc, zw, 0, lo.dzb, pbf + lp.oW, 2 * H, 0, lo.pbf, dn, H, nullptr, 0, gs);
                if (rc == CRNN_OK) rc = gemm_tc(h, rows, H, H, dzc + 2 * H, zw, 0, lo.dzb, pbf + lp.oWc, H, 0, lo.pbf, dn, H, nullptr, 1, gs);
            } else {
                const int GH = h->G * H;
                rc = gemm_tc(h, rows, H, GH, dzc, zw, 0, lo.dzb, pbf + lp.oW, GH, 0, lo.pbf, dn, H, nullptr, 0, gs);
            }
            if (rc == CRNN_OK && drop) rc = drop_apply(h, call, l, dn, dn, nullptr, (int64_t)t0 * BH, (int64_t)rows * H, gs);
            if (rc != CRNN_OK) return rc;
            if (persist && l > 0) {                  // the layer below reads its dy chunk once this is written
                rc = wf_write_value(gs, ws_pflag(ws, 1, l - 1), (unsigned)(C - ck));
                if (rc != CRNN_OK) return rc;
            }
            if (wf) {
                rc = wf_record(h, wf_id(h, 3, l, ck), gs);
                if (rc != CRNN_OK) return rc;
            }
        }
        s = gs;                                      // weight gradients follow the layer's last input-gradient GEMM
        if (!wf) prof_mark(h, st, CRNN_PHASE_WEIGHT_GRADS);
        float* gW = grads + lp.oW;
        float* gb = grads + lp.ob;
        if (c.model == CRNN_MODEL_GRU) {
            float* gWc = grads + lp.oWc;
            float* gbc = grads + lp.obc;
            const bf16_t* rsb = wsb(ws, h->ws.rsb[l]);
            rc = gemm_tc(h, H, 2 * H, (int)SB, inb, H, 1, in_lo, dzb, zw, 1, lo.dzb, gW, 2 * H, nullptr, 1, s);
            if (rc == CRNN_OK) rc = gemm_tc(h, H, 2 * H, (int)SB, hxb, H, 1, lo.hxb, dzb, zw, 1, lo.dzb, gW + (int64_t)H * 2 * H, 2 * H, nullptr, 1, s);
            if (rc == CRNN_OK) rc = colsum_f32(dz, zw, (int)SB, 2 * H, gb, s, wsf(ws, h->ws.csum[l]));
            if (rc == CRNN_OK) rc = gemm_tc(h, H, H, (int)SB, inb, H, 1, in_lo, dzb + 2 * H, zw, 1, lo.dzb, gWc, H, nullptr, 1, s);
            if (rc == CRNN_OK) rc = gemm_tc(h, H, H, (int)SB, rsb, H, 1, lo.act, dzb + 2 * H, zw, 1, lo.dzb, gWc + (int64_t)H * H, H, nullptr, 1, s);
            if (rc == CRNN_OK) rc = colsum_f32(dz + 2 * H, zw, (int)SB, H, gbc, s, wsf(ws, h->ws.csum[l]));
        } else {
            const int GH = h->G * H;
            rc = gemm_tc(h, H, GH, (int)SB, inb, H, 1, in_lo, dzb, zw, 1, lo.dzb, gW, GH, nullptr, 1, s);
            if (rc == CRNN_OK) rc = gemm_tc(h, H, GH, (int)SB, hxb, H, 1, lo.hxb, dzb, zw, 1, lo.dzb, gW + (int64_t)H * GH, GH, nullptr, 1, s);
            if (rc == CRNN_OK) rc = colsum_f32(dz, zw, (int)SB, GH, gb, s, wsf(ws, h->ws.csum[l]));
        }
        if (rc != CRNN_OK) return rc;
        if (h->bucket_event && l >= h->bucket_layer) {
            // crnn_set_bucket_event: cells bucket_layer .. L-1 are final once each of their streams has got this far
            const int slot = 4 * L * C + 64;
            if (l > h->bucket_layer) {
                rc = wf_record(h, slot + l, s);
                if (rc != CRNN_OK) return rc;
            } else {
                for (int lu = l + 1; lu < L; ++lu) {
                    rc = wf_wait(h, slot + lu, s);
                    if (rc != CRNN_OK) return rc;
                }
                CRNN_CUDA(cudaEventRecord(h->bucket_event, s));
            }
        }
    }
    // The embedding scatter needs layer 0's input gradients only, which are complete a whole set of weight-gradient GEMMs
    // before the join: it runs on the side stream beside those GEMMs (HBM-bound next to tensor-bound work).
    const bool early_scatter = wf && !hoist0 && h->ws.es_sorted;
    if (early_scatter) {
        rc = wf_wait(h, wf_id(h, 3, 0, 0), h->aux_stream);           // layer 0's last input-gradient chunk (chunk 0)
        if (rc == CRNN_OK) rc = scatter_and_finalize(h, x, ld, wsf(ws, h->ws.gx0), grads, ws, -1, h->aux_stream);
        if (rc == CRNN_OK) rc = wf_record(h, ev_sort + 1, h->aux_stream);
        if (rc != CRNN_OK) return rc;
    }
    rc = wf_join(h, st, 2 + 2 * h->wf_W);
    if (rc != CRNN_OK) return rc;
    if (hoist0) {                                    // layer 0's input gradients for the whole sequence
        prof_mark(h, st, CRNN_PHASE_INPUT_GRADS);
        LayerParams lp = layer_params(h, params, 0);
        const bf16_t* dzb0 = wsb(ws, h->ws.dzbv[0]);
        float* dn = wsf(ws, h->ws.gx0);
        if (c.model == CRNN_MODEL_GRU) {
            rc = gemm_tc(h, (int)SB, H, 2 * H, dzb0, zw, 0, lo.dzb, pbf + lp.oW, 2 * H, 0, lo.pbf, dn, H, nullptr, 0, st);
            if (rc == CRNN_OK) rc = gemm_tc(h, (int)SB, H, H, dzb0 + 2 * H, zw, 0, lo.dzb, pbf + lp.oWc, H, 0, lo.pbf, dn, H, nullptr, 1, st);
        } else {
            rc = gemm_tc(h, (int)SB, H, h->G * H, dzb0, zw, 0, lo.dzb, pbf + lp.oW, h->G * H, 0, lo.pbf, dn, H, nullptr, 0, st);
        }
        if (rc == CRNN_OK && drop) rc = drop_apply(h, call, 0, dn, dn, nullptr, 0, SB * H, st);
        if (rc != CRNN_OK) return rc;
    }
    if (h->ws.es_sorted) { rc = wf_wait(h, ev_sort + 1, st); if (rc != CRNN_OK) return rc; }   // the token order, or the whole early scatter
    if (early_scatter) {
        prof_mark(h, st, CRNN_PHASE_EMBED_SCATTER);
        finalize_scalars_kernel<<<1, 256, 0, st>>>(ws_sq(ws), ws_loss(ws), (double)SB, grads + h->P, ws_loss_part(ws), nloss);
        CRNN_LAUNCH_CHECK("finalize_scalars_kernel");
    } else {
        rc = scatter_and_finalize(h, x, ld, wsf(ws, h->ws.gx0), grads, ws, nloss, st);
        if (rc != CRNN_OK) return rc;
    }
    prof_mark(h, st, -1);
    return CRNN_OK;
}

static int check_common(crnn_handle* h, const void* a, const void* b, const void* c) {
    if (!h) return set_err(CRNN_ERR_ARG, "null handle");
    if (!a || !b || !c) return set_err(CRNN_ERR_ARG, "null device pointer");
    CRNN_CUDA(cudaSetDevice(h->cfg.device));
    if (h->cfg.batch_size > 2 * 1024 * 1024) return set_err(CRNN_ERR_ARG, "batch too large");
    if (h->cfg.num_layers > 32) return set_err(CRNN_ERR_ARG, "at most 32 layers");
    if (h->err_word && *reinterpret_cast<volatile unsigned*>(h->err_word))
        return set_err(CRNN_ERR_CUDA, "a recurrence kernel of an earlier call gave up waiting for a peer slice's rows (results of that call are invalid)");
    return CRNN_OK;
}

extern "C" {

int crnn_forward_backward(crnn_handle* h, const int32_t* x, const int32_t* y, int64_t ld,
                          const float* state_in, float* state_out, const float* params, float* grads,
                          void* ws, void* stream) {
    int rc = check_common(h, x, params, ws);
    if (rc != CRNN_OK) return rc;
    if (!y || !state_in || !state_out || !grads) return set_err(CRNN_ERR_ARG, "null device pointer");
    const crnn_config& c = h->cfg;
    if (ld < c.seq_length) return set_err(CRNN_ERR_ARG, "ld_xy %lld < seq_length %d", (long long)ld, c.seq_length);
    cudaStream_t st = (cudaStream_t)stream;
    const long long call = h->fb_calls++;               // index of this training call (dropout masks)
    const bool drop = h->drop_on;
    if (h->tc) return forward_backward_bf16(h, x, y, ld, state_in, state_out, params, grads, ws, st, call);
    const int H = c.rnn_size, V = c.vocab_size, L = c.num_layers, S = c.seq_length, B = c.batch_size;
    const int64_t BH = (int64_t)B * H, SB = (int64_t)S * B;
    const int gw = h->GS * H, zw = h->GZ * H;

    prof_mark(h, st, CRNN_PHASE_MISC);
    CRNN_CUDA(cudaMemsetAsync(grads, 0, (h->P + CRNN_GRAD_TAIL) * sizeof(float), st));
    CRNN_CUDA(cudaMemsetAsync(ws, 0, kScratchBytes, st));
    rc = run_forward(h, x, ld, state_in, state_out, params, ws, st, call);
    if (rc != CRNN_OK) return rc;
    prof_mark(h, st, CRNN_PHASE_SOFTMAX_XENT);

    // loss + dlogits in place, then the three projection-layer gradients
    float* logits = wsf(ws, h->ws.logits);
    const int nloss = grid_for(SB * 32, h->sm_count);
    softmax_xent_kernel<<<nloss, 256, 0, st>>>(logits, B, S, V, y, ld, ws_loss(ws), nullptr, 0, 0, ws_loss_part(ws));
    CRNN_LAUNCH_CHECK("softmax_xent_kernel");
    const float* top = drop ? wsf(ws, h->ws.topd) : wsf(ws, h->ws.hx[L - 1]) + BH;
    const float* Wsm = params + h->tensors[0].offset;
    float* dya = wsf(ws, h->ws.dya);
    float* dyb = wsf(ws, h->ws.dyb);
    prof_mark(h, st, CRNN_PHASE_PROJ_GRADS);
    rc = gemm_f32(1, 0, H, V, (int)SB, top, H, logits, V, grads + h->tensors[0].offset, V, nullptr, 1, st, h->sm_count);
    if (rc == CRNN_OK) rc = colsum_f32(logits, V, (int)SB, V, grads + h->tensors[1].offset, st, wsf(ws, h->ws.csum[L]));
    if (rc == CRNN_OK) rc = gemm_f32(0, 1, (int)SB, H, V, logits, V, Wsm, V, dya, H, nullptr, 0, st, h->sm_count);
    if (rc == CRNN_OK && drop) rc = drop_apply(h, call, L, dya, dya, nullptr, 0, SB * H, st);   // back through the output mask
    if (rc != CRNN_OK) return rc;

    float* dz = wsf(ws, h->ws.dz);
    float* dcur = dya;
    float* dnext = dyb;
    for (int l = L - 1; l >= 0; --l) {
        LayerParams lp = layer_params(h, params, l);
        float* hx = wsf(ws, h->ws.hx[l]);
        float* cx = c.model == CRNN_MODEL_LSTM ? wsf(ws, h->ws.cx[l]) : nullptr;
        float* gates = wsf(ws, h->ws.gates[l]);
        const float* in = drop ? wsf(ws, h->ws.xin[l]) : (l == 0 ? wsf(ws, h->ws.x0) : wsf(ws, h->ws.hx[l - 1]) + BH);
        RecArgs a{};
        a.B = B; a.H = H; a.S = S; a.GS = h->GS;
        a.W = lp.W; a.Wc = lp.Wc; a.gates = gates; a.hx = hx; a.cx = cx;
        a.dy = dcur; a.dz = dz; a.carry0 = wsf(ws, h->ws.carry0); a.carry1 = wsf(ws, h->ws.carry1);
        a.barrier = ws_barrier(ws, 2 * l + 1);
        prof_mark(h, st, CRNN_PHASE_REC_BWD);
        rc = launch_rec_model(h, true, a, st);
        if (rc != CRNN_OK) return rc;
        prof_mark(h, st, CRNN_PHASE_WEIGHT_GRADS);
        float* gW = grads + lp.oW;
        float* gb = grads + lp.ob;
        const int sc = h->sm_count;
        if (c.model == CRNN_MODEL_GRU) {
            float* gWc = grads + lp.oWc;
            float* gbc = grads + lp.obc;
            // gates kernel [2H,2H]: rows 0..H from inputs, rows H..2H from h_{t-1}
            rc = gemm_f32(1, 0, H, 2 * H, (int)SB, in, H, dz, zw, gW, 2 * H, nullptr, 1, st, sc);
            if (rc == CRNN_OK) rc = gemm_f32(1, 0, H, 2 * H, (int)SB, hx, H, dz, zw, gW + (int64_t)H * 2 * H, 2 * H, nullptr, 1, st, sc);
            if (rc == CRNN_OK) rc = colsum_f32(dz, zw, (int)SB, 2 * H, gb, st, wsf(ws, h->ws.csum[l]));
            // candidate kernel [2H,H]: rows H..2H see r*h_{t-1}
            if (rc == CRNN_OK) rc = gemm_f32(1, 0, H, H, (int)SB, in, H, dz + 2 * H, zw, gWc, H, nullptr, 1, st, sc);
            if (rc == CRNN_OK) rc = gemm_f32(1, 0, H, H, (int)SB, gates + 3 * H, gw, dz + 2 * H, zw, gWc + (int64_t)H * H, H, nullptr, 1, st, sc);
            if (rc == CRNN_OK) rc = colsum_f32(dz + 2 * H, zw, (int)SB, H, gbc, st, wsf(ws, h->ws.csum[l]));
            prof_mark(h, st, CRNN_PHASE_INPUT_GRADS);
            if (rc == CRNN_OK) rc = gemm_f32(0, 1, (int)SB, H, 2 * H, dz, zw, lp.W, 2 * H, dnext, H, nullptr, 0, st, sc);
            if (rc == CRNN_OK) rc = gemm_f32(0, 1, (int)SB, H, H, dz + 2 * H, zw, lp.Wc, H, dnext, H, nullptr, 1, st, sc);
        } else {
            const int GH = h->G * H;
            rc = gemm_f32(1, 0, H, GH, (int)SB, in, H, dz, zw, gW, GH, nullptr, 1, st, sc);
            if (rc == CRNN_OK) rc = gemm_f32(1, 0, H, GH, (int)SB, hx, H, dz, zw, gW + (int64_t)H * GH, GH, nullptr, 1, st, sc);
            if (rc == CRNN_OK) rc = colsum_f32(dz, zw, (int)SB, GH, gb, st, wsf(ws, h->ws.csum[l]));
            prof_mark(h, st, CRNN_PHASE_INPUT_GRADS);
            if (rc == CRNN_OK) rc = gemm_f32(0, 1, (int)SB, H, GH, dz, zw, lp.W, GH, dnext, H, nullptr, 0, st, sc);
        }
        if (rc == CRNN_OK && drop) rc = drop_apply(h, call, l, dnext, dnext, nullptr, 0, SB * H, st);   // back through this cell's input masks
        if (rc != CRNN_OK) return rc;
        if (h->bucket_event && l == h->bucket_layer) CRNN_CUDA(cudaEventRecord(h->bucket_event, st));   // cells l .. L-1 are final
        std::swap(dcur, dnext);
    }
    // dcur = dX [S,B,H] = the IndexedSlices values of the embedding gradient
    rc = scatter_sort(h, x, ld, ws, st);
    if (rc == CRNN_OK) rc = scatter_and_finalize(h, x, ld, dcur, grads, ws, nloss, st);
    if (rc != CRNN_OK) return rc;
    prof_mark(h, st, -1);
    return CRNN_OK;
}

int crnn_set_bucket_event(crnn_handle* h, int32_t first_layer, void* event) {
    if (!h) return set_err(CRNN_ERR_ARG, "null handle");
    if (event && (first_layer < 0 || first_layer >= h->cfg.num_layers)) return set_err(CRNN_ERR_ARG, "first_layer must be in [0, num_layers)");
    h->bucket_event = (cudaEvent_t)event;
    h->bucket_layer = first_layer;
    return CRNN_OK;
}

int crnn_clip_adam(crnn_handle* h, float* params, const float* grads, float* m, float* v, float lr, float clip,
                   int64_t step, float mult, float* gnorm_out, void* ws, void* stream) {
    int rc = check_common(h, params, grads, ws);
    if (rc != CRNN_OK) return rc;
    if (!m || !v) return set_err(CRNN_ERR_ARG, "null device pointer");
    if (step < 1) return set_err(CRNN_ERR_ARG, "Adam step must be >= 1");
    cudaStream_t st = (cudaStream_t)stream;
    // beta powers as TF keeps them: fp32 variables multiplied once per step
    if (step < h->pow_step) { h->pow_step = 0; h->b1p = 1.0f; h->b2p = 1.0f; }
    while (h->pow_step < step) { h->b1p *= kBeta1; h->b2p *= kBeta2; ++h->pow_step; }
    const int64_t emb_end = h->emb_off + (int64_t)h->cfg.vocab_size * h->cfg.rnn_size;
    int nblk = (int)std::min<int64_t>(NORM_BLOCKS, std::max<int64_t>(1, h->P / 1024));
    prof_mark(h, st, CRNN_PHASE_CLIP_ADAM);
    sqnorm_partials_kernel<<<nblk, 256, 0, st>>>(grads, h->P, h->emb_off, emb_end, ws_partials(ws));
    CRNN_LAUNCH_CHECK("sqnorm_partials_kernel");
    AdamArgs a{};
    a.p = params; a.g = grads; a.m = m; a.v = v; a.p_bf16 = nullptr; a.P = h->P;
    a.partial = ws_partials(ws); a.npartial = nblk;
    a.lr_t = lr * sqrtf(1.0f - h->b2p) / (1.0f - h->b1p);
    a.beta1 = kBeta1; a.beta2 = kBeta2; a.eps = kEps; a.clip = clip; a.mult = mult; a.gnorm_out = gnorm_out;
    clip_adam_kernel<<<grid_for(h->P / 4, h->sm_count), 256, 0, st>>>(a);
    CRNN_LAUNCH_CHECK("clip_adam_kernel");
    prof_mark(h, st, -1);
    return CRNN_OK;
}

// ---- optimiser step for callers that cannot hand us a workspace -------------------------------
// The partial sums of the global norm normally live in the caller's workspace.  This variant keeps
// them in a fixed array that is part of the library image, so the call needs no buffer from the
// caller and performs no allocation of its own.  One partial per CRNN_NOALLOC_SPAN gradient floats.
#define CRNN_NOALLOC_MAX_PARTIALS 8192
#define CRNN_NOALLOC_SPAN 16384
static __device__ __align__(256) double g_noalloc_partials[CRNN_NOALLOC_MAX_PARTIALS];   // 64 KB on every GPU that loads the library
// In host code `g_noalloc_partials` is only a stand-in for the symbol, not a device address: the
// address each device gave the array is looked up once (below) and cached per device ordinal.
static double* g_noalloc_partials_addr[64] = {nullptr};

int crnn_clip_adam_noalloc(crnn_handle* h, float* params, const float* grads, float* m, float* v, float lr, float clip,
                           int64_t step, float mult, float* gnorm_out, void* stream) {
    if (!h) return set_err(CRNN_ERR_ARG, "null handle");
    if (!params || !grads || !m || !v) return set_err(CRNN_ERR_ARG, "null device pointer");
    if (step < 1) return set_err(CRNN_ERR_ARG, "Adam step must be >= 1");
    const int dev = h->cfg.device;
    if (dev < 0 || dev >= 64) return set_err(CRNN_ERR_ARG, "crnn_clip_adam_noalloc: device ordinal %d outside [0,64)", dev);
    const int64_t need = ceil_div64(h->P, CRNN_NOALLOC_SPAN);
    if (need > CRNN_NOALLOC_MAX_PARTIALS)
        return set_err(CRNN_ERR_ARG,
                       "crnn_clip_adam_noalloc: %lld parameter floats need %lld partial sums, the built-in scratch holds %d "
                       "(limit %lld floats); use crnn_clip_adam with a caller workspace",
                       (long long)h->P, (long long)need, CRNN_NOALLOC_MAX_PARTIALS,
                       (long long)CRNN_NOALLOC_MAX_PARTIALS * CRNN_NOALLOC_SPAN);
    CRNN_CUDA(cudaSetDevice(dev));
    if (!g_noalloc_partials_addr[dev]) {
        // first call on this device: resolve the array's device address (a driver query, not an allocation)
        double* addr = nullptr;
        CRNN_CUDA(cudaGetSymbolAddress((void**)&addr, g_noalloc_partials));
        if (!addr) return set_err(CRNN_ERR_CUDA, "cudaGetSymbolAddress returned a null address for the built-in scratch");
        g_noalloc_partials_addr[dev] = addr;
    }
    double* partials = g_noalloc_partials_addr[dev];
    cudaStream_t st = (cudaStream_t)stream;
    if (step < h->pow_step) { h->pow_step = 0; h->b1p = 1.0f; h->b2p = 1.0f; }
    while (h->pow_step < step) { h->b1p *= kBeta1; h->b2p *= kBeta2; ++h->pow_step; }
    const int64_t emb_end = h->emb_off + (int64_t)h->cfg.vocab_size * h->cfg.rnn_size;
    // Every partial the update kernel reads is rewritten here first, so nothing left in the array by
    // an earlier call (it is shared by all calls) needs clearing.
    sqnorm_partials_kernel<<<(int)need, 256, 0, st>>>(grads, h->P, h->emb_off, emb_end, partials);
    CRNN_LAUNCH_CHECK("sqnorm_partials_kernel");
    AdamArgs a{};
    a.p = params; a.g = grads; a.m = m; a.v = v; a.p_bf16 = nullptr; a.P = h->P;
    a.partial = partials; a.npartial = (int)need;
    a.lr_t = lr * sqrtf(1.0f - h->b2p) / (1.0f - h->b1p);
    a.beta1 = kBeta1; a.beta2 = kBeta2; a.eps = kEps; a.clip = clip; a.mult = mult; a.gnorm_out = gnorm_out;
    clip_adam_kernel<<<grid_for(h->P / 4, h->sm_count), 256, 0, st>>>(a);
    CRNN_LAUNCH_CHECK("clip_adam_kernel");
    return CRNN_OK;
}

int crnn_train_step(crnn_handle* h, const int32_t* x, const int32_t* y, int64_t ld, float* state, float* params,
                    float* grads, float* m, float* v, float lr, float clip, int64_t step, void* ws, void* stream) {
    int rc = crnn_forward_backward(h, x, y, ld, state, state, params, grads, ws, stream);
    if (rc != CRNN_OK) return rc;
    return crnn_clip_adam(h, params, grads, m, v, lr, clip, step, 1.0f, nullptr, ws, stream);
}

int crnn_forward(crnn_handle* h, const int32_t* x, int64_t ld, const float* state_in, float* state_out,
                 const float* params, float* logits_out, float* probs_out, void* ws, void* stream) {
    int rc = check_common(h, x, params, ws);
    if (rc != CRNN_OK) return rc;
    if (!state_in || !state_out) return set_err(CRNN_ERR_ARG, "null device pointer");
    const crnn_config& c = h->cfg;
    if (ld < c.seq_length) return set_err(CRNN_ERR_ARG, "ld_x %lld < seq_length %d", (long long)ld, c.seq_length);
    cudaStream_t st = (cudaStream_t)stream;
    CRNN_CUDA(cudaMemsetAsync(ws, 0, kScratchBytes, st));
    rc = h->tc ? run_forward_bf16(h, x, ld, state_in, state_out, params, ws, st)
                                       : run_forward(h, x, ld, state_in, state_out, params, ws, st);
    if (rc != CRNN_OK) return rc;
    if (logits_out || probs_out) {
        const int64_t SB = (int64_t)c.seq_length * c.batch_size;
        logits_out_kernel<<<grid_for(SB * 32, h->sm_count), 256, 0, st>>>(wsf(ws, h->ws.logits), c.batch_size,
                                                                         c.seq_length, c.vocab_size, logits_out, probs_out);
        CRNN_LAUNCH_CHECK("logits_out_kernel");
    }
    return CRNN_OK;
}

// ---- single operators ------------------------------------------------------------------

int crnn_embedding_gather(const float* table, int32_t V, int32_t H, const int32_t* x, int64_t ldx, int32_t B, int32_t S,
                          float* out, void* stream) {
    if (!table || !x || !out) return set_err(CRNN_ERR_ARG, "null device pointer");
    if (V < 1 || H < 1 || B < 0 || S < 0) return set_err(CRNN_ERR_ARG, "bad shape");
    if ((int64_t)B * S == 0) return CRNN_OK;
    embed_gather_kernel<<<grid_for((int64_t)S * B * H / 4 + 1, 148), 256, 0, (cudaStream_t)stream>>>(table, V, H, x, ldx, B, S, out, nullptr);
    CRNN_LAUNCH_CHECK("embed_gather_kernel");
    return CRNN_OK;
}

int crnn_embedding_scatter_add(float* dtable, int32_t V, int32_t H, const int32_t* x, int64_t ldx, int32_t B, int32_t S,
                               const float* dout, double* sqnorm, void* stream) {
    if (!dtable || !x || !dout) return set_err(CRNN_ERR_ARG, "null device pointer");
    if (V < 1 || H < 1 || B < 0 || S < 0) return set_err(CRNN_ERR_ARG, "bad shape");
    if ((int64_t)B * S == 0) return CRNN_OK;
    embed_scatter_kernel<<<grid_for((int64_t)S * B * H, 148), 256, 0, (cudaStream_t)stream>>>(dtable, V, H, x, ldx, B, S, dout, sqnorm);
    CRNN_LAUNCH_CHECK("embed_scatter_kernel");
    return CRNN_OK;
}

int crnn_softmax_xent(float* logits, int32_t B, int32_t S, int32_t V, const int32_t* y, int64_t ldy, double* loss_sum, void* stream) {
    if (!logits || !y) return set_err(CRNN_ERR_ARG, "null device pointer");
    if (V < 1 || B < 0 || S < 0) return set_err(CRNN_ERR_ARG, "bad shape");
    if ((int64_t)B * S == 0) return CRNN_OK;
    softmax_xent_kernel<<<grid_for((int64_t)S * B * 32, 148), 256, 0, (cudaStream_t)stream>>>(logits, B, S, V, y, ldy, loss_sum, nullptr, 0);
    CRNN_LAUNCH_CHECK("softmax_xent_kernel");
    return CRNN_OK;
}

int crnn_gemm_f32(int32_t ta, int32_t tb, int32_t M, int32_t N, int32_t K, const float* A, int64_t lda, const float* B,
                  int64_t ldb, float* C, int64_t ldc, const float* bias, int32_t accumulate, void* stream) {
    if (!A || !B || !C) return set_err(CRNN_ERR_ARG, "null device pointer");
    if (M < 0 || N < 0 || K < 0) return set_err(CRNN_ERR_ARG, "bad shape");
    return gemm_f32(ta, tb, M, N, K, A, lda, B, ldb, C, ldc, bias, accumulate, (cudaStream_t)stream);
}

int crnn_dropout_mask(const crnn_handle* h, int64_t call_index, int32_t site, float* out, void* stream) {
    if (!h || !out) return set_err(CRNN_ERR_ARG, "null argument");
    if (site < 0 || site > h->cfg.num_layers || call_index < 0) return set_err(CRNN_ERR_ARG, "site must be in [0, num_layers], call_index >= 0");
    const int64_t n = (int64_t)h->cfg.seq_length * h->cfg.batch_size * h->cfg.rnn_size;
    return drop_apply(h, call_index, site, nullptr, out, nullptr, 0, n, (cudaStream_t)stream);
}

int crnn_gemm_bf16(int32_t a_mn, int32_t b_mn, int32_t M, int32_t N, int32_t K, const void* A, int64_t lda,
                   const void* B, int64_t ldb, float* C, int64_t ldc, const float* bias, int32_t accumulate, void* stream) {
    if (!A || !B || !C) return set_err(CRNN_ERR_ARG, "null device pointer");
    if (M < 0 || N < 0 || K < 1) return set_err(CRNN_ERR_ARG, "bad shape");
    return tc::gemm_bf16_tc(M, N, K, (const __nv_bfloat16*)A, lda, a_mn, (const __nv_bfloat16*)B, ldb, b_mn, C, ldc, bias,
                            accumulate, (cudaStream_t)stream);
}

int crnn_gemm_bf16x3(int32_t a_mn, int32_t b_mn, int32_t M, int32_t N, int32_t K, const void* A_hi, const void* A_lo, int64_t lda,
                     const void* B_hi, const void* B_lo, int64_t ldb, float* C, int64_t ldc, const float* bias, int32_t accumulate,
                     void* stream) {
    if (!A_hi || !A_lo || !B_hi || !B_lo || !C) return set_err(CRNN_ERR_ARG, "null device pointer");
    if (M < 0 || N < 0 || K < 1) return set_err(CRNN_ERR_ARG, "bad shape");
    return tc::gemm_bf16_tc(M, N, K, (const __nv_bfloat16*)A_hi, lda, a_mn, (const __nv_bfloat16*)B_hi, ldb, b_mn, C, ldc, bias,
                            accumulate, (cudaStream_t)stream, 148, (const __nv_bfloat16*)A_lo, (const __nv_bfloat16*)B_lo);
}

}  // extern "C"
