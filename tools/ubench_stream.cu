// ubench_stream.cu — how fast can ONE SM pull an L2-resident operand tile into shared memory, and
// does the answer depend on the shape of the request?  (Decides the layout of the rows the
// recurrence kernels stream every step.)
//   mode 0: 3-D tensor map [kblock][row][64 bf16], 128-byte swizzle, rows 1 KB apart (what the kernels do)
//   mode 1: 1-D bulk copy of the same number of bytes, contiguous in global memory
//   mode 2: 2-D tensor map, one K block of `rows` rows per instruction
// Every CTA of a group of `share` CTAs reads the same bytes (slices of one batch tile do).
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tools/ubench_stream tools/ubench_stream.cu
#include "../distributed_char_rnn_b200/csrc/tc_common.cuh"
#include <stdarg.h>
#include <vector>

namespace crnn {
thread_local char g_err[512] = "";
long long g_launches = 0;
inline int set_err(int code, const char* fmt, ...) { va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof g_err, fmt, ap); va_end(ap); return code; }
inline int fail_cuda(cudaError_t e, const char* what) { return set_err(CRNN_ERR_CUDA, "%s: %s", what, cudaGetErrorString(e)); }
}
using namespace crnn;
using namespace crnn::tc;

__device__ __forceinline__ void bulk_load_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ void mbar_poll(uint64_t* bar, uint32_t parity) {       // test_wait: never suspends
    uint32_t ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    } while (!ok);
}

struct Args { int nw; int poll; int mode, rows, ckb, nkb, nst, rounds, share; const __nv_bfloat16* src; long long* out; };

__global__ void __launch_bounds__(128, 1) stream_kernel(const __grid_constant__ CUtensorMap tm3, const __grid_constant__ CUtensorMap tm2, Args a) {
    extern __shared__ uint8_t raw[];
    uint8_t* ring = reinterpret_cast<uint8_t*>(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t full_all[4][16];
    const int tile = a.rows * 128, stage = tile * a.ckb;
    if (threadIdx.x == 0) { for (int w = 0; w < 4; ++w) for (int i = 0; i < 16; ++i) mbar_init(&full_all[w][i], 1); fence_barrier_init(); }
    __syncthreads();
    const int grp = blockIdx.x / a.share;
    const int row0 = grp * a.rows;
    const int nstage = a.nkb / a.ckb;                  // stages per round (one round = one phase's K extent)
    __shared__ uint64_t go_bar;
    __shared__ long long seen[16];
    if (threadIdx.x == 0) { mbar_init(&go_bar, 1); fence_barrier_init(); }
    __syncthreads();
    const int wid = threadIdx.x >> 5;
    __shared__ uint32_t tm_slot;
    __shared__ uint64_t mma_done;
    if (a.mode == 6) {                                  // mode 5 + the consumer warp issues MMAs on every stage that arrives
        if (threadIdx.x == 0) { mbar_init(&mma_done, 1); fence_barrier_init(); }
        if (wid == 1) tmem_alloc(&tm_slot, 64);
        tc_fence_before();
        __syncthreads();
        tc_fence_after();
    }
    if (a.mode == 5 || a.mode == 6) {
        uint64_t* full = full_all[0];
        if (a.mode == 6 && wid == 1) {
            const uint32_t tmem = __shfl_sync(0xffffffffu, tm_slot, 0);
            const uint32_t idesc = idesc_bf16(a.rows >= 128 ? 128 : 64, 32, 0, 0);
            uint8_t* B = ring + 8 * a.rows * 128;            // 4 KB of zeros behind the ring (never written)
            uint32_t ncommit = 0;
            for (int r = 0; r < a.rounds; ++r) {
                const long long t0 = clock64();
                for (int i = 0; i < nstage; ++i) {
                    mbar_wait(&full[i], r & 1);
                    tc_fence_after();
                    if (r == a.rounds - 1 && threadIdx.x == 32) seen[i] = clock64() - t0;
                    uint64_t da = desc_kmajor(smem_u32(ring + i * stage)), db = desc_kmajor(smem_u32(B));
                    for (int j = 0; j < a.ckb; ++j, da += (uint64_t)(tile >> 4))
                        for (int rep = 0; rep < 3; ++rep) {      // three products per K block, as the split-precision kernels issue
                            mma_bf16_elect(tmem, da, db, idesc, 1u); mma_bf16_elect(tmem, da + 2, db + 2, idesc, 1u);
                            mma_bf16_elect(tmem, da + 4, db + 4, idesc, 1u); mma_bf16_elect(tmem, da + 6, db + 6, idesc, 1u);
                        }
                }
                mma_commit_elect(&mma_done);
                mbar_wait(&mma_done, ncommit & 1); ++ncommit;
                if (threadIdx.x == 32) mbar_arrive(&go_bar);
                __syncwarp();
            }
            if (blockIdx.x == 0 && threadIdx.x == 32) { printf("with MMAs: stage arrival:"); for (int i = 0; i < nstage; ++i) printf(" %lld", seen[i]); printf("\n"); }
            tc_fence_before();
        }
        if (threadIdx.x == 0) {
            long long t0 = clock64();
            for (int r = 0; r < a.rounds; ++r) {
                if (r > 0) mbar_wait(&go_bar, (r - 1) & 1);          // the consumer has seen every stage of the previous round
                for (int i = 0; i < nstage; ++i) {
                    mbar_expect_tx(&full[i], (uint32_t)stage);
                    tma_load_3d(ring + i * stage, &tm3, &full[i], 0, row0, i * a.ckb);
                }
            }
            mbar_wait(&go_bar, (a.rounds - 1) & 1);
            a.out[blockIdx.x] = clock64() - t0;
        } else if (a.mode == 5 && threadIdx.x == 32) {
            for (int r = 0; r < a.rounds; ++r) {
                const long long t0 = clock64();
                for (int i = 0; i < nstage; ++i) { mbar_wait(&full[i], r & 1); if (r == a.rounds - 1) seen[i] = clock64() - t0; }
                mbar_arrive(&go_bar);
            }
            if (blockIdx.x == 0) { printf("stage arrival (cycles after the consumer started waiting):"); for (int i = 0; i < nstage; ++i) printf(" %lld", seen[i]); printf("\n"); }
        }
        return;
    }
    if ((threadIdx.x & 31) == 0 && wid < a.nw) {
        uint64_t* full = full_all[wid];
        ring += (size_t)wid * a.nst * stage;
        long long t0 = clock64();
        uint32_t issued = 0, waited = 0;
        const uint32_t total = (uint32_t)(a.rounds * nstage);
        if (a.mode == 3 || a.mode == 4) {
            // round-at-once: mode 3 arms every stage barrier first and then issues the loads back to back;
            // mode 4 interleaves arm / issue per stage.  All stages of the round are in flight together.
            for (int r = 0; r < a.rounds; ++r) {
                if (a.mode == 3) for (int i = 0; i < nstage; ++i) mbar_expect_tx(&full[i], (uint32_t)stage);
                for (int i = 0; i < nstage; ++i) {
                    if (a.mode == 4) mbar_expect_tx(&full[i], (uint32_t)stage);
                    tma_load_3d(ring + i * stage, &tm3, &full[i], 0, row0, i * a.ckb);
                }
                for (int i = 0; i < nstage; ++i) mbar_wait(&full[i], r & 1);
            }
            waited = total;
        }
        while (waited < total) {
            while (issued < total && issued - waited < (uint32_t)a.nst) {
                const int st = issued % a.nst, kb = (issued % nstage) * a.ckb;
                mbar_expect_tx(&full[st], (uint32_t)stage);
                if (a.mode == 0) tma_load_3d(ring + st * stage, &tm3, &full[st], 0, row0, kb);
                else if (a.mode == 1) bulk_load_1d(ring + st * stage, a.src + ((size_t)grp * a.nkb + kb) * (size_t)(tile / 2), (uint32_t)stage, &full[st]);
                else for (int j = 0; j < a.ckb; ++j) tma_load_2d(ring + st * stage + j * tile, &tm2, &full[st], (kb + j) * 64, row0);
                ++issued;
            }
            if (a.poll) mbar_poll(&full[waited % a.nst], (waited / a.nst) & 1);
            else        mbar_wait(&full[waited % a.nst], (waited / a.nst) & 1);
            ++waited;
        }
        if (wid == 0) a.out[blockIdx.x] = clock64() - t0;
    }
}

int main(int argc, char** argv) {
    const int H = 512, rows_total = 8192;
    __nv_bfloat16* d; cudaMalloc(&d, (size_t)rows_total * H * 2); cudaMemset(d, 0, (size_t)rows_total * H * 2);
    long long* out; cudaMalloc(&out, 148 * 8);
    cudaFuncSetAttribute(stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    const int poll = argc > 1 ? atoi(argv[1]) : 0;
    const int nw = argc > 2 ? atoi(argv[2]) : 1;
    printf("issuing warps = %d (bytes below are PER WARP)\n", nw);
    printf("wait = %s\n", poll ? "test_wait poll" : "try_wait");
    printf("mode rows ckb nst ctas share | KB/round  cyc/round  B/clk/SM  chip TB/s@1.9GHz\n");
    for (int mode : {5, 6})
        for (int rows : {64, 128})
            for (int ckb : {1, 2})
                for (int nst : {8})
                    for (int ctas : {64})
                        for (int share : {16}) {
                            const int nkb = 8, rounds = 200;
                            if ((size_t)rows * 128 * ckb * (nst > 8 / ckb ? nst : 8 / ckb) * nw > 190 * 1024) continue;
                            if (mode == 2 && ckb != 1 && ckb != 4) continue;
                            Args a{nw, poll, mode, rows, ckb, nkb, nst, rounds, share, d, out};
                            CUtensorMap tm3, tm2;
                            if (make_tmap_bf16_kblocks(&tm3, d, rows_total, H, H, rows, ckb) != CRNN_OK) { printf("tmap3 failed %s\n", g_err); return 1; }
                            if (make_tmap_bf16(&tm2, d, rows_total, H, H, rows) != CRNN_OK) { printf("tmap2 failed %s\n", g_err); return 1; }
                            stream_kernel<<<ctas, 128, 200 * 1024>>>(tm3, tm2, a);
                            stream_kernel<<<ctas, 128, 200 * 1024>>>(tm3, tm2, a);
                            if (cudaDeviceSynchronize() != cudaSuccess) { printf("kernel failed: %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
                            std::vector<long long> h(ctas);
                            cudaMemcpy(h.data(), out, ctas * 8, cudaMemcpyDeviceToHost);
                            long long mx = 0; for (long long v : h) mx = v > mx ? v : mx;
                            const double bytes = (double)rows * 128 * nkb, cyc = (double)mx / rounds;
                            printf("%d %4d %2d %2d %4d %3d | %6.0f %9.0f %8.1f %8.2f\n", mode, rows, ckb, nst, ctas, share, bytes / 1024, cyc,
                                   bytes / cyc, bytes / cyc * ctas * 1.9e9 / 1e12);
                        }
    return 0;
}
