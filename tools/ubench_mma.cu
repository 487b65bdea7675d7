// ubench_mma.cu — cost of a short chain of small tcgen05.mma instructions (operands in shared memory,
// K-major SWIZZLE_128B, bf16) from first issue to the commit's mbarrier arrival, as the recurrence
// kernels issue them: one thread, descriptors advanced by adds, one commit at the end.
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tools/ubench_mma tools/ubench_mma.cu
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

struct Args { int M, N, nkb, rows, rounds, commits; long long* out; };

__global__ void __launch_bounds__(128, 1) mma_kernel(Args a) {
    extern __shared__ uint8_t raw[];
    uint8_t* sm = reinterpret_cast<uint8_t*>(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
    __shared__ uint64_t done, go;
    __shared__ uint32_t slot;
    __shared__ long long t_first, t_issued;
    for (int i = threadIdx.x; i < 160 * 1024 / 16; i += 128) reinterpret_cast<uint4*>(sm)[i] = make_uint4(0, 0, 0, 0);
    fence_proxy_async();
    if (threadIdx.x == 0) { mbar_init(&done, 1); mbar_init(&go, 1); fence_barrier_init(); }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 1) tmem_alloc(&slot, 256);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = slot;
    const int tileA = a.rows * 128, tileB = a.N * 128;          // bytes per K block
    uint8_t* A = sm;
    uint8_t* B = sm + 64 * 1024;
    long long sum_issue = 0, sum_done = 0;
    if (warp == 1 && lane == 0) {
        const uint32_t idesc = idesc_bf16(a.M, a.N, 0, 0);
        for (int r = 0; r < a.rounds; ++r) {
            mbar_wait(&go, (r & 1) ^ 1);                         // consumer has seen the previous accumulator
            tc_fence_after();
            const long long t0 = clock64();
            uint64_t da = desc_kmajor(smem_u32(A)), db = desc_kmajor(smem_u32(B));
            const uint64_t sa = (uint64_t)(tileA >> 4), sb = (uint64_t)(tileB >> 4);
            for (int kb = 0; kb < a.nkb; ++kb, da += sa, db += sb) {
                mma_bf16(tmem, da, db, idesc, (uint32_t)(kb != 0));
                mma_bf16(tmem, da + 2, db + 2, idesc, 1u);
                mma_bf16(tmem, da + 4, db + 4, idesc, 1u);
                mma_bf16(tmem, da + 6, db + 6, idesc, 1u);
            }
            mma_commit(&done);
            t_first = t0;
            t_issued = clock64();
        }
    } else if (warp == 2 && lane == 0) {
        for (int r = 0; r < a.rounds; ++r) {
            mbar_wait(&done, r & 1);
            tc_fence_after();
            const long long t1 = clock64();
            if (r >= 4) { sum_issue += t_issued - t_first; sum_done += t1 - t_first; }
            tc_fence_before();
            mbar_arrive(&go);
        }
        a.out[blockIdx.x * 2] = sum_issue / (a.rounds - 4);
        a.out[blockIdx.x * 2 + 1] = sum_done / (a.rounds - 4);
    }
    __syncthreads();
    if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem, 256); }
}

int main() {
    long long* out; cudaMalloc(&out, 148 * 16);
    cudaFuncSetAttribute(mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 170 * 1024);
    printf("   M    N  nkb (MMAs) | issue cyc  (per MMA) | first issue -> commit arrival  (per MMA)   floor/MMA\n");
    struct C { int M, N, nkb; };
    for (C c : std::vector<C>{{64, 64, 8}, {128, 64, 8}, {64, 64, 2}, {128, 64, 2}, {64, 64, 1}, {128, 32, 8}, {128, 16, 8}, {128, 128, 8},
                              {128, 256, 4}, {64, 128, 8}, {64, 32, 8}, {128, 64, 4}, {128, 32, 24}, {128, 64, 16}}) {
        Args a{c.M, c.N, c.nkb, c.M, 64, 1, out};
        if ((size_t)c.nkb * c.M * 128 > 64 * 1024 || (size_t)c.nkb * c.N * 128 > 96 * 1024) { printf("skip\n"); continue; }
        mma_kernel<<<64, 128, 170 * 1024>>>(a);
        if (cudaDeviceSynchronize() != cudaSuccess) { printf("kernel failed: %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
        long long h[2]; cudaMemcpy(h, out, 16, cudaMemcpyDeviceToHost);
        const int n = c.nkb * 4;
        printf("%4d %4d %4d (%3d)  | %8lld  (%6.1f) | %8lld  (%6.1f)   %d\n", c.M, c.N, c.nkb, n, h[0], (double)h[0] / n, h[1], (double)h[1] / n,
               (c.M > 128 ? c.M : 128) * c.N / 256);
    }
    return 0;
}
