// common.cuh — shared helpers for the char-RNN kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <math.h>

#define CRNN_OK            0
#define CRNN_ERR_ARG      -1
#define CRNN_ERR_CUDA     -2
#define CRNN_ERR_UNSUP    -3

namespace crnn {

extern thread_local char g_err[512];
extern long long g_launches;   // kernels launched by this library in this process

inline int set_err(int code, const char* fmt, ...) __attribute__((format(printf, 2, 3)));
inline int fail_cuda(cudaError_t e, const char* what);

#define CRNN_CUDA(call)                                                     \
    do {                                                                    \
        cudaError_t e__ = (call);                                           \
        if (e__ != cudaSuccess) return crnn::fail_cuda(e__, #call);         \
    } while (0)

#define CRNN_LAUNCH_CHECK(name)                                             \
    do {                                                                    \
        cudaError_t e__ = cudaGetLastError();                               \
        if (e__ != cudaSuccess) return crnn::fail_cuda(e__, name);          \
        ++crnn::g_launches;                                                 \
    } while (0)

__host__ __device__ __forceinline__ int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }
__host__ __device__ __forceinline__ int ceil_div(int a, int b) { return (a + b - 1) / b; }
__host__ __device__ __forceinline__ int64_t round_up64(int64_t a, int64_t b) { return ceil_div64(a, b) * b; }

// Split-bf16 operands (CRNN_PREC_BF16X3): a float is carried as hi = bf16(v) and lo = bf16(v - hi), 16
// significand bits together; a product keeps hi*hi + lo*hi + hi*lo (the dropped lo*lo term is ~2^-18
// relative), so tensor-core contractions land within ~1e-5 of fp32.  `lo_ofs` is the element offset of
// the lo twin from the hi array (0: plain bf16, no twin).
__device__ __forceinline__ void store_bf16_split(__nv_bfloat16* __restrict__ p, int64_t lo_ofs, float v) {
    const __nv_bfloat16 hi = __float2bfloat16_rn(v);
    p[0] = hi;
    if (lo_ofs) p[lo_ofs] = __float2bfloat16_rn(v - __bfloat162float(hi));
}

// fp32-accurate activations for the parity mode (no fast-math approximations).
__device__ __forceinline__ float sigmoid_acc(float x) { return 1.0f / (1.0f + expf(-x)); }
__device__ __forceinline__ float tanh_acc(float x) { return tanhf(x); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Block-wide double sum; result valid in thread 0.  `sh` must hold 32 doubles.
__device__ __forceinline__ double block_sum_d(double v, double* sh) {
    v = warp_sum_d(v);
    int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) sh[w] = v;
    __syncthreads();
    double r = 0.0;
    if (w == 0) {
        int nw = (blockDim.x + 31) >> 5;
        r = lane < nw ? sh[lane] : 0.0;
        r = warp_sum_d(r);
    }
    __syncthreads();
    return r;
}

// ---- grid-wide barrier for the persistent recurrence kernels --------------------------
// Launched with cudaLaunchCooperativeKernel so all CTAs are co-resident.  `counter` is a
// zero-initialised device word; `target` is the monotonically growing arrival count the
// caller expects after this barrier ((n_barriers_so_far) * gridDim.x).
__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void grid_barrier(unsigned* counter, unsigned target) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(counter) : "memory");
        while (ld_acquire_u32(counter) < target) { }
    }
    __syncthreads();
}

}  // namespace crnn
