// ubench_dsmem.cu — what does an all-to-all exchange of row slices cost INSIDE a thread-block cluster?
// Every CTA of a cluster of CS pushes `bytes` from its own shared memory into the same slot of every peer's
// landing buffer and waits until its own buffer has received the CS slices of the round (the exchange a
// weight-stationary recurrence would do once per contraction if a batch tile's slices lived in one cluster).
//   mode 0: one thread issues CS bulk copies  cp.async.bulk.shared::cluster.shared::cta  that complete_tx on the
//           receiver's mbarrier (no L2, no flag, no poll)
//   mode 1: 256 threads write 16-byte pieces with st.shared::cluster, then one remote mbarrier arrive per peer
//   mode 2: the L2 way for comparison: 8-byte global stores, bar.sync, red.release.gpu, ld.acquire.gpu poll, one 1-D bulk
//           load of CS*bytes from global memory
// Prints cycles per round (steady state, slowest CTA) and how many such clusters can be resident.
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tools/ubench_dsmem tools/ubench_dsmem.cu
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t mapa(uint32_t a, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t n) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(b)), "r"(n) : "memory"); }
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(s32(b)), "r"(parity) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t cta_rank() { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }

struct Args { int mode, cs, bytes, rounds; long long* out; uint8_t* gbuf; unsigned* flags; };

__global__ void __launch_bounds__(256, 1) xchg_kernel(Args a) {
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* landing = smem;                                   // cs * bytes
    uint8_t* stage = smem + (size_t)a.cs * a.bytes;            // bytes
    __shared__ uint64_t full[2];
    __shared__ long long t_first, t_last;
    const uint32_t rank = cta_rank();
    const int cluster = blockIdx.x / a.cs;
    if (threadIdx.x == 0) {
        mbar_init(&full[0], a.mode == 1 ? a.cs + 0 : 1);
        mbar_init(&full[1], a.mode == 1 ? a.cs + 0 : 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int i = threadIdx.x; i < a.bytes / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(stage)[i] = rank * 1000 + i;
    __syncthreads();
    cluster_sync();
    uint8_t* gslot = a.gbuf + (size_t)cluster * 2 * a.cs * a.bytes;          // two round buffers per cluster
    unsigned* flag = a.flags + cluster * 32;
    for (int r = 0; r < a.rounds; ++r) {
        const int b = r & 1;
        if (r == 8 && threadIdx.x == 0) t_first = clock64();
        if (a.mode == 0) {
            // (the slice would be written into `stage` by the epilogue threads here)
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncthreads();
            if (threadIdx.x == 0) {
                mbar_expect(&full[b], (uint32_t)(a.cs * a.bytes));
                for (int p = 0; p < a.cs; ++p) {
                    const uint32_t peer = (rank + p) % a.cs;
                    const uint32_t dst = mapa(s32(landing + (size_t)rank * a.bytes), peer);
                    const uint32_t bar = mapa(s32(&full[b]), peer);
                    asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                 ::"r"(dst), "r"(s32(stage)), "r"((uint32_t)a.bytes), "r"(bar) : "memory");
                }
            }
            mbar_wait(&full[b], (r >> 1) & 1);
        } else if (a.mode == 1) {
            const int per = a.bytes / 16;                      // 16-byte pieces per slice
            for (int p = 0; p < a.cs; ++p) {
                const uint32_t peer = (rank + p) % a.cs;
                for (int i = threadIdx.x; i < per; i += blockDim.x) {
                    const uint4 v = reinterpret_cast<const uint4*>(stage)[i];
                    const uint32_t dst = mapa(s32(landing + (size_t)rank * a.bytes + (size_t)i * 16), peer);
                    asm volatile("st.shared::cluster.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
                }
            }
            __syncthreads();
            if (threadIdx.x < a.cs) {
                const uint32_t bar = mapa(s32(&full[b]), threadIdx.x);
                asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(bar) : "memory");
            }
            mbar_wait(&full[b], (r >> 1) & 1);
        } else {
            uint8_t* dst = gslot + (size_t)b * a.cs * a.bytes + (size_t)rank * a.bytes;
            for (int i = threadIdx.x; i < a.bytes / 8; i += blockDim.x)
                reinterpret_cast<uint2*>(dst)[i] = reinterpret_cast<const uint2*>(stage)[i];
            __syncthreads();
            if (threadIdx.x == 0) {
                asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(flag) : "memory");
                unsigned v;
                const unsigned want = (unsigned)(r + 1) * a.cs;
                do { asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory"); } while (v < want);
                asm volatile("fence.proxy.async;" ::: "memory");
                mbar_expect(&full[b], (uint32_t)(a.cs * a.bytes));
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(s32(landing)), "l"(gslot + (size_t)b * a.cs * a.bytes), "r"((uint32_t)(a.cs * a.bytes)), "r"(s32(&full[b])) : "memory");
            }
            mbar_wait(&full[b], (r >> 1) & 1);
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) t_last = clock64();
    __syncthreads();
    // checksum so the copies are real: word 1 of every landed slice
    if (threadIdx.x == 0) {
        long long bad = 0;
        for (int p = 0; p < a.cs; ++p) bad += reinterpret_cast<uint32_t*>(landing + (size_t)p * a.bytes)[1] != (uint32_t)(p * 1000 + 1);
        a.out[blockIdx.x * 2 + 0] = (t_last - t_first) / (a.rounds - 8);
        a.out[blockIdx.x * 2 + 1] = bad;
    }
    cluster_sync();
}

int main(int argc, char** argv) {
    const int nclusters = argc > 1 ? atoi(argv[1]) : 4;
    const int rounds = 208;
    long long* out;
    CK(cudaMalloc(&out, 4096 * 16));
    uint8_t* gbuf; unsigned* flags;
    CK(cudaMalloc(&gbuf, 64u << 20));
    CK(cudaMalloc(&flags, 64 * 128));
    CK(cudaFuncSetAttribute(xchg_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
    CK(cudaFuncSetAttribute(xchg_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
    for (int cs : {8, 16}) {
        for (int extra_kb : {0, 100, 200}) {       // how many clusters fit beside a large resident slice?
            cudaLaunchConfig_t cfg{};
            cfg.gridDim = dim3(cs * nclusters); cfg.blockDim = dim3(256);
            cfg.dynamicSmemBytes = (size_t)extra_kb * 1024 + 16384;
            cudaLaunchAttribute at[1];
            at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
            cfg.attrs = at; cfg.numAttrs = 1;
            int n = -1;
            cudaError_t e = cudaOccupancyMaxActiveClusters(&n, xchg_kernel, &cfg);
            printf("cluster %2d, %3zu KB smem: max active clusters %d (%s)\n", cs, cfg.dynamicSmemBytes / 1024, n, cudaGetErrorString(e));
        }
    }
    for (int mode = 0; mode < 3; ++mode)
        for (int cs : {8, 16})
            for (int bytes : {2048, 4096, 8192}) {
                for (int ncl : {1, nclusters}) {
                    Args a{mode, cs, bytes, rounds, out, gbuf, flags};
                    CK(cudaMemset(flags, 0, 64 * 128));
                    cudaLaunchConfig_t cfg{};
                    cfg.gridDim = dim3(cs * ncl); cfg.blockDim = dim3(256);
                    cfg.dynamicSmemBytes = (size_t)(cs + 1) * bytes + 1024;
                    cudaLaunchAttribute at[1];
                    at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = cs; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
                    cfg.attrs = at; cfg.numAttrs = 1;
                    cudaError_t e = cudaLaunchKernelEx(&cfg, xchg_kernel, a);
                    if (e != cudaSuccess) { printf("mode %d cs %d bytes %d: launch %s\n", mode, cs, bytes, cudaGetErrorString(e)); cudaGetLastError(); continue; }
                    CK(cudaDeviceSynchronize());
                    std::vector<long long> h(cs * ncl * 2);
                    CK(cudaMemcpy(h.data(), out, h.size() * 8, cudaMemcpyDeviceToHost));
                    long long mx = 0, mn = 1ll << 60, bad = 0;
                    for (int i = 0; i < cs * ncl; ++i) { mx = h[2 * i] > mx ? h[2 * i] : mx; mn = h[2 * i] < mn ? h[2 * i] : mn; bad += h[2 * i + 1]; }
                    printf("mode %d  cluster %2d x %d  slice %5d B (%3d KB in per CTA): %6lld .. %6lld cycles per round  bad=%lld\n",
                           mode, cs, ncl, bytes, cs * bytes / 1024, mn, mx, bad);
                }
            }
    return 0;
}
