// Write-pattern probe 3: which launch shapes reach the ~7.4 TB/s that torch's fill_ gets?
//   nvcc -O3 -arch=sm_100a -o tools/probe_write3.bin tools/probe_write3.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void st4(float4* p, float4 v) {
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// persistent grid-stride
__global__ void k_fill(float4* out, size_t n4) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x)
        st4(out + i, make_float4(1.f, 0.f, 1.f, 0.f));
}
// one-shot: block b owns a contiguous chunk of blockDim * U float4, written as U passes of blockDim float4
template <int U>
__global__ void k_oneshot(float4* out) {
    float4* o = out + (size_t)blockIdx.x * blockDim.x * U + threadIdx.x;
#pragma unroll
    for (int u = 0; u < U; ++u) st4(o + (size_t)u * blockDim.x, make_float4(1.f, 0.f, 1.f, 0.f));
}
// one-shot, warp-owned tiles: warp w of the grid writes tile4 float4 contiguous
__global__ void k_oneshot_tiles(float4* out, int tile4) {
    const int lane = threadIdx.x & 31;
    const size_t w = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    float4* o = out + w * tile4;
    for (int q = lane; q < tile4; q += 32) st4(o + q, make_float4(1.f, 0.f, 1.f, 0.f));
}
// persistent, warp-owned adjacent tiles, CTAs of `blockDim` threads
__global__ void k_persist_tiles(float4* out, int tile4, int reps, int nwarps) {
    const int lane = threadIdx.x & 31;
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    for (int j = 0; j < reps; ++j) {
        float4* o = out + ((size_t)j * nwarps + w) * tile4;
        for (int q = lane; q < tile4; q += 32) st4(o + q, make_float4(1.f, 0.f, (float)j, 0.f));
    }
}
// persistent, but the warps of the whole grid re-synchronise every `sync_every` tiles through a global counter
// (keeps the write window compact); cooperative-launch-free spin barrier, all CTAs resident by construction
__global__ void k_persist_tiles_sync(float4* out, int tile4, int reps, int nwarps, int sync_every, unsigned* ctr) {
    const int lane = threadIdx.x & 31;
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    for (int j = 0; j < reps; ++j) {
        float4* o = out + ((size_t)j * nwarps + w) * tile4;
        for (int q = lane; q < tile4; q += 32) st4(o + q, make_float4(1.f, 0.f, (float)j, 0.f));
        if ((j + 1) % sync_every == 0) {
            __syncthreads();
            if (threadIdx.x == 0) {
                atomicAdd(ctr, 1u);
                const unsigned target = (unsigned)((j + 1) / sync_every) * gridDim.x;
                while (*((volatile unsigned*)ctr) < target) {}
            }
            __syncthreads();
        }
    }
}

template <class F>
float timeit(F f, int iters) {
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    for (int i = 0; i < 2; ++i) f();
    cudaEventRecord(a);
    for (int i = 0; i < iters; ++i) f();
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms;
    cudaEventElapsedTime(&ms, a, b);
    return ms / iters;
}

int main() {
    const size_t total = (size_t)4 << 30;
    const size_t n4 = total / 16;
    float4* buf;
    unsigned* ctr;
    cudaMalloc(&buf, total);
    cudaMalloc(&ctr, 4);
    float ms;
    for (int bs : {64, 256, 1024})
        for (int per_sm : {8, 16, 32}) {
            if (per_sm * bs > 2048) continue;
            ms = timeit([&] { k_fill<<<148 * per_sm, bs>>>(buf, n4); }, 5);
            printf("grid-stride fill  %2d CTAs/SM x %4d thr : %.0f GB/s\n", per_sm, bs, total / ms / 1e6);
        }
    for (int bs : {64, 128, 256}) {
        ms = timeit([&] { k_oneshot<1><<<(unsigned)(n4 / bs), bs>>>(buf); }, 5);
        printf("one-shot  U=1 block %3d                  : %.0f GB/s\n", bs, total / ms / 1e6);
        ms = timeit([&] { k_oneshot<4><<<(unsigned)(n4 / bs / 4), bs>>>(buf); }, 5);
        printf("one-shot  U=4 block %3d                  : %.0f GB/s\n", bs, total / ms / 1e6);
        ms = timeit([&] { k_oneshot<16><<<(unsigned)(n4 / bs / 16), bs>>>(buf); }, 5);
        printf("one-shot  U=16 block %3d                 : %.0f GB/s\n", bs, total / ms / 1e6);
    }
    for (int tile_bytes : {2048, 32768}) {
        const int tile4 = tile_bytes / 16;
        for (int bs : {64, 256}) {
            ms = timeit([&] { k_oneshot_tiles<<<(unsigned)(total / tile_bytes / (bs / 32)), bs>>>(buf, tile4); }, 5);
            printf("one-shot warp tiles %6d B block %3d    : %.0f GB/s\n", tile_bytes, bs, total / ms / 1e6);
        }
    }
    for (int tile_bytes : {2048, 32768})
        for (int bs : {64, 256}) {
            const int nwarps = 148 * 16;  // 16 warps/SM like the rollout
            const int tile4 = tile_bytes / 16;
            const int reps = (int)(total / ((size_t)nwarps * tile_bytes));
            const double bytes = (double)reps * nwarps * tile_bytes;
            ms = timeit([&] { k_persist_tiles<<<nwarps * 32 / bs, bs>>>(buf, tile4, reps, nwarps); }, 5);
            printf("persistent warp tiles %6d B block %3d  : %.0f GB/s\n", tile_bytes, bs, bytes / ms / 1e6);
            for (int se : {1, 4}) {
                ms = timeit([&] {
                    cudaMemsetAsync(ctr, 0, 4);
                    k_persist_tiles_sync<<<nwarps * 32 / bs, bs>>>(buf, tile4, reps, nwarps, se, ctr);
                }, 5);
                printf("   + grid re-sync every %d tiles           : %.0f GB/s\n", se, bytes / ms / 1e6);
            }
        }
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
