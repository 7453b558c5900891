// Write-pattern probe 2: why does a grid-stride fill reach ~7.4 TB/s while the rollout's per-warp tile pattern
// tops out near 6.3 TB/s?  Variants: one tile stream per warp (tile bytes swept), number of resident warps,
// three streams per warp (obs / scalar / mask, the rollout's shape), and warps sharing tiles in address order.
//   nvcc -O3 -arch=sm_100a -o tools/probe_write2.bin tools/probe_write2.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void st4(float4* p, float4 v) {
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void st4_plain(float4* p, float4 v) { *p = v; }
__device__ __forceinline__ void st4_cs(float4* p, float4 v) { __stcs(p, v); }

__global__ void k_fill(float4* out, size_t n4) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x)
        st4(out + i, make_float4(1.f, 0.f, 1.f, 0.f));
}

// every warp writes `reps` tiles of tile4 float4s; tile j of warp w lives at ((j * nwarps + w) * tile4): at any time
// the nwarps active tiles are adjacent in memory (like the rollout's waves: one tile per warp per ply)
template <int MODE>
__global__ void k_tiles(float4* out, int tile4, int reps, int nwarps) {
    const int lane = threadIdx.x & 31;
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    for (int j = 0; j < reps; ++j) {
        float4* o = out + ((size_t)j * nwarps + w) * tile4;
        for (int q = lane; q < tile4; q += 32) {
            const float4 v = make_float4(1.f, 0.f, (float)j, 0.f);
            if (MODE == 0) st4(o + q, v);
            else if (MODE == 1) st4_plain(o + q, v);
            else st4_cs(o + q, v);
        }
    }
}

// same bytes, but each warp's tiles are far apart: warp w owns the contiguous region [w * reps * tile4, ...)
__global__ void k_tiles_own_region(float4* out, int tile4, int reps, int nwarps) {
    const int lane = threadIdx.x & 31;
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    for (int j = 0; j < reps; ++j) {
        float4* o = out + ((size_t)w * reps + j) * tile4;
        for (int q = lane; q < tile4; q += 32) st4(o + q, make_float4(1.f, 0.f, (float)j, 0.f));
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
    const size_t total = (size_t)4 << 30;  // 4 GiB written per launch
    float4* buf;
    cudaMalloc(&buf, total);
    float ms = timeit([&] { k_fill<<<148 * 16, 256>>>(buf, total / 16); }, 10);
    printf("grid-stride fill                         : %.3f ms  %.0f GB/s\n", ms, total / ms / 1e6);
    for (int nwarps : {1184, 2048, 4096, 8192}) {
        for (int tile_bytes : {2048, 8192, 32768, 131072}) {
            const int tile4 = tile_bytes / 16;
            const int reps = (int)(total / ((size_t)nwarps * tile_bytes));
            const double bytes = (double)reps * nwarps * tile_bytes;
            ms = timeit([&] { k_tiles<0><<<nwarps / 2, 64>>>(buf, tile4, reps, nwarps); }, 5);
            float ms1 = timeit([&] { k_tiles<1><<<nwarps / 2, 64>>>(buf, tile4, reps, nwarps); }, 5);
            float ms2 = timeit([&] { k_tiles_own_region<<<nwarps / 2, 64>>>(buf, tile4, reps, nwarps); }, 5);
            printf("warps %5d tile %6d B : adjacent tiles %5.0f GB/s (plain st %5.0f) | own region %5.0f GB/s\n", nwarps,
                   tile_bytes, bytes / ms / 1e6, bytes / ms1 / 1e6, bytes / ms2 / 1e6);
        }
    }
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
