// Write-pattern probe 4: is the persistent-kernel deficit (6.3 vs 7.5 TB/s) a load-balance effect?  Persistent warps
// that take tiles from a global atomic counter (dynamic) instead of a fixed stride (static).
//   nvcc -O3 -arch=sm_100a -o tools/probe_write4.bin tools/probe_write4.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void st4(float4* p, float4 v) {
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ unsigned smid() {
    unsigned r;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(r));
    return r;
}

__global__ void k_static(float4* out, int tile4, int ntiles, int nwarps, unsigned long long* sm_tiles) {
    const int lane = threadIdx.x & 31;
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    for (int t = w; t < ntiles; t += nwarps) {
        float4* o = out + (size_t)t * tile4;
        for (int q = lane; q < tile4; q += 32) st4(o + q, make_float4(1.f, 0.f, (float)t, 0.f));
    }
}
__global__ void k_dynamic(float4* out, int tile4, int ntiles, unsigned* ctr, unsigned long long* sm_tiles) {
    const int lane = threadIdx.x & 31;
    unsigned mine = 0;
    for (;;) {
        unsigned t = 0;
        if (lane == 0) t = atomicAdd(ctr, 1u);
        t = __shfl_sync(0xffffffffu, t, 0);
        if (t >= (unsigned)ntiles) break;
        float4* o = out + (size_t)t * tile4;
        for (int q = lane; q < tile4; q += 32) st4(o + q, make_float4(1.f, 0.f, (float)t, 0.f));
        ++mine;
    }
    if (lane == 0 && sm_tiles) atomicAdd(sm_tiles + smid(), (unsigned long long)mine);
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
    float4* buf;
    unsigned* ctr;
    unsigned long long* sm_tiles;
    cudaMalloc(&buf, total);
    cudaMalloc(&ctr, 4);
    cudaMalloc(&sm_tiles, 256 * 8);
    for (int tile_bytes : {2048, 8192, 65536}) {
        const int tile4 = tile_bytes / 16, ntiles = (int)(total / tile_bytes);
        for (int wps : {8, 16, 32}) {
            const int nwarps = 148 * wps, bs = 64;
            float ms = timeit([&] { k_static<<<nwarps * 32 / bs, bs>>>(buf, tile4, ntiles, nwarps, nullptr); }, 5);
            float md = timeit([&] {
                cudaMemsetAsync(ctr, 0, 4);
                k_dynamic<<<nwarps * 32 / bs, bs>>>(buf, tile4, ntiles, ctr, nullptr);
            }, 5);
            printf("tile %6d B, %2d warps/SM : static %5.0f GB/s | dynamic %5.0f GB/s\n", tile_bytes, wps, total / ms / 1e6,
                   total / md / 1e6);
        }
    }
    // how unequal are the SMs?  tiles taken per SM under dynamic assignment (8 KB tiles, 16 warps/SM)
    {
        const int tile_bytes = 8192, tile4 = tile_bytes / 16, ntiles = (int)(total / tile_bytes), nwarps = 148 * 16;
        cudaMemset(sm_tiles, 0, 256 * 8);
        cudaMemset(ctr, 0, 4);
        k_dynamic<<<nwarps / 2, 64>>>(buf, tile4, ntiles, ctr, sm_tiles);
        unsigned long long h[256];
        cudaMemcpy(h, sm_tiles, sizeof(h), cudaMemcpyDeviceToHost);
        unsigned long long mn = ~0ull, mx = 0, sum = 0;
        int n = 0;
        for (int i = 0; i < 256; ++i)
            if (h[i]) { mn = h[i] < mn ? h[i] : mn; mx = h[i] > mx ? h[i] : mx; sum += h[i]; ++n; }
        printf("tiles per SM under dynamic assignment: %d SMs, min %llu, mean %.0f, max %llu\n", n, mn, (double)sum / n, mx);
        for (int i = 0; i < 256; ++i)
            if (h[i]) printf("%llu%s", h[i], (i % 16 == 15) ? "\n" : " ");
        printf("\n");
    }
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
