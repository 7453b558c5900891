// Write-pattern probe: how fast can HBM absorb (a) a grid-stride fill and (b) the rollout's pattern — 2048 warps,
// each streaming its own contiguous tiles (obs 32,256 B + scalar 2,304 B + action_mask 27,392 B per ply) —
// with trivial data?  Build and run on the GPU box:  nvcc -O3 -arch=sm_100a -o /tmp/pw tools/probe_write.cu && /tmp/pw
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void st4(float4* p, float4 v) {
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

__global__ void k_fill(float4* out, size_t n4) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x)
        st4(out + i, make_float4(1.f, 0.f, 1.f, 0.f));
}

// tiles: per ply, warp w writes t0 bytes at A + w*t0, t1 at B + w*t1, t2 at C + w*t2 (bases rotate over `waves`)
template <int COOP>  // COOP = warps of a CTA that write one tile together (1 = each warp its own tile)
__global__ void k_tiles(float* base, int plies, int waves, size_t wave_floats, int nwarps) {
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5, wpb = blockDim.x >> 5;
    const int tf[3] = {8064, 576, 6848};  // floats per warp tile
    size_t off[3] = {0, (size_t)nwarps * 8064, (size_t)nwarps * (8064 + 576)};
    for (int p = 0; p < plies; ++p) {
        float* wb = base + (size_t)(p % waves) * wave_floats;
        if (COOP == 1) {
            const int w = blockIdx.x * wpb + wib;
            for (int k = 0; k < 3; ++k) {
                float4* o = reinterpret_cast<float4*>(wb + off[k] + (size_t)w * tf[k]);
                for (int q = lane; q < tf[k] / 4; q += 32) st4(o + q, make_float4(1.f, 0.f, (float)p, 0.f));
            }
        } else {
            for (int t = 0; t < wpb; ++t) {
                const int w = blockIdx.x * wpb + t;
                for (int k = 0; k < 3; ++k) {
                    float4* o = reinterpret_cast<float4*>(wb + off[k] + (size_t)w * tf[k]);
                    for (int q = threadIdx.x; q < tf[k] / 4; q += blockDim.x) st4(o + q, make_float4(1.f, 0.f, (float)p, 0.f));
                }
            }
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
    const int nwarps = 2048, plies = 32, waves = 2;
    const size_t wave_floats = (size_t)nwarps * (8064 + 576 + 6848);
    float* buf;
    cudaMalloc(&buf, wave_floats * waves * sizeof(float));
    const double bytes = (double)wave_floats * 4 * plies;
    float ms = timeit([&] { k_fill<<<148 * 16, 256>>>((float4*)buf, wave_floats * waves / 4); }, 20);
    printf("grid-stride fill            : %.3f ms  %.0f GB/s\n", ms, wave_floats * waves * 4.0 / ms / 1e6);
    for (int bs : {32, 64, 128, 256}) {
        ms = timeit([&] { k_tiles<1><<<nwarps * 32 / bs, bs>>>(buf, plies, waves, wave_floats, nwarps); }, 10);
        printf("warp-own tiles, CTA %3d     : %.3f ms  %.0f GB/s\n", bs, ms, bytes / ms / 1e6);
    }
    for (int bs : {64, 128, 256}) {
        ms = timeit([&] { k_tiles<2><<<nwarps * 32 / bs, bs>>>(buf, plies, waves, wave_floats, nwarps); }, 10);
        printf("CTA-cooperative tiles, %3d  : %.3f ms  %.0f GB/s\n", bs, ms, bytes / ms / 1e6);
    }
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
