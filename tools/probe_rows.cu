// Row-pattern probe for the target recursion: what bandwidth does HBM give a kernel that moves data in the
// recursion's access pattern but does no arithmetic?  Each warp owns 32 rows of RB bytes (row stride = RB, as in a
// contiguous (B, T, P) tensor) and walks them from the END backwards in chunks of C bytes per row: per chunk it reads
// 32 x C bytes from each of two inputs and writes the same to each of two outputs (the UPGO/TD traffic shape,
// 16 B per element).  Sweeps C (bytes per row per fetch) and resident warps per SM; also the plain streaming copy.
//   nvcc -O3 -arch=sm_100a -o /tmp/pr tools/probe_rows.cu && /tmp/pr [T] [B]
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ float4 ld4(const float4* p) {
    float4 v;
    asm volatile("ld.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ void st4(float4* p, float4 v) {
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

__global__ void k_stream(const float4* a, const float4* b, float4* oa, float4* ob, size_t n4) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        float4 x = ld4(a + i), y = ld4(b + i);
        st4(oa + i, x);
        st4(ob + i, y);
    }
}

// C16 = 16-byte pieces per row per chunk; U = independent pieces in flight per lane per tensor
template <int U>
__global__ void __launch_bounds__(32) k_rows(const float4* a, const float4* b, float4* oa, float4* ob, int rows,
                                             int rb16, int c16, int ntiles) {
    const int lane = threadIdx.x;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const size_t base = (size_t)tile * 32 * rb16;
        for (int c0 = rb16 - c16; c0 >= 0; c0 -= c16) {
            const int pieces = 32 * c16;
            for (int j = lane; j < pieces; j += 32 * U) {
                float4 x[U], y[U];
                size_t off[U];
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const int idx = j + 32 * u;
                    const int r = idx / c16, o = idx - r * c16;
                    off[u] = base + (size_t)r * rb16 + c0 + o;
                    if (idx < pieces) {
                        x[u] = ld4(a + off[u]);
                        y[u] = ld4(b + off[u]);
                    }
                }
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    if (j + 32 * u < pieces) {
                        st4(oa + off[u], x[u]);
                        st4(ob + off[u], y[u]);
                    }
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

int main(int argc, char** argv) {
    const int T = argc > 1 ? atoi(argv[1]) : 1024, B = argc > 2 ? atoi(argv[2]) : 65536, P = 2;
    const int rb16 = T * P * 4 / 16;
    const size_t n4 = (size_t)B * rb16;
    float4 *a, *b, *oa, *ob;
    cudaMalloc(&a, n4 * 16);
    cudaMalloc(&b, n4 * 16);
    cudaMalloc(&oa, n4 * 16);
    cudaMalloc(&ob, n4 * 16);
    cudaMemset(a, 0, n4 * 16);
    cudaMemset(b, 0, n4 * 16);
    const double bytes = 4.0 * n4 * 16;
    printf("T=%d B=%d P=%d  row = %d B, %.2f GB moved per pass\n", T, B, P, rb16 * 16, bytes / 1e9);
    float ms = timeit([&] { k_stream<<<148 * 16, 256>>>(a, b, oa, ob, n4); }, 10);
    printf("streaming 2-in 2-out copy        : %.3f ms  %.0f GB/s\n", ms, bytes / ms / 1e6);
    const int ntiles = B / 32;
    for (int cbytes : {64, 128, 256, 512, 1024, 2048}) {
        const int c16 = cbytes / 16;
        if (rb16 % c16) continue;
        for (int wps : {4, 8, 16, 32}) {
            const int grid = 148 * wps < ntiles ? 148 * wps : ntiles;
            ms = timeit([&] { k_rows<4><<<grid, 32>>>(a, b, oa, ob, B, rb16, c16, ntiles); }, 5);
            float ms8 = timeit([&] { k_rows<8><<<grid, 32>>>(a, b, oa, ob, B, rb16, c16, ntiles); }, 5);
            printf("rows: %4d B/row/fetch, %2d warps/SM : U=4 %.3f ms %5.0f GB/s | U=8 %.3f ms %5.0f GB/s\n", cbytes, wps, ms,
                   bytes / ms / 1e6, ms8, bytes / ms8 / 1e6);
        }
    }
    printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    return 0;
}
