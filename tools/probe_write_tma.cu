// Probe: can the emit path of k_rollout / k_make_batch get closer to the write-stream ceiling
// by staging the expanded fp32 tile in shared memory and letting the TMA engine write it
// (cp.async.bulk.global.shared::cta.bulk_group, SASS UBLKCP) instead of one STG.128 per 16 bytes?
//
// Every variant produces the same bytes: per warp a tile of 32 rows x NB two-valued floats generated from a flat bit
// string in shared memory (the product's emit.cuh layout: row r at bit r*NB), tiles handed out one CTA (4 warps)
// per 4 tiles by the hardware dispatcher like k_make_batch.
//   mode 0  STG   : lane q expands nibble q of the flat string -> st.global.L1::no_allocate.v4.f32 (today's path)
//   mode 1  TMA   : the warp expands CH bytes into one of NBUF staging buffers (st.shared.v4, conflict-free),
//                   fence.proxy.async, one lane issues the bulk copy, commit_group; a buffer is reused after
//                   cp.async.bulk.wait_group.read
//   mode 2  FILL  : no bit expansion at all, plain STG.128 of a constant (the write ceiling of this launch shape)
//
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o tools/probe_write_tma.bin tools/probe_write_tma.cu
//   tools/probe_write_tma.bin [tiles]          (default 196608 tiles = 6.3 GB per launch, > L2)
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

constexpr int NB = 252;                    // floats per row (Geister observation board 7x6x6)
constexpr int ROWS = 32;
constexpr int TILE_F = ROWS * NB;          // 8064 floats = 32,256 B
constexpr int TILE_B = TILE_F * 4;
constexpr int FLAT_W = TILE_F / 32;        // 252 words
constexpr int WARPS = 4;

__device__ __forceinline__ uint32_t mix(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}

__device__ __forceinline__ void st_stream_f4(float4* p, float4 v) {
    asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                 : "memory");
}

__device__ __forceinline__ float4 expand(uint32_t nib) {
    return make_float4((nib & 1u) ? 1.f : 0.f, (nib & 2u) ? 1.f : 0.f, (nib & 4u) ? 1.f : 0.f, (nib & 8u) ? 1.f : 0.f);
}

__device__ __forceinline__ void fill_flat(uint32_t* flat, int64_t tile, int lane) {
    // stands for the per-lane game logic + flat_pack_warp: ~8 words per lane
    for (int w = lane; w < FLAT_W; w += 32) flat[w] = mix((uint32_t)tile * 977u + (uint32_t)w);
    __syncwarp();
}

template <int MODE, int CH, int NBUF>
__global__ void __launch_bounds__(WARPS * 32) k_probe(float* __restrict__ out, int64_t ntiles) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t tile = (int64_t)blockIdx.x * WARPS + warp;
    if (tile >= ntiles) return;
    uint32_t* flat = reinterpret_cast<uint32_t*>(smem) + warp * (FLAT_W + 4);
    float* dst = out + tile * TILE_F;
    if (MODE == 2) {
        float4* o4 = reinterpret_cast<float4*>(dst);
        const float4 v = make_float4(1.f, 0.f, 1.f, 0.f);
        for (int q = lane; q < TILE_F / 4; q += 32) st_stream_f4(o4 + q, v);
        return;
    }
    fill_flat(flat, tile, lane);
    const int sh = (lane & 7) << 2;
    if (MODE == 0) {
        float4* o4 = reinterpret_cast<float4*>(dst);
        const uint32_t* fw = flat + (lane >> 3);
        int q = lane;
#pragma unroll 1
        for (; q + 96 < TILE_F / 4; q += 128) {
            uint32_t w[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) w[u] = fw[(q >> 3) - (lane >> 3) + 4 * u];
#pragma unroll
            for (int u = 0; u < 4; ++u) st_stream_f4(o4 + q + 32 * u, expand(w[u] >> sh));
        }
        for (; q < TILE_F / 4; q += 32) st_stream_f4(o4 + q, expand(flat[q >> 3] >> sh));
        return;
    }
    // ---- MODE 1: stage in shared memory, bulk-copy out ----
    constexpr int CH4 = CH / 16;                                   // float4 per chunk
    float4* stage = reinterpret_cast<float4*>(smem + WARPS * (FLAT_W + 4) * 4 + (size_t)warp * NBUF * CH);
    constexpr int NCH = (TILE_B + CH - 1) / CH;
#pragma unroll 1
    for (int c = 0; c < NCH; ++c) {
        float4* buf = stage + (c % NBUF) * CH4;
        if (c >= NBUF) {                                            // the copy that read this buffer must be done
            if (lane == 0) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(NBUF - 1) : "memory");
            __syncwarp();
        }
        const int q0 = c * CH4;
        const int nq = min(CH4, TILE_F / 4 - q0);
#pragma unroll 4
        for (int q = lane; q < nq; q += 32) buf[q] = expand(flat[(q0 + q) >> 3] >> (((q0 + q) & 7) << 2));
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) {
            const uint32_t src = (uint32_t)__cvta_generic_to_shared(buf);
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst + (size_t)q0 * 4),
                         "r"(src), "r"(nq * 16)
                         : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    __syncwarp();
}

template <int MODE, int CH, int NBUF>
static void run(const char* name, float* out, int64_t ntiles, int reps) {
    const size_t smem = (size_t)WARPS * (FLAT_W + 4) * 4 + (MODE == 1 ? (size_t)WARPS * NBUF * CH : 0);
    auto kern = k_probe<MODE, CH, NBUF>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    int per_sm = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, WARPS * 32, smem);
    const int grid = (int)((ntiles + WARPS - 1) / WARPS);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    kern<<<grid, WARPS * 32, smem>>>(out, ntiles);
    cudaDeviceSynchronize();
    float best = 1e30f, sum = 0.f;
    for (int r = 0; r < reps; ++r) {
        cudaEventRecord(e0);
        kern<<<grid, WARPS * 32, smem>>>(out, ntiles);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        best = ms < best ? ms : best;
        sum += ms;
    }
    const double bytes = (double)ntiles * TILE_B;
    cudaError_t err = cudaGetLastError();
    printf("%-28s smem/CTA %6zu B  CTAs/SM %2d  mean %.3f ms  best %.3f ms  %.2f TB/s (mean) %s\n", name, smem, per_sm,
           sum / reps, best, bytes / (sum / reps * 1e-3) / 1e12, err == cudaSuccess ? "" : cudaGetErrorString(err));
}

__global__ void k_check(const float* a, const float* b, int64_t n, unsigned long long* bad) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        if (a[i] != b[i]) atomicAdd(bad, 1ull);
}

int main(int argc, char** argv) {
    const int64_t ntiles = argc > 1 ? atoll(argv[1]) : 196608;
    float* out;
    cudaMalloc(&out, (size_t)ntiles * TILE_B);
    // correctness of the TMA variants against the STG variant on a small prefix
    {
        const int64_t nt = 2048;
        float *a, *b;
        unsigned long long* bad;
        cudaMalloc(&a, nt * TILE_B);
        cudaMalloc(&b, nt * TILE_B);
        cudaMalloc(&bad, 8);
        cudaMemset(bad, 0, 8);
        cudaMemset(b, 0xFF, nt * TILE_B);
        k_probe<0, 4096, 2><<<(int)(nt / WARPS), WARPS * 32, WARPS * (FLAT_W + 4) * 4>>>(a, nt);
        const size_t smem = (size_t)WARPS * (FLAT_W + 4) * 4 + (size_t)WARPS * 2 * 4096;
        cudaFuncSetAttribute(k_probe<1, 4096, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        k_probe<1, 4096, 2><<<(int)(nt / WARPS), WARPS * 32, smem>>>(b, nt);
        k_check<<<592, 256>>>(a, b, nt * TILE_F, bad);
        unsigned long long h = 1;
        cudaMemcpy(&h, bad, 8, cudaMemcpyDeviceToHost);
        printf("TMA variant vs STG variant: %llu mismatching floats (%s)\n", h, cudaGetErrorString(cudaGetLastError()));
        cudaFree(a); cudaFree(b); cudaFree(bad);
    }
    const int reps = 10;
    run<2, 4096, 2>("fill STG.128 (ceiling)", out, ntiles, reps);
    run<0, 4096, 2>("emit STG.128 (today)", out, ntiles, reps);
    run<1, 2048, 2>("emit TMA 2 KB x2", out, ntiles, reps);
    run<1, 4096, 2>("emit TMA 4 KB x2", out, ntiles, reps);
    run<1, 4096, 3>("emit TMA 4 KB x3", out, ntiles, reps);
    run<1, 8192, 2>("emit TMA 8 KB x2", out, ntiles, reps);
    run<1, 16384, 2>("emit TMA 16 KB x2", out, ntiles, reps);
    run<1, 32256, 1>("emit TMA whole tile x1", out, ntiles, reps);
    cudaFree(out);
    return 0;
}
