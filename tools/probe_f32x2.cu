// Issue cost of scalar and packed fp32 instructions for ONE warp (the situation of the compute warp of
// k_target_rows_ws: it runs alone on its scheduler).  For each instruction: cycles per instruction with 8 independent
// chains (pipe occupancy, "rt") and with 1 dependent chain (latency).  clock64() around an unrolled loop.
//   nvcc -gencode arch=compute_100a,code=sm_100a --fmad=false -o tools/probe_f32x2.bin tools/probe_f32x2.cu
#include <cstdio>
#include <cuda_runtime.h>

typedef unsigned long long u64;
#define ITERS 4096

template <int OP, int CH>
__global__ void k(const float* in, float* out, long long* cyc, float bparam) {
    float a[CH], b = OP >= 6 ? bparam : in[33];   // OP 6 / 7: the multiplier comes from a kernel parameter (uniform register)
    u64 p[CH], q;
    for (int i = 0; i < CH; ++i) {
        a[i] = in[threadIdx.x + i];
        asm("mov.b64 %0, {%1,%2};" : "=l"(p[i]) : "f"(a[i]), "f"(a[i] + 1.f));
    }
    asm("mov.b64 %0, {%1,%2};" : "=l"(q) : "f"(b), "f"(b));
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS / 8; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
#pragma unroll
            for (int i = 0; i < CH; ++i) {
                if (OP == 0) a[i] = __fmul_rn(a[i], b);
                if (OP == 1) a[i] = __fadd_rn(a[i], b);
                if (OP == 2) asm volatile("mul.rn.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(q));
                if (OP == 3) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(p[i]) : "l"(q));
                if (OP == 4) asm volatile("fma.rn.f32x2 %0, %0, %1, %1;" : "+l"(p[i]) : "l"(q));
                if (OP == 5) a[i] = fmaxf(a[i], b);
                if (OP == 6) a[i] = __fmul_rn(a[i], b);
                if (OP == 7) a[i] = __fadd_rn(a[i], b);
            }
        }
    }
    long long t1 = clock64();
    float s = 0;
    for (int i = 0; i < CH; ++i) {
        float x, y;
        asm("mov.b64 {%0,%1}, %2;" : "=f"(x), "=f"(y) : "l"(p[i]));
        s += a[i] + x + y;
    }
    out[threadIdx.x] = s;
    if (threadIdx.x == 0) cyc[0] = t1 - t0;
}

template <int OP, int CH>
void run(const char* name, float* in, float* out, long long* cyc) {
    k<OP, CH><<<1, 32>>>(in, out, cyc, 1.0001f);
    k<OP, CH><<<1, 32>>>(in, out, cyc, 1.0001f);
    cudaDeviceSynchronize();
    long long c;
    cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%-10s chains=%d  %.2f cycles per instruction\n", name, CH, (double)c / (double)(ITERS * CH));
}

int main() {
    float *in, *out;
    long long* cyc;
    cudaMalloc(&in, 4096);
    cudaMalloc(&out, 4096);
    cudaMalloc(&cyc, 8);
    cudaMemset(in, 0, 4096);
    run<0, 1>("FMUL", in, out, cyc);  run<0, 8>("FMUL", in, out, cyc);
    run<1, 1>("FADD", in, out, cyc);  run<1, 8>("FADD", in, out, cyc);
    run<2, 1>("FMUL2", in, out, cyc); run<2, 8>("FMUL2", in, out, cyc);
    run<3, 1>("FADD2", in, out, cyc); run<3, 8>("FADD2", in, out, cyc);
    run<4, 1>("FFMA2", in, out, cyc); run<4, 8>("FFMA2", in, out, cyc);
    run<5, 1>("FMNMX", in, out, cyc); run<5, 8>("FMNMX", in, out, cyc);
    run<6, 1>("FMUL.UR", in, out, cyc); run<6, 8>("FMUL.UR", in, out, cyc); run<6, 2>("FMUL.UR", in, out, cyc); run<6, 4>("FMUL.UR", in, out, cyc);
    run<7, 1>("FADD.UR", in, out, cyc); run<7, 8>("FADD.UR", in, out, cyc);
    run<0, 2>("FMUL", in, out, cyc); run<0, 4>("FMUL", in, out, cyc);
    return 0;
}
