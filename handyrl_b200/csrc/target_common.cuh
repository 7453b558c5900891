// Shared definitions of the target-recursion kernels (target_scan.cu: generic strided path,
// target_rows.cu: row-per-thread cp.async pipeline for the contiguous P<=2 layouts).
#pragma once
#include "common.cuh"
#include "../../include/b200rl.h"

struct View3 {
    const float* p;
    int64_t sb, st, sp;
    int px;   // 1 when the player dim is broadcast (stride 0), else P
    int vec;  // 128-bit path allowed (contiguous along (t,p), 16B aligned rows)
};

struct ScanParams {
    View3 v, rew, rho, c;
    const float* boot;
    int64_t boot_sb, boot_sp;
    float* out_t;
    float* out_a;
    int B, T, P;
    int rows;  // batch rows per CTA
    float oml, lmb, gamma, gl;
};

__device__ __forceinline__ float max_nanprop(float a, float b) {
    // torch.max propagates NaN; fmaxf does not.  max.NaN.f32 is a single FMNMX.NAN.
    float r;
    asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
    return r;
}

// Per-sequence recurrence state.  `step` is called for t = T-1 down to 0 and
// returns (target, advantage) for that t.
template <int MODE>
struct Rec;

template <>
struct Rec<B200RL_TARGET_TD> {  // losses.py:20-28
    float G, v1;
    __device__ __forceinline__ void step(bool last, float v, float r, float, float, float R, const ScanParams& k,
                                         float& tgt, float& adv) {
        if (last) {
            G = R;
        } else {
            float a = __fmul_rn(k.oml, v1);
            float b = __fmul_rn(k.lmb, G);
            float c = __fadd_rn(a, b);
            G = __fadd_rn(r, __fmul_rn(k.gamma, c));
        }
        v1 = v;
        tgt = G;
        adv = __fsub_rn(G, v);
    }
};

template <>
struct Rec<B200RL_TARGET_UPGO> {  // losses.py:31-40
    float G, v1;
    __device__ __forceinline__ void step(bool last, float v, float r, float, float, float R, const ScanParams& k,
                                         float& tgt, float& adv) {
        if (last) {
            G = R;
        } else {
            float a = __fmul_rn(k.oml, v1);
            float b = __fmul_rn(k.lmb, G);
            float c = __fadd_rn(a, b);
            G = __fadd_rn(r, __fmul_rn(k.gamma, max_nanprop(v1, c)));
        }
        v1 = v;
        tgt = G;
        adv = __fsub_rn(G, v);
    }
};

template <>
struct Rec<B200RL_TARGET_VTRACE> {  // losses.py:43-58
    float acc, v1, vs1;
    __device__ __forceinline__ void step(bool last, float v, float r, float rho, float c, float R,
                                         const ScanParams& k, float& tgt, float& adv) {
        float vt1 = last ? R : v1;
        float vst1 = last ? R : vs1;
        float delta = __fmul_rn(rho, __fsub_rn(__fadd_rn(r, __fmul_rn(k.gamma, vt1)), v));
        acc = last ? delta : __fadd_rn(delta, __fmul_rn(__fmul_rn(k.gl, c), acc));
        float vs = __fadd_rn(acc, v);
        tgt = vs;
        adv = __fsub_rn(__fadd_rn(r, __fmul_rn(k.gamma, vst1)), v);
        v1 = v;
        vs1 = vs;
    }
};


// Row-per-thread fast path (target_rows.cu).  Returns -1 when the layout is not eligible
// (caller falls back to the generic kernel), else a b200rl status.
int b200rl_launch_target_rows(int mt, int ma, const ScanParams& k, bool has_rew, cudaStream_t st);
