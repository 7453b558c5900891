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

// Per-sequence recurrence state.  `first` handles t = T-1 (the bootstrap step), `next` every earlier t in
// descending order; both return (target, advantage) of that t.  Splitting the two keeps the hot loop free of
// the `t == T-1` selects.
template <int MODE>
struct Rec;

template <>
struct Rec<B200RL_TARGET_TD> {  // losses.py:20-28
    float G, v1;
    __device__ __forceinline__ void first(float v, float, float, float, float R, const ScanParams&, float& tgt,
                                          float& adv) {
        G = R;
        v1 = v;
        tgt = G;
        adv = __fsub_rn(G, v);
    }
    __device__ __forceinline__ void next(float v, float r, float, float, const ScanParams& k, float& tgt, float& adv) {
        float a = __fmul_rn(k.oml, v1);
        float b = __fmul_rn(k.lmb, G);
        float c = __fadd_rn(a, b);
        G = __fadd_rn(r, __fmul_rn(k.gamma, c));
        v1 = v;
        tgt = G;
        adv = __fsub_rn(G, v);
    }
};

template <>
struct Rec<B200RL_TARGET_UPGO> {  // losses.py:31-40
    float G, v1;
    __device__ __forceinline__ void first(float v, float, float, float, float R, const ScanParams&, float& tgt,
                                          float& adv) {
        G = R;
        v1 = v;
        tgt = G;
        adv = __fsub_rn(G, v);
    }
    __device__ __forceinline__ void next(float v, float r, float, float, const ScanParams& k, float& tgt, float& adv) {
        float a = __fmul_rn(k.oml, v1);
        float b = __fmul_rn(k.lmb, G);
        float c = __fadd_rn(a, b);
        G = __fadd_rn(r, __fmul_rn(k.gamma, max_nanprop(v1, c)));
        v1 = v;
        tgt = G;
        adv = __fsub_rn(G, v);
    }
};

template <>
struct Rec<B200RL_TARGET_VTRACE> {  // losses.py:43-58
    float acc, v1, vs1;
    __device__ __forceinline__ void first(float v, float r, float rho, float, float R, const ScanParams& k,
                                          float& tgt, float& adv) {
        acc = __fmul_rn(rho, __fsub_rn(__fadd_rn(r, __fmul_rn(k.gamma, R)), v));   // delta[T-1]
        float vs = __fadd_rn(acc, v);
        tgt = vs;
        adv = __fsub_rn(__fadd_rn(r, __fmul_rn(k.gamma, R)), v);
        v1 = v;
        vs1 = vs;
    }
    __device__ __forceinline__ void next(float v, float r, float rho, float c, const ScanParams& k, float& tgt,
                                         float& adv) {
        float delta = __fmul_rn(rho, __fsub_rn(__fadd_rn(r, __fmul_rn(k.gamma, v1)), v));
        acc = __fadd_rn(delta, __fmul_rn(__fmul_rn(k.gl, c), acc));
        float vs = __fadd_rn(acc, v);
        tgt = vs;
        adv = __fsub_rn(__fadd_rn(r, __fmul_rn(k.gamma, vs1)), v);
        v1 = v;
        vs1 = vs;
    }
    // kept for the generic kernel's call sites
};

template <int MODE>
__device__ __forceinline__ void rec_step(Rec<MODE>& rec, bool last, float v, float r, float rho, float c, float R,
                                         const ScanParams& k, float& tgt, float& adv) {
    if (last) rec.first(v, r, rho, c, R, k, tgt, adv);
    else rec.next(v, r, rho, c, k, tgt, adv);
}

// Row-per-thread fast path (target_rows.cu).  Returns -1 when the layout is not eligible
// (caller falls back to the generic kernel), else a b200rl status.
int b200rl_launch_target_rows(int mt, int ma, const ScanParams& k, bool has_rew, cudaStream_t st);
// Two heads (same B, T, P; rows of <= 64 steps) in one launch; -1 when not eligible.
int b200rl_launch_target_rows2(int mt, int ma, const ScanParams& k0, const ScanParams& k1, cudaStream_t st);
