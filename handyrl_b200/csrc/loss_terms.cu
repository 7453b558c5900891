// Loss pre-processing around the target recursion (handyrl/train.py:224-239) as fused row kernels.
//
// compute_loss needs, per (b, t) row of A logits,
//     log pi_b(a) = log_softmax(batch['policy'])[a] * emask          (behaviour policy, train.py:225)
//     log pi_t(a) = log_softmax(outputs['policy'])[a] * emask        (current policy,  train.py:226)
//     rho = exp(log pi_t - log pi_b),  clipped_rho = clamp(rho, 0, 1),  c = clamp(rho, 0, 1)   (228-231)
// which the reference computes with ~8 elementwise / reduction launches and three passes over the rows.
// Here one warp reads both rows once.  The backward of log pi_t w.r.t. the current logits,
//     d/dx_j = g * emask * ([j == a] - softmax(x)_j),
// is a second row kernel, wired in by a torch.autograd.Function on the Python side.
// The value-head symmetrisation + episode-mask blend (train.py:236-239) is a small elementwise kernel
// with the reference's op order (bit-identical to the torch expression).
#include "common.cuh"
#include "../../include/b200rl.h"

namespace {

constexpr int WPB = 8;  // warps (rows) per CTA

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// log_softmax(row)[a] for one warp; also returns max and log-sum-exp pieces for reuse
__device__ __forceinline__ float row_log_selected(const float* __restrict__ row, int A, int a, int lane, float& mx,
                                                  float& sum) {
    float m = -INFINITY;
    for (int j = lane; j < A; j += 32) m = fmaxf(m, row[j]);
    m = warp_max(m);
    float s = 0.f;
    for (int j = lane; j < A; j += 32) s += expf(row[j] - m);
    s = warp_sum(s);
    mx = m;
    sum = s;
    return (row[a] - m) - logf(s);
}

__global__ void __launch_bounds__(WPB * 32)
k_policy_terms(const float* __restrict__ pol_b, const float* __restrict__ pol_t, const long long* __restrict__ action,
               const float* __restrict__ emask, int64_t rows, int A, float clip_rho, float clip_c,
               float* __restrict__ logp_t, float* __restrict__ rho, float* __restrict__ c) {
    const int lane = threadIdx.x & 31;
    const int64_t r = blockIdx.x * (int64_t)WPB + (threadIdx.x >> 5);
    if (r >= rows) return;
    const int a = (int)action[r];
    const float em = emask[r];
    float m, s;
    const float lb = row_log_selected(pol_b + r * A, A, a, lane, m, s) * em;
    const float lt = row_log_selected(pol_t + r * A, A, a, lane, m, s) * em;
    if (lane == 0) {
        const float ratio = expf(lt - lb);
        logp_t[r] = lt;
        rho[r] = fminf(fmaxf(ratio, 0.f), clip_rho);
        c[r] = fminf(fmaxf(ratio, 0.f), clip_c);
    }
}

__global__ void __launch_bounds__(WPB * 32)
k_policy_terms_bwd(const float* __restrict__ pol_t, const long long* __restrict__ action,
                   const float* __restrict__ emask, const float* __restrict__ grad_out, int64_t rows, int A,
                   float* __restrict__ grad_pol) {
    const int lane = threadIdx.x & 31;
    const int64_t r = blockIdx.x * (int64_t)WPB + (threadIdx.x >> 5);
    if (r >= rows) return;
    const float* row = pol_t + r * A;
    const int a = (int)action[r];
    const float g = grad_out[r] * emask[r];
    float m = -INFINITY;
    for (int j = lane; j < A; j += 32) m = fmaxf(m, row[j]);
    m = warp_max(m);
    float s = 0.f;
    for (int j = lane; j < A; j += 32) s += expf(row[j] - m);
    s = warp_sum(s);
    const float inv = 1.f / s;
    for (int j = lane; j < A; j += 32) grad_pol[r * A + j] = g * ((j == a ? 1.f : 0.f) - expf(row[j] - m) * inv);
}

// train.py:236-239 for (N, P) rows: P == 2 and symmetric -> v <- (v + (-swap(v))) / (omask.sum + 1e-8);
// then v <- v * emask + outcome * (1 - emask)
__global__ void k_value_blend(const float* __restrict__ v, const float* __restrict__ omask,
                              const float* __restrict__ emask, const float* __restrict__ outcome, int64_t n_bt, int T,
                              int P, int symmetric, float* __restrict__ out) {
    const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (i >= n_bt * P) return;
    const int64_t bt = i / P;
    const int p = (int)(i - bt * P);
    const int64_t b = bt / T;
    float x = v[i];
    if (symmetric && P == 2) {
        const float other = v[bt * 2 + (1 - p)];
        const float den = __fadd_rn(__fadd_rn(omask[bt * 2], omask[bt * 2 + 1]), 1e-8f);
        x = __fdiv_rn(__fadd_rn(x, -other), den);
    }
    const float em = emask[bt];
    out[i] = __fadd_rn(__fmul_rn(x, em), __fmul_rn(outcome[b * P + p], __fsub_rn(1.f, em)));
}

}  // namespace

extern "C" int b200rl_policy_terms(const float* policy_behaviour, const float* policy_target, const int64_t* action,
                                   const float* emask, int64_t rows, int64_t A, double clip_rho, double clip_c,
                                   float* out_logp_target, float* out_rho, float* out_c, void* stream) {
    B200RL_REQUIRE(policy_behaviour && policy_target && action && emask && out_logp_target && out_rho && out_c,
                   "policy_terms: null pointer");
    B200RL_REQUIRE(rows >= 0 && A >= 1 && A < (1 << 20), "policy_terms: bad rows/A");
    if (rows == 0) return B200RL_OK;
    k_policy_terms<<<(unsigned)((rows + WPB - 1) / WPB), WPB * 32, 0, (cudaStream_t)stream>>>(
        policy_behaviour, policy_target, (const long long*)action, emask, rows, (int)A, (float)clip_rho, (float)clip_c,
        out_logp_target, out_rho, out_c);
    return b200rl_check_launch("k_policy_terms");
}

extern "C" int b200rl_policy_terms_backward(const float* policy_target, const int64_t* action, const float* emask,
                                            const float* grad_logp, int64_t rows, int64_t A, float* out_grad_policy,
                                            void* stream) {
    B200RL_REQUIRE(policy_target && action && emask && grad_logp && out_grad_policy, "policy_terms_backward: null pointer");
    B200RL_REQUIRE(rows >= 0 && A >= 1 && A < (1 << 20), "policy_terms_backward: bad rows/A");
    if (rows == 0) return B200RL_OK;
    k_policy_terms_bwd<<<(unsigned)((rows + WPB - 1) / WPB), WPB * 32, 0, (cudaStream_t)stream>>>(
        policy_target, (const long long*)action, emask, grad_logp, rows, (int)A, out_grad_policy);
    return b200rl_check_launch("k_policy_terms_bwd");
}

extern "C" int b200rl_value_blend(const float* values, const float* omask, const float* emask, const float* outcome,
                                  int64_t B, int64_t T, int64_t P, int symmetric, float* out, void* stream) {
    B200RL_REQUIRE(values && omask && emask && outcome && out, "value_blend: null pointer");
    B200RL_REQUIRE(B >= 0 && T >= 1 && P >= 1, "value_blend: bad shape");
    const int64_t n = B * T * P;
    if (n == 0) return B200RL_OK;
    k_value_blend<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(values, omask, emask, outcome, B * T,
                                                                                 (int)T, (int)P, symmetric, out);
    return b200rl_check_launch("k_value_blend");
}
