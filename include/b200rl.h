/*
 * b200rl.h — C-ABI of the B200-native HandyRL hot path (libb200rl.so).
 *
 * Conventions:
 *  - every function returns 0 on success, non-zero on error; b200rl_last_error()
 *    returns a thread-local message.  Nothing throws across this boundary.
 *  - all data pointers are DEVICE pointers owned by the caller (torch-allocated)
 *    unless the name ends in `_host`; sizes/strides are int64 in ELEMENTS.
 *  - `stream` is a cudaStream_t passed as void* (torch.cuda.current_stream().cuda_stream).
 *  - no hidden allocation and no synchronisation inside the launchers.
 *
 * Each entry point cites the reference interface (file:line under the HandyRL
 * tree) that it replaces.
 */
#ifndef B200RL_H
#define B200RL_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* b200rl_last_error(void);
int b200rl_version(void);

/* ------------------------------------------------------------------ */
/* Part 3: off-policy target recursion — handyrl/losses.py:61-74       */
/* ------------------------------------------------------------------ */

enum { B200RL_TARGET_MC = 0, B200RL_TARGET_TD = 1, B200RL_TARGET_UPGO = 2, B200RL_TARGET_VTRACE = 3 };

/*
 * Time-reversed scan over (B, T, P) fp32 sequences, replaces
 * temporal_difference / upgo / vtrace (handyrl/losses.py:20-58).
 *   values   (B,T,P) with element strides v_sb, v_st, v_sp
 *   boot     (B,P)   = returns[:, -1]  (the bootstrap), strides boot_sb, boot_sp
 *   rewards  (B,T,P) or NULL (reference passes None for the value head)
 *   rhos, cs (B,T,P) views (stride 0 on broadcast dims); may be NULL unless mode==VTRACE
 *   out_targets, out_adv  contiguous (B,T,P) fp32
 *   lmb, gamma are the Python doubles of the reference signature; the scalar
 *   roundings ((float)(1-lmb), (float)(gamma*lmb)) follow torch's scalar rule.
 */
int b200rl_target_scan(int mode, const float* values, int64_t v_sb, int64_t v_st, int64_t v_sp,
                       const float* boot, int64_t boot_sb, int64_t boot_sp,
                       const float* rewards, int64_t r_sb, int64_t r_st, int64_t r_sp,
                       const float* rhos, int64_t rho_sb, int64_t rho_st, int64_t rho_sp,
                       const float* cs, int64_t c_sb, int64_t c_st, int64_t c_sp,
                       float* out_targets, float* out_adv, int64_t B, int64_t T, int64_t P,
                       double lmb, double gamma, void* stream);

/*
 * One pass producing targets of `mode_targets` and advantages of `mode_adv`
 * (train.py:248-253 calls compute_target twice per head when
 * policy_target != value_target and keeps targets of the first call and
 * advantages of the second).  Same argument meaning as b200rl_target_scan.
 */
int b200rl_target_scan_fused(int mode_targets, int mode_adv, const float* values, int64_t v_sb, int64_t v_st,
                             int64_t v_sp, const float* boot, int64_t boot_sb, int64_t boot_sp,
                             const float* rewards, int64_t r_sb, int64_t r_st, int64_t r_sp,
                             const float* rhos, int64_t rho_sb, int64_t rho_st, int64_t rho_sp,
                             const float* cs, int64_t c_sb, int64_t c_st, int64_t c_sp,
                             float* out_targets, float* out_adv, int64_t B, int64_t T, int64_t P,
                             double lmb, double gamma, void* stream);

/*
 * monte_carlo (handyrl/losses.py:16-17): out_adv = returns - values with
 * broadcasting; the target is `returns` itself (aliased by the Python shim).
 */
int b200rl_target_mc(const float* values, int64_t v_sb, int64_t v_st, int64_t v_sp,
                     const float* returns, int64_t ret_sb, int64_t ret_st, int64_t ret_sp,
                     float* out_adv, int64_t B, int64_t T, int64_t P, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200RL_H */
