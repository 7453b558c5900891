// Time-reversed target recursions (MC / TD(lambda) / UPGO / V-Trace) for sm_100a.
//
// Replaces handyrl/losses.py:16-74.  Layout is the reference's batch-first
// (B, T, P, 1): time is dim 1, so one (b, p) sequence is strided by P in memory
// and the natural "thread per sequence" walk is uncoalesced.  Each CTA therefore
// stages a [rows x TC timesteps x P] tile per input through shared memory with
// coalesced 128-bit loads, every thread then walks ONE sequence backward in time
// inside the tile (carry in registers across tiles), results go back through the
// same shared tile and leave with coalesced 128-bit stores.
//
// Arithmetic follows the reference's evaluation order exactly, one IEEE rounding
// per torch op (no FMA contraction), so results are bit-identical to torch fp32
// on CPU, not merely within 1e-5.
#include "target_common.cuh"
#include <stdlib.h>

namespace {

constexpr int TC = 32;  // timesteps per staged tile

// Stage rows x (tl*px) floats of a strided view into shared memory.
template <int NT>
__device__ __forceinline__ void load_tile(float* __restrict__ s, const View3& x, int rows, int tl, int b0, int t0,
                                          int rs) {
    const int tid = threadIdx.x;
    const int npr = tl * x.px;
    if (x.vec && (npr & 3) == 0) {
        const int n4 = npr >> 2;
        const int total = rows * n4;
        for (int e0 = tid; e0 < total; e0 += 4 * NT) {
            float4 buf[4];
            int rr[4], qq[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                int e = e0 + u * NT;
                if (e < total) {
                    rr[u] = e / n4;
                    qq[u] = e - rr[u] * n4;
                    buf[u] = ld_stream_f4(
                        reinterpret_cast<const float4*>(x.p + (int64_t)(b0 + rr[u]) * x.sb + (int64_t)t0 * x.st) +
                        qq[u]);
                }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                int e = e0 + u * NT;
                if (e < total) {
                    float* d = s + rr[u] * rs + qq[u] * 4;
                    d[0] = buf[u].x;
                    d[1] = buf[u].y;
                    d[2] = buf[u].z;
                    d[3] = buf[u].w;
                }
            }
        }
    } else {
        const int total = rows * npr;
        for (int e = tid; e < total; e += NT) {
            int r = e / npr;
            int j = e - r * npr;
            int tt = j / x.px;
            int p = j - tt * x.px;
            s[r * rs + j] = x.p[(int64_t)(b0 + r) * x.sb + (int64_t)(t0 + tt) * x.st + (int64_t)p * x.sp];
        }
    }
}

// Write rows x (tl*P) floats from shared memory to a contiguous (B,T,P) output.
template <int NT>
__device__ __forceinline__ void store_tile(const float* __restrict__ s, float* __restrict__ out, int rows, int tl,
                                           int b0, int t0, int rs, int T, int P, bool vec) {
    const int tid = threadIdx.x;
    const int npr = tl * P;
    const int64_t row_stride = (int64_t)T * P;
    if (vec && (npr & 3) == 0) {
        const int n4 = npr >> 2;
        const int total = rows * n4;
        for (int e = tid; e < total; e += NT) {
            int r = e / n4;
            int q = e - r * n4;
            const float* d = s + r * rs + q * 4;
            float4 val = make_float4(d[0], d[1], d[2], d[3]);
            st_stream_f4(reinterpret_cast<float4*>(out + (int64_t)(b0 + r) * row_stride + (int64_t)t0 * P) + q, val);
        }
    } else {
        const int total = rows * npr;
        for (int e = tid; e < total; e += NT) {
            int r = e / npr;
            int j = e - r * npr;
            out[(int64_t)(b0 + r) * row_stride + (int64_t)t0 * P + j] = s[r * rs + j];
        }
    }
}

template <int MT, int MA, int NT, bool HAS_REW, bool NEED_RHO>
__global__ void __launch_bounds__(NT) k_target_scan(const ScanParams k) {
    extern __shared__ float smem[];
    const int P = k.P;
    const int rs_v = (TC + 1) * P;
    const int rs_rho = (TC + 1) * k.rho.px;
    const int rs_c = (TC + 1) * k.c.px;
    float* s_v = smem;                    // values in, advantages out
    float* s_r = s_v + k.rows * rs_v;     // rewards in, targets out
    float* s_rho = s_r + k.rows * rs_v;
    float* s_c = s_rho + (NEED_RHO ? k.rows * rs_rho : 0);

    const int b0 = blockIdx.x * k.rows;
    const int rows = min(k.rows, k.B - b0);
    const int sl = threadIdx.x;
    const int r = sl / P;
    const int p = sl - r * P;
    const bool active = r < rows;
    const bool out_vec = ((int64_t)k.T * P) % 4 == 0;

    float R = 0.f;
    if (active) R = k.boot[(int64_t)(b0 + r) * k.boot_sb + (int64_t)p * k.boot_sp];

    Rec<MT> rec_t;
    Rec<MA> rec_a;

    const int nchunks = (k.T + TC - 1) / TC;
    for (int ch = nchunks - 1; ch >= 0; --ch) {
        const int t0 = ch * TC;
        const int tl = min(TC, k.T - t0);
        load_tile<NT>(s_v, k.v, rows, tl, b0, t0, rs_v);
        if (HAS_REW) load_tile<NT>(s_r, k.rew, rows, tl, b0, t0, rs_v);
        if (NEED_RHO) {
            load_tile<NT>(s_rho, k.rho, rows, tl, b0, t0, rs_rho);
            load_tile<NT>(s_c, k.c, rows, tl, b0, t0, rs_c);
        }
        __syncthreads();
        if (active) {
            float* pv = s_v + r * rs_v + p;
            float* pr = s_r + r * rs_v + p;
            const float* prho = s_rho + r * rs_rho + (k.rho.px == 1 ? 0 : p);
            const float* pc = s_c + r * rs_c + (k.c.px == 1 ? 0 : p);
            for (int tt = tl - 1; tt >= 0; --tt) {
                const bool last = (t0 + tt) == k.T - 1;
                float v = pv[tt * P];
                // `rewards is None` -> the reference adds the literal 0 (losses.py:23,35,44)
                float rw = HAS_REW ? pr[tt * P] : 0.f;
                float rho = 0.f, c = 0.f;
                if (NEED_RHO) {
                    rho = prho[tt * k.rho.px];
                    c = pc[tt * k.c.px];
                }
                float tgt, adv, tgt2, adv2;
                rec_step(rec_t, last, v, rw, rho, c, R, k, tgt, adv);
                if (MT != MA) {
                    rec_step(rec_a, last, v, rw, rho, c, R, k, tgt2, adv2);
                    adv = adv2;
                }
                pv[tt * P] = adv;
                pr[tt * P] = tgt;
            }
        }
        __syncthreads();
        store_tile<NT>(s_r, k.out_t, rows, tl, b0, t0, rs_v, k.T, P, out_vec);
        store_tile<NT>(s_v, k.out_a, rows, tl, b0, t0, rs_v, k.T, P, out_vec);
        __syncthreads();
    }
}

__global__ void k_target_mc(const float* __restrict__ v, int64_t v_sb, int64_t v_st, int64_t v_sp,
                            const float* __restrict__ ret, int64_t ret_sb, int64_t ret_st, int64_t ret_sp,
                            float* __restrict__ out, int64_t n, int T, int P) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        int p = (int)(i % P);
        int64_t bt = i / P;
        int t = (int)(bt % T);
        int64_t b = bt / T;
        out[i] = __fsub_rn(ret[b * ret_sb + t * ret_st + p * ret_sp], v[b * v_sb + t * v_st + p * v_sp]);
    }
}

// monte_carlo for contiguous values rows (T*P % 4 == 0): one float4 per thread; `returns` is either
// contiguous like values (RET_FULL) or constant over time, (B,1,P) (the outcome, train.py:245).
template <bool RET_FULL>
__global__ void __launch_bounds__(256)
k_target_mc_rows(const float4* __restrict__ v, int64_t v_sb4, const float* __restrict__ ret, int64_t ret_sb,
                 int64_t ret_sp, float4* __restrict__ out, int64_t n4, int row4, int P) {
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t b = i / row4;
        const int j4 = (int)(i - b * row4);
        const float4 a = ld_stream_f4(v + b * v_sb4 + j4);
        float4 r;
        if (RET_FULL) {
            r = ld_stream_f4(reinterpret_cast<const float4*>(ret + b * ret_sb) + j4);
        } else {
            const float* rb = ret + b * ret_sb;
            const int p0 = (j4 * 4) % P;
            int p1 = p0 + 1; if (p1 >= P) p1 -= P;
            int p2 = p1 + 1; if (p2 >= P) p2 -= P;
            int p3 = p2 + 1; if (p3 >= P) p3 -= P;
            r = make_float4(rb[p0 * ret_sp], rb[p1 * ret_sp], rb[p2 * ret_sp], rb[p3 * ret_sp]);
        }
        st_stream_f4(out + i, make_float4(__fsub_rn(r.x, a.x), __fsub_rn(r.y, a.y), __fsub_rn(r.z, a.z),
                                          __fsub_rn(r.w, a.w)));
    }
}

View3 make_view(const float* p, int64_t sb, int64_t st, int64_t sp, int P) {
    View3 x;
    x.p = p;
    x.sb = sb;
    x.st = st;
    x.sp = sp;
    x.px = (sp == 0 && P > 1) ? 1 : P;
    bool contig = (st == x.px) && (x.px == 1 || sp == 1);
    x.vec = p != nullptr && contig && (sb % 4 == 0) && ((uintptr_t)p % 16 == 0);
    return x;
}

template <int MT, int MA, int NT>
int launch_cfg(const ScanParams& k, bool has_rew, cudaStream_t st) {
    const bool need_rho = (MT == B200RL_TARGET_VTRACE || MA == B200RL_TARGET_VTRACE);
    size_t fl = (size_t)2 * k.rows * (TC + 1) * k.P;
    if (need_rho) fl += (size_t)k.rows * (TC + 1) * (k.rho.px + k.c.px);
    size_t smem = fl * sizeof(float);
    int grid = (k.B + k.rows - 1) / k.rows;
#define LAUNCH(HR, NR)                                                                                    \
    do {                                                                                                  \
        auto kern = k_target_scan<MT, MA, NT, HR, NR>;                                                    \
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);               \
        kern<<<grid, NT, smem, st>>>(k);                                                                  \
    } while (0)
    if (need_rho) {
        if (has_rew) LAUNCH(true, true); else LAUNCH(false, true);
    } else {
        if (has_rew) LAUNCH(true, false); else LAUNCH(false, false);
    }
#undef LAUNCH
    return b200rl_check_launch("k_target_scan");
}

template <int MT, int MA>
int launch_nt(ScanParams& k, bool has_rew, cudaStream_t st) {
    // Pick the CTA width so that small batches still spread over the 148 SMs.
    int64_t seqs = (int64_t)k.B * k.P;
    int nt = 128;
    if (seqs < (int64_t)148 * 2 * 128) nt = 64;
    if (seqs < (int64_t)148 * 2 * 64) nt = 32;
    if (nt < k.P) nt = 128;
    k.rows = nt / k.P;
    if (nt == 128) return launch_cfg<MT, MA, 128>(k, has_rew, st);
    if (nt == 64) return launch_cfg<MT, MA, 64>(k, has_rew, st);
    return launch_cfg<MT, MA, 32>(k, has_rew, st);
}

int dispatch(int mt, int ma, ScanParams& k, bool has_rew, cudaStream_t st) {
#define CASE(A, B_) \
    if (mt == A && ma == B_) return launch_nt<A, B_>(k, has_rew, st);
    CASE(B200RL_TARGET_TD, B200RL_TARGET_TD)
    CASE(B200RL_TARGET_UPGO, B200RL_TARGET_UPGO)
    CASE(B200RL_TARGET_VTRACE, B200RL_TARGET_VTRACE)
    CASE(B200RL_TARGET_TD, B200RL_TARGET_UPGO)
    CASE(B200RL_TARGET_TD, B200RL_TARGET_VTRACE)
    CASE(B200RL_TARGET_UPGO, B200RL_TARGET_TD)
    CASE(B200RL_TARGET_UPGO, B200RL_TARGET_VTRACE)
    CASE(B200RL_TARGET_VTRACE, B200RL_TARGET_TD)
    CASE(B200RL_TARGET_VTRACE, B200RL_TARGET_UPGO)
#undef CASE
    b200rl_set_error("target_scan: unsupported mode pair (%d, %d)", mt, ma);
    return B200RL_ERR_UNSUPPORTED;
}

}  // namespace

namespace {

ScanParams make_params(const float* values, int64_t v_sb, int64_t v_st, int64_t v_sp, const float* boot,
                       int64_t boot_sb, int64_t boot_sp, const float* rewards, int64_t r_sb, int64_t r_st,
                       int64_t r_sp, const float* rhos, int64_t rho_sb, int64_t rho_st, int64_t rho_sp,
                       const float* cs, int64_t c_sb, int64_t c_st, int64_t c_sp, float* out_targets, float* out_adv,
                       int64_t B, int64_t T, int64_t P, double lmb, double gamma) {
    ScanParams k;
    k.v = make_view(values, v_sb, v_st, v_sp, (int)P);
    k.v.px = (int)P;  // values always carry the player dim
    k.v.vec = k.v.vec && v_st == P && (P == 1 || v_sp == 1);
    k.rew = make_view(rewards, r_sb, r_st, r_sp, (int)P);
    k.rew.px = (int)P;  // the reward tile is reused for the targets, keep the player dim
    k.rew.vec = k.rew.vec && r_st == P && (P == 1 || r_sp == 1);
    k.rho = make_view(rhos, rho_sb, rho_st, rho_sp, (int)P);
    k.c = make_view(cs, c_sb, c_st, c_sp, (int)P);
    k.boot = boot;
    k.boot_sb = boot_sb;
    k.boot_sp = boot_sp;
    k.out_t = out_targets;
    k.out_a = out_adv;
    k.B = (int)B;
    k.T = (int)T;
    k.P = (int)P;
    k.rows = 0;
    // torch multiplies an fp32 tensor by a Python double by rounding the scalar to
    // fp32 once; (1 - lmb) and gamma * lmb are formed in double first (losses.py:24,36,51).
    k.oml = (float)(1.0 - lmb);
    k.lmb = (float)lmb;
    k.gamma = (float)gamma;
    k.gl = (float)(gamma * lmb);
    return k;
}

int launch_one(int mode_targets, int mode_adv, ScanParams& k, cudaStream_t stream) {
    // contiguous P<=2 layouts (what compute_loss produces): row-per-thread cp.async pipeline
    const bool force_generic = b200rl_tuning(B200RL_TUNE_TARGET_GENERIC, nullptr);  // A/B switch for tools/sweep_targets.py
    const bool has_rew = k.rew.p != nullptr;
    const int fast = force_generic ? -1 : b200rl_launch_target_rows(mode_targets, mode_adv, k, has_rew, stream);
    if (fast >= 0) return fast;
    return dispatch(mode_targets, mode_adv, k, has_rew, stream);
}

}  // namespace

extern "C" int b200rl_target_scan_fused(int mode_targets, int mode_adv, const float* values, int64_t v_sb,
                                        int64_t v_st, int64_t v_sp, const float* boot, int64_t boot_sb,
                                        int64_t boot_sp, const float* rewards, int64_t r_sb, int64_t r_st,
                                        int64_t r_sp, const float* rhos, int64_t rho_sb, int64_t rho_st,
                                        int64_t rho_sp, const float* cs, int64_t c_sb, int64_t c_st, int64_t c_sp,
                                        float* out_targets, float* out_adv, int64_t B, int64_t T, int64_t P,
                                        double lmb, double gamma, void* stream) {
    B200RL_REQUIRE(values && boot && out_targets && out_adv, "target_scan: null pointer");
    B200RL_REQUIRE(((uintptr_t)out_targets % 16 == 0) && ((uintptr_t)out_adv % 16 == 0), "target_scan: outputs must be 16B aligned");
    B200RL_REQUIRE(B >= 0 && T >= 1 && P >= 1 && P <= 128, "target_scan: bad shape B=%lld T=%lld P=%lld",
                   (long long)B, (long long)T, (long long)P);
    B200RL_REQUIRE(B < (1ll << 31) && T < (1ll << 31), "target_scan: shape too large");
    const bool need_rho = mode_targets == B200RL_TARGET_VTRACE || mode_adv == B200RL_TARGET_VTRACE;
    B200RL_REQUIRE(!need_rho || (rhos && cs), "target_scan: VTRACE needs rhos and cs");
    if (B == 0) return B200RL_OK;
    ScanParams k = make_params(values, v_sb, v_st, v_sp, boot, boot_sb, boot_sp, rewards, r_sb, r_st, r_sp, rhos, rho_sb,
                               rho_st, rho_sp, cs, c_sb, c_st, c_sp, out_targets, out_adv, B, T, P, lmb, gamma);
    return launch_one(mode_targets, mode_adv, k, (cudaStream_t)stream);
}

extern "C" int b200rl_target_scan_heads(int mode_targets, int mode_adv, const b200rl_target_head* heads, int n_heads,
                                        int64_t B, int64_t T, int64_t P, double lmb, void* stream) {
    B200RL_REQUIRE(heads && n_heads >= 1 && n_heads <= 2, "target_scan_heads: 1 or 2 heads");
    B200RL_REQUIRE(B >= 0 && T >= 1 && P >= 1 && P <= 128, "target_scan_heads: bad shape B=%lld T=%lld P=%lld",
                   (long long)B, (long long)T, (long long)P);
    B200RL_REQUIRE(B < (1ll << 31) && T < (1ll << 31), "target_scan_heads: shape too large");
    const bool need_rho = mode_targets == B200RL_TARGET_VTRACE || mode_adv == B200RL_TARGET_VTRACE;
    ScanParams k[2];
    for (int h = 0; h < n_heads; ++h) {
        const b200rl_target_head& x = heads[h];
        B200RL_REQUIRE(x.values && x.boot && x.out_targets && x.out_adv, "target_scan_heads: null pointer in head %d", h);
        B200RL_REQUIRE(((uintptr_t)x.out_targets % 16 == 0) && ((uintptr_t)x.out_adv % 16 == 0),
                       "target_scan_heads: outputs must be 16B aligned");
        B200RL_REQUIRE(!need_rho || (x.rhos && x.cs), "target_scan_heads: VTRACE needs rhos and cs");
        k[h] = make_params(x.values, x.v_sb, x.v_st, x.v_sp, x.boot, x.boot_sb, x.boot_sp, x.rewards, x.r_sb, x.r_st,
                           x.r_sp, x.rhos, x.rho_sb, x.rho_st, x.rho_sp, x.cs, x.c_sb, x.c_st, x.c_sp, x.out_targets,
                           x.out_adv, B, T, P, lmb, x.gamma);
    }
    if (B == 0) return B200RL_OK;
    const bool no_dual = b200rl_tuning(B200RL_TUNE_TARGET_GENERIC, nullptr) || b200rl_tuning(B200RL_TUNE_TARGET_NO_DUAL, nullptr);
    if (n_heads == 2 && !no_dual) {
        const int rc = b200rl_launch_target_rows2(mode_targets, mode_adv, k[0], k[1], (cudaStream_t)stream);
        if (rc >= 0) return rc;
    }
    for (int h = 0; h < n_heads; ++h) {
        const int rc = launch_one(mode_targets, mode_adv, k[h], (cudaStream_t)stream);
        if (rc != B200RL_OK) return rc;
    }
    return B200RL_OK;
}

extern "C" int b200rl_target_scan(int mode, const float* values, int64_t v_sb, int64_t v_st, int64_t v_sp,
                                  const float* boot, int64_t boot_sb, int64_t boot_sp, const float* rewards,
                                  int64_t r_sb, int64_t r_st, int64_t r_sp, const float* rhos, int64_t rho_sb,
                                  int64_t rho_st, int64_t rho_sp, const float* cs, int64_t c_sb, int64_t c_st,
                                  int64_t c_sp, float* out_targets, float* out_adv, int64_t B, int64_t T,
                                  int64_t P, double lmb, double gamma, void* stream) {
    return b200rl_target_scan_fused(mode, mode, values, v_sb, v_st, v_sp, boot, boot_sb, boot_sp, rewards, r_sb,
                                    r_st, r_sp, rhos, rho_sb, rho_st, rho_sp, cs, c_sb, c_st, c_sp, out_targets,
                                    out_adv, B, T, P, lmb, gamma, stream);
}

extern "C" int b200rl_target_mc(const float* values, int64_t v_sb, int64_t v_st, int64_t v_sp,
                                const float* returns, int64_t ret_sb, int64_t ret_st, int64_t ret_sp,
                                float* out_adv, int64_t B, int64_t T, int64_t P, void* stream) {
    B200RL_REQUIRE(values && returns && out_adv, "target_mc: null pointer");
    int64_t n = B * T * P;
    if (n == 0) return B200RL_OK;
    const int64_t TP = T * P;
    const bool v_rows = (T == 1 || v_st == P) && (P == 1 || v_sp == 1) && TP % 4 == 0 && v_sb % 4 == 0 &&
                        ((uintptr_t)values % 16) == 0 && ((uintptr_t)out_adv % 16) == 0 && P <= 1024;
    const bool ret_full = (T == 1 || ret_st == P) && (P == 1 || ret_sp == 1) && ret_sb % 4 == 0 &&
                          ((uintptr_t)returns % 16) == 0;
    const bool ret_const = ret_st == 0;
    if (v_rows && (ret_full || ret_const)) {
        const int64_t n4 = n / 4;
        int64_t blocks = (n4 + 255) / 256;
        if (blocks > 148 * 16) blocks = 148 * 16;
        if (ret_full)
            k_target_mc_rows<true><<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(
                (const float4*)values, v_sb / 4, returns, ret_sb, ret_sp, (float4*)out_adv, n4, (int)(TP / 4), (int)P);
        else
            k_target_mc_rows<false><<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(
                (const float4*)values, v_sb / 4, returns, ret_sb, ret_sp, (float4*)out_adv, n4, (int)(TP / 4), (int)P);
        return b200rl_check_launch("k_target_mc_rows");
    }
    int64_t blocks = (n + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    k_target_mc<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(values, v_sb, v_st, v_sp, returns, ret_sb, ret_st,
                                                               ret_sp, out_adv, n, (int)T, (int)P);
    return b200rl_check_launch("k_target_mc");
}
