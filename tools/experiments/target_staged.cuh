// NOT PART OF THE BUILD.  An experiment kept for the record (profiles/r2_target_ws_rowlane.md section 3): bit-exact, but at
// best break-even with k_target_rows_ws.  Commit b78b3bd has it wired into handyrl_b200/csrc/target_rows.cu (launcher
// `launch_staged`, knob target_ws_staged, 60 long-row test cases).
//
// Staged pipeline for the target recursion when tiles are FEW (small B, long rows) - included by target_rows.cu
// inside its anonymous namespace (it uses that file's cp.async / mbarrier / LaneIO helpers).
//
// At B = 4,096 every tile of the warp-specialised kernel runs concurrently, so the kernel lasts T times what ONE compute
// warp needs per timestep: 44 cycles for TD, 64 for V-Trace (profiles/r2_target_staged.md) - 12-15 instructions, every
// fp32 one holding its pipe for 2 cycles, 4 scalar LDS / STS of 3-4 issue cycles each, and the dependent chain on top.
// But only a small part of each recurrence is sequential:
//     TD      G[t]   = r[t] + gamma * ((1-lmb) * v[t+1] + lmb * G[t+1])              4 dependent fp32 ops
//     UPGO    G[t]   = r[t] + gamma * max(v[t+1], (1-lmb) * v[t+1] + lmb * G[t+1])   5
//     V-Trace acc[t] = delta[t] + (gamma*lmb*c[t]) * acc[t+1]                         2
// while delta[t] = rho[t] * ((r[t] + gamma * v[t+1]) - v[t]), gamma*lmb*c[t], vs = acc + v, every advantage and the
// (t, player) de-interleaving are ELEMENTWISE: any lane can do them for any t.  So a CTA is four specialised warps around
// a ring of slots (one chunk of RTC timesteps of 32/P batch rows each), four mbarriers per slot:
//     warp 0  loader   16-byte cp.async of the raw chunk (global layout [row][t][p]); arrives on `full` when landed
//     warp 1  pre      lane = 16-byte unit of a row: v[t+1] by shuffle / carry, delta, gamma*lmb*c; writes the scan
//                      inputs DE-INTERLEAVED, [sequence][t], so that ...                                   -> `ready`
//     warp 2  scan     lane = (row, player): one LDS.128 per input and 4 timesteps, only the dependent chain above,
//                      one STS.128 per output and 4 timesteps                                             -> `done`
//     warp 3  post     lane = 16-byte unit: targets / advantages from the scan output and the raw chunk, written
//                      to global memory with coalesced 128-bit streaming stores                          -> `empty`
// Every fp32 operation of the reference's evaluation order is kept, per element, with the same operands (Rec<> of
// target_common.cuh spelled out stage by stage), so results are bit-identical to the other kernels and to torch.
#pragma once

constexpr int ST_BAR_ONLY = 4 * WS_MAX_STAGES * 2;     // 4 x 8 mbarriers of 8 bytes, in floats
// + chunk-boundary carries of the pre and the post warp: [warp][buffer][sequence]; buffer 0 starts as the bootstrap R
constexpr int ST_BAR_FLOATS = ST_BAR_ONLY + 2 * 2 * 32;
constexpr int ST_TD = B200RL_TARGET_TD, ST_UPGO = B200RL_TARGET_UPGO, ST_VTRACE = B200RL_TARGET_VTRACE;

// shared-memory layout of one slot, in floats (used by the kernel and by its launcher)
template <int MT, int MA, int P, bool HAS_REW, bool RHO_PP, int RTC>
struct StagedLayout {
    static constexpr bool VT_T = MT == ST_VTRACE, VT_A = MA == ST_VTRACE;
    static constexpr bool NEED_RHO = VT_T || VT_A;
    static constexpr bool NEED_G = !VT_T || !VT_A;   // a TD / UPGO recurrence runs
    static constexpr bool TWO = MT != MA;             // two recurrences per sequence
    static constexpr int TR = 32 / P;                 // batch rows per tile: one scan lane per (row, player)
    static constexpr int WR = RHO_PP ? P : 1;
    static constexpr int RSV = RTC * P + 4;           // raw row strides: odd numbers of 16-byte units
    static constexpr int RSR = RTC * WR + 4;
    static constexpr int XS = RTC + 4;                // stride of a de-interleaved sequence
    static constexpr int UPR = RTC * P / 4;           // 16-byte units per raw row
    static constexpr int RPP = 32 / UPR;              // rows a warp covers per pass in pre / post
    static_assert(UPR == 32 || UPR == 16, "a raw row is one or half a warp of 16-byte units");
    static constexpr int OFF_V = 0;
    static constexpr int OFF_R = TR * RSV;
    static constexpr int OFF_RHO = TR * RSV * (HAS_REW ? 2 : 1);
    // X: scan inputs [v1][r] (TD / UPGO), [delta][g] (V-Trace); Y: scan outputs of MT (and of MA when it differs)
    static constexpr int NXG = NEED_G ? (HAS_REW ? 2 : 1) : 0;
    static constexpr int NX = NXG + (NEED_RHO ? 2 : 0);
    static constexpr int NY = TWO ? 2 : 1;
    // the area of `cs` is reserved even when cs aliases rhos: every offset is then a compile-time constant, which
    // lets the compiler see that a pass's stores (X) never alias the next pass's loads (raw) and overlap them
    static constexpr int OFF_X = OFF_RHO + (NEED_RHO ? 2 * TR * RSR : 0);
    static constexpr int OFF_Y = OFF_X + NX * 32 * XS;
    static constexpr int STAGE = OFF_Y + NY * 32 * XS;
    __host__ __device__ static constexpr int off_x(bool) { return OFF_X; }
    __host__ __device__ static constexpr int off_y(bool) { return OFF_Y; }
    __host__ __device__ static constexpr int stage(bool) { return STAGE; }
};

// One timestep of a recurrence in the scan warp.  G-modes: x0 = v[t+1], x1 = r[t]; V-Trace: x0 = delta[t], x1 = g[t].
template <int MODE>
__device__ __forceinline__ float staged_step(float state, float x0, float x1, const ScanParams& k) {
    if (MODE == ST_VTRACE) return __fadd_rn(x0, __fmul_rn(x1, state));                 // losses.py:51
    const float a = __fmul_rn(k.oml, x0);
    const float c = __fadd_rn(a, __fmul_rn(k.lmb, state));
    if (MODE == ST_UPGO) return __fadd_rn(x1, __fmul_rn(k.gamma, max_nanprop(x0, c)));  // losses.py:36
    return __fadd_rn(x1, __fmul_rn(k.gamma, c));                                       // losses.py:24
}

// ---------------- pre warp: elementwise inputs of the scan, de-interleaved ----------------
// lane = (row within a pass, 16-byte unit of the raw row).  Element e of a unit: P = 2 -> (timestep e >> 1, player
// e & 1); P = 1 -> (timestep e, player 0); its successor in time is element e + P of the same unit, else element
// e + P - 4 of the NEXT unit (lane + 1), else - last unit of the chunk - the row's carry: v at the first timestep of
// the chunk processed before, kept in a double-buffered shared array whose first buffer starts as the bootstrap R
// (= v[T] for the V-Trace delta at T-1, losses.py:45).
template <int MT, int MA, int P, bool HAS_REW, bool RHO_PP, int RTC>
__device__ __forceinline__ void staged_pre_warp(const ScanParams& k, float* ring, float* carry, bool alias, uint64_t* full, uint64_t* ready, int ns, int n, int b0,
                                                int rows, int lane) {
    using L = StagedLayout<MT, MA, P, HAS_REW, RHO_PP, RTC>;
    constexpr bool NEED_RHO = L::NEED_RHO, NEED_G = L::NEED_G;
    constexpr int TR = L::TR, RSV = L::RSV, RSR = L::RSR, XS = L::XS, UPR = L::UPR, RPP = L::RPP;
    constexpr int UPT = 4 / P, STAGE = L::STAGE, OFF_X = L::OFF_X;
    const int rr = lane / UPR, u = lane % UPR;
    {
        const int r = lane / P, p = lane % P;
        carry[lane] = r < rows ? k.boot[(int64_t)(b0 + r) * k.boot_sb + (int64_t)p * k.boot_sp] : 0.f;
        __syncwarp();
    }
    int slot = 0;
    uint32_t par = 0;
    for (int i = 0; i < n; ++i) {
        float* st = ring + slot * STAGE;
        const int tl = min(RTC, k.T - (n - 1 - i) * RTC);
        const int upr = tl / UPT;                  // units per row in this chunk
        const float* cin = carry + (i & 1) * 32;
        float* cout = carry + ((i + 1) & 1) * 32;
        mbar_wait(full + slot, par);
        // per-lane bases; a pass advances them by RPP rows
        const float* pv = st + L::OFF_V + rr * RSV + 4 * u;
        const float* prw = st + L::OFF_R + rr * RSV + 4 * u;
        const float* prho = st + L::OFF_RHO + rr * RSR + (RHO_PP ? 4 * u : UPT * u);
        float* px = st + OFF_X + rr * P * XS + UPT * u;
        struct In { float v[4], rw[4], rho[4], cc[4], cr[P]; };
        auto load = [&](int r0, In& in) {
            const int row = r0 + rr;
            const bool on = row < rows && u < upr;
#pragma unroll
            for (int e = 0; e < 4; ++e) in.v[e] = in.rw[e] = in.rho[e] = in.cc[e] = 0.f;
            if (on) {
                const float4 a = *reinterpret_cast<const float4*>(pv + r0 * RSV);
                in.v[0] = a.x; in.v[1] = a.y; in.v[2] = a.z; in.v[3] = a.w;
                if (HAS_REW) {
                    const float4 b = *reinterpret_cast<const float4*>(prw + r0 * RSV);
                    in.rw[0] = b.x; in.rw[1] = b.y; in.rw[2] = b.z; in.rw[3] = b.w;
                }
                if (NEED_RHO) {
                    if (RHO_PP || P == 1) {   // (t, p) interleaved like the values / 4 timesteps of the one player
                        const float4 a4 = *reinterpret_cast<const float4*>(prho + r0 * RSR);
                        in.rho[0] = a4.x; in.rho[1] = a4.y; in.rho[2] = a4.z; in.rho[3] = a4.w;
                        const float4 b4 = alias ? a4 : *reinterpret_cast<const float4*>(prho + TR * RSR + r0 * RSR);
                        in.cc[0] = b4.x; in.cc[1] = b4.y; in.cc[2] = b4.z; in.cc[3] = b4.w;
                    } else {                   // one value per timestep, shared by the 2 players: 2 timesteps per unit
                        const float2 a2 = *reinterpret_cast<const float2*>(prho + r0 * RSR);
                        in.rho[0] = in.rho[1] = a2.x; in.rho[2] = in.rho[3] = a2.y;
                        const float2 b2 = alias ? a2 : *reinterpret_cast<const float2*>(prho + TR * RSR + r0 * RSR);
                        in.cc[0] = in.cc[1] = b2.x; in.cc[2] = in.cc[3] = b2.y;
                    }
                }
            }
#pragma unroll
            for (int p = 0; p < P; ++p) in.cr[p] = (on && u == upr - 1) ? cin[row * P + p] : 0.f;
        };
        auto finish = [&](int r0, const In& in) {
            const int row = r0 + rr;
            const bool on = row < rows && u < upr;
            float vn[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                if (e + P < 4) {
                    vn[e] = in.v[e + P];
                } else {
                    const float nb = __shfl_down_sync(0xffffffffu, in.v[e + P - 4], 1);
                    vn[e] = u == upr - 1 ? in.cr[P == 2 ? (e & 1) : 0] : nb;
                }
            }
            float dl[4], g[4];
            if (NEED_RHO) {
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const float r_ = HAS_REW ? in.rw[e] : 0.f;   // `rewards is None`: the literal 0 (losses.py:44)
                    dl[e] = __fmul_rn(in.rho[e], __fsub_rn(__fadd_rn(r_, __fmul_rn(k.gamma, vn[e])), in.v[e]));  // losses.py:46
                    g[e] = __fmul_rn(k.gl, in.cc[e]);                                                               // losses.py:51
                }
            }
            if (on) {
                // sequence of element e: row * P + player; its local timestep: UPT * u + e / P
                auto put = [&](float* arr, const float (&val)[4]) {
                    if (P == 2) {
                        *reinterpret_cast<float2*>(arr + (r0 * 2 + 0) * XS) = make_float2(val[0], val[2]);
                        *reinterpret_cast<float2*>(arr + (r0 * 2 + 1) * XS) = make_float2(val[1], val[3]);
                    } else {
                        *reinterpret_cast<float4*>(arr + r0 * XS) = make_float4(val[0], val[1], val[2], val[3]);
                    }
                };
                if (NEED_G) {
                    put(px, vn);
                    if (HAS_REW) put(px + 32 * XS, in.rw);
                }
                if (NEED_RHO) {
                    put(px + L::NXG * 32 * XS, dl);
                    put(px + (L::NXG + 1) * 32 * XS, g);
                }
                if (u == 0) {   // carry for the next chunk: this chunk's first timestep
#pragma unroll
                    for (int p = 0; p < P; ++p) cout[row * P + p] = in.v[p];
                }
            }
        };
        if (rows == TR) {   // full tile: the loads of NB passes are issued before the first of them is finished
            constexpr int NPASS = TR / RPP, NB = NPASS % 4 == 0 ? 4 : (NPASS % 2 == 0 ? 2 : 1);
#pragma unroll
            for (int b = 0; b < NPASS; b += NB) {
                In in[NB];
#pragma unroll
                for (int j = 0; j < NB; ++j) load((b + j) * RPP, in[j]);
#pragma unroll
                for (int j = 0; j < NB; ++j) finish((b + j) * RPP, in[j]);
            }
        } else {
            for (int r0 = 0; r0 < rows; r0 += RPP) {
                In in;
                load(r0, in);
                finish(r0, in);
            }
        }
        __syncwarp();   // carries written by the lanes u == 0 are read by the lanes u == upr-1 of the next chunk
        mbar_arrive(ready + slot);
        if (++slot == ns) {
            slot = 0;
            par ^= 1;
        }
    }
}

// ---------------- scan warp: lane = (row, player), only the dependent chains ----------------
template <int MT, int MA, int P, bool HAS_REW, bool RHO_PP, int RTC>
__device__ __forceinline__ void staged_scan_warp(const ScanParams& k, float* ring, uint64_t* ready, uint64_t* done, int ns, int n, int b0, int rows,
                                                 int lane) {
    using L = StagedLayout<MT, MA, P, HAS_REW, RHO_PP, RTC>;
    constexpr bool VT_T = L::VT_T, VT_A = L::VT_A, TWO = L::TWO;
    constexpr int XS = L::XS, STAGE = L::STAGE, OFF_X = L::OFF_X, OFF_Y = L::OFF_Y;
    // which X arrays a recurrence reads: G-modes (v1, r), V-Trace (delta, g)
    constexpr int XT0 = VT_T ? L::NXG : 0, XA0 = VT_A ? L::NXG : 0;
    const int r = lane / P, p = lane % P;
    const bool active = r < rows;
    const float R = active ? k.boot[(int64_t)(b0 + r) * k.boot_sb + (int64_t)p * k.boot_sp] : 0.f;
    float st_t = 0.f, st_a = 0.f;   // G or acc of the two recurrences
    struct Grp { float4 t0, t1, a0, a1; };
    int slot = 0;
    uint32_t par = 0;
    for (int i = 0; i < n; ++i) {
        float* st = ring + slot * STAGE;
        const int tl = min(RTC, k.T - (n - 1 - i) * RTC);
        mbar_wait(ready + slot, par);
        if (active) {
            const float* x = st + OFF_X + lane * XS;
            float* y = st + OFF_Y + lane * XS;
            auto load_group = [&](Grp& g, int q) {
                g.t0 = *reinterpret_cast<const float4*>(x + XT0 * 32 * XS + 4 * q);
                if (VT_T || HAS_REW) g.t1 = *reinterpret_cast<const float4*>(x + (XT0 + 1) * 32 * XS + 4 * q);
                if (TWO && XA0 != XT0) {   // the other recurrence reads other arrays (one is V-Trace, one a G-mode)
                    g.a0 = *reinterpret_cast<const float4*>(x + XA0 * 32 * XS + 4 * q);
                    if (VT_A || HAS_REW) g.a1 = *reinterpret_cast<const float4*>(x + (XA0 + 1) * 32 * XS + 4 * q);
                }
            };
            auto compute_group = [&](const Grp& g, int q, auto first_tag) {
                constexpr bool FIRST = decltype(first_tag)::value;
                const float t0[4] = {g.t0.x, g.t0.y, g.t0.z, g.t0.w};
                float t1[4] = {0.f, 0.f, 0.f, 0.f};   // rewards = None: the literal 0 (losses.py:23,35)
                if (VT_T || HAS_REW) { t1[0] = g.t1.x; t1[1] = g.t1.y; t1[2] = g.t1.z; t1[3] = g.t1.w; }
                float a0[4], a1[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) { a0[j] = t0[j]; a1[j] = t1[j]; }
                if (TWO && XA0 != XT0) {
                    a0[0] = g.a0.x; a0[1] = g.a0.y; a0[2] = g.a0.z; a0[3] = g.a0.w;
                    a1[0] = a1[1] = a1[2] = a1[3] = 0.f;
                    if (VT_A || HAS_REW) { a1[0] = g.a1.x; a1[1] = g.a1.y; a1[2] = g.a1.z; a1[3] = g.a1.w; }
                }
                float yt[4], ya[4];
#pragma unroll
                for (int j = 3; j >= 0; --j) {
                    if (FIRST && j == 3) {   // t = T-1: G = R (losses.py:21,32), acc = delta (losses.py:49)
                        st_t = VT_T ? t0[j] : R;
                        if (TWO) st_a = VT_A ? a0[j] : R;
                    } else {
                        st_t = staged_step<MT>(st_t, t0[j], t1[j], k);
                        if (TWO) st_a = staged_step<MA>(st_a, a0[j], a1[j], k);
                    }
                    yt[j] = st_t;
                    ya[j] = st_a;
                }
                *reinterpret_cast<float4*>(y + 4 * q) = make_float4(yt[0], yt[1], yt[2], yt[3]);
                if (TWO) *reinterpret_cast<float4*>(y + 32 * XS + 4 * q) = make_float4(ya[0], ya[1], ya[2], ya[3]);
            };
            // two register sets (ping-pong): the loads of group q-1 are issued before the chain of group q
            Grp ga, gb;
            int q = (tl >> 2) - 1;
            load_group(ga, q);
            if (i == 0) {   // warp-uniform: the chunk ends with the bootstrap step
                if (q >= 1) load_group(gb, q - 1);
                compute_group(ga, q, std::true_type{});
                --q;
                if (q >= 0) ga = gb;
            }
            for (; q >= 1; q -= 2) {
                load_group(gb, q - 1);
                compute_group(ga, q, std::false_type{});
                if (q >= 2) load_group(ga, q - 2);
                compute_group(gb, q - 1, std::false_type{});
            }
            if (q == 0) compute_group(ga, 0, std::false_type{});
        }
        mbar_arrive(done + slot);
        if (++slot == ns) {
            slot = 0;
            par ^= 1;
        }
    }
}

// ---------------- post warp: lane = 16-byte unit of a row; targets / advantages -> global memory ----------------
// The V-Trace advantage needs vs[t+1] (losses.py:55-56): inside the unit, from the next unit (lane + 1), or the row's
// carry = vs at the first timestep of the chunk processed before (double-buffered like the pre warp's; starts as R).
template <int MT, int MA, int P, bool HAS_REW, bool RHO_PP, int RTC>
__device__ __forceinline__ void staged_post_warp(const ScanParams& k, float* ring, float* carry, uint64_t* done, uint64_t* empty, int ns, int n, int b0, int rows,
                                                 int lane) {
    using L = StagedLayout<MT, MA, P, HAS_REW, RHO_PP, RTC>;
    constexpr bool VT_T = L::VT_T, VT_A = L::VT_A, TWO = L::TWO;
    constexpr int TR = L::TR, RSV = L::RSV, XS = L::XS, UPR = L::UPR, RPP = L::RPP;
    constexpr int UPT = 4 / P, STAGE = L::STAGE, OFF_Y = L::OFF_Y;
    const int rr = lane / UPR, u = lane % UPR;
    const int64_t row_out = (int64_t)k.T * P;
    if (VT_A) {
        const int r = lane / P, p = lane % P;
        carry[lane] = r < rows ? k.boot[(int64_t)(b0 + r) * k.boot_sb + (int64_t)p * k.boot_sp] : 0.f;
        __syncwarp();
    }
    int slot = 0;
    uint32_t par = 0;
    for (int i = 0; i < n; ++i) {
        const float* st = ring + slot * STAGE;
        const int t0 = (n - 1 - i) * RTC, tl = min(RTC, k.T - t0);
        const int upr = tl / UPT;
        const float* cin = carry + (i & 1) * 32;
        float* cout = carry + ((i + 1) & 1) * 32;
        mbar_wait(done + slot, par);
        const float* pv = st + L::OFF_V + rr * RSV + 4 * u;
        const float* prw = st + L::OFF_R + rr * RSV + 4 * u;
        const float* py = st + OFF_Y + rr * P * XS + UPT * u;
        const int64_t o0 = (int64_t)(b0 + rr) * row_out + (int64_t)t0 * P + 4 * u;
        float* gt = k.out_t + o0;
        float* ga = k.out_a + o0;
        auto pass = [&](int r0) {
            const int row = r0 + rr;
            const bool on = row < rows && u < upr;
            float v[4] = {0.f, 0.f, 0.f, 0.f}, rw[4] = {0.f, 0.f, 0.f, 0.f}, yt[4] = {0.f, 0.f, 0.f, 0.f};
            float ya[4] = {0.f, 0.f, 0.f, 0.f}, cr[P];
            if (on) {
                const float4 a = *reinterpret_cast<const float4*>(pv + r0 * RSV);
                v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
                if (HAS_REW && VT_A) {   // rewards enter the V-Trace advantage only (losses.py:56)
                    const float4 b = *reinterpret_cast<const float4*>(prw + r0 * RSV);
                    rw[0] = b.x; rw[1] = b.y; rw[2] = b.z; rw[3] = b.w;
                }
                auto get = [&](const float* arr, float (&val)[4]) {   // inverse of the pre warp's `put`
                    if (P == 2) {
                        const float2 a0 = *reinterpret_cast<const float2*>(arr + (r0 * 2 + 0) * XS);
                        const float2 a1 = *reinterpret_cast<const float2*>(arr + (r0 * 2 + 1) * XS);
                        val[0] = a0.x; val[2] = a0.y; val[1] = a1.x; val[3] = a1.y;
                    } else {
                        const float4 a0 = *reinterpret_cast<const float4*>(arr + r0 * XS);
                        val[0] = a0.x; val[1] = a0.y; val[2] = a0.z; val[3] = a0.w;
                    }
                };
                get(py, yt);
                if (TWO) get(py + 32 * XS, ya);
            }
            if (!TWO) {
#pragma unroll
                for (int e = 0; e < 4; ++e) ya[e] = yt[e];
            }
            float tg[4], ad[4], vsa[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                tg[e] = VT_T ? __fadd_rn(yt[e], v[e]) : yt[e];               // vs = acc + v (losses.py:54) | G
                vsa[e] = VT_A ? ((TWO || !VT_T) ? __fadd_rn(ya[e], v[e]) : tg[e]) : 0.f;
            }
            if (VT_A) {
#pragma unroll
                for (int p = 0; p < P; ++p) cr[p] = (on && u == upr - 1) ? cin[row * P + p] : 0.f;
            }
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                if (VT_A) {   // (r + gamma * vs[t+1]) - v, vs[T] = bootstrap (losses.py:55-56)
                    float vn;
                    if (e + P < 4) {
                        vn = vsa[e + P];
                    } else {
                        const float nb = __shfl_down_sync(0xffffffffu, vsa[e + P - 4], 1);
                        vn = u == upr - 1 ? cr[P == 2 ? (e & 1) : 0] : nb;
                    }
                    const float r_ = HAS_REW ? rw[e] : 0.f;
                    ad[e] = __fsub_rn(__fadd_rn(r_, __fmul_rn(k.gamma, vn)), v[e]);
                } else {
                    ad[e] = __fsub_rn(ya[e], v[e]);                          // G - v (losses.py:26,38)
                }
            }
            if (on) {
                st_stream_f4(reinterpret_cast<float4*>(gt + (int64_t)r0 * row_out), make_float4(tg[0], tg[1], tg[2], tg[3]));
                st_stream_f4(reinterpret_cast<float4*>(ga + (int64_t)r0 * row_out), make_float4(ad[0], ad[1], ad[2], ad[3]));
                if (VT_A && u == 0) {
#pragma unroll
                    for (int p = 0; p < P; ++p) cout[row * P + p] = vsa[p];
                }
            }
        };
        if (rows == TR) {
#pragma unroll
            for (int r0 = 0; r0 < TR; r0 += RPP) pass(r0);
        } else {
            for (int r0 = 0; r0 < rows; r0 += RPP) pass(r0);
        }
        __syncwarp();
        mbar_arrive(empty + slot);   // in program order after the loads of the raw chunk and of the scan output
        if (++slot == ns) {
            slot = 0;
            par ^= 1;
        }
    }
}

template <int MT, int MA, int P, bool HAS_REW, bool RHO_PP, int RTC>
__global__ void __launch_bounds__(128) k_target_rows_st(const ScanParams k, int c_alias, int ns) {
    using L = StagedLayout<MT, MA, P, HAS_REW, RHO_PP, RTC>;
    constexpr bool VT_T = L::VT_T, VT_A = L::VT_A, NEED_RHO = L::NEED_RHO, NEED_G = L::NEED_G, TWO = L::TWO;
    constexpr int TR = L::TR, WR = L::WR, RSV = L::RSV, RSR = L::RSR, XS = L::XS, UPR = L::UPR, RPP = L::RPP;
    constexpr int UPT = 4 / P;                        // timesteps per 16-byte unit of a raw row
    extern __shared__ __align__(16) float smem[];
    const bool alias = c_alias != 0;
    constexpr int OFF_X = L::OFF_X, OFF_Y = L::OFF_Y, STAGE = L::STAGE;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem);
    uint64_t* ready = full + WS_MAX_STAGES;
    uint64_t* done = ready + WS_MAX_STAGES;
    uint64_t* empty = done + WS_MAX_STAGES;
    float* ring = smem + ST_BAR_FLOATS;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int b0 = blockIdx.x * TR;
    const int rows = min(TR, k.B - b0);
    const int n = (k.T + RTC - 1) / RTC;              // chunks, processed in order i = 0..n-1 <-> t0 = (n-1-i) * RTC
    if (threadIdx.x == 0) {
        for (int s = 0; s < ns; ++s) {
            mbar_init(full + s, 32);
            mbar_init(ready + s, 32);
            mbar_init(done + s, 32);
            mbar_init(empty + s, 32);
        }
    }
    __syncthreads();

    if (warp == 0) {
        // ---------------- loader ----------------
        LaneIO<P, TR, RTC> io_v, io_r;
        LaneIO<WR, TR, RTC> io_rho, io_c;
        io_v.init(k.v.sb, b0, lane);
        if (HAS_REW) io_r.init(k.rew.sb, b0, lane);
        if (NEED_RHO) {
            io_rho.init(k.rho.sb, b0, lane);
            io_c.init(k.c.sb, b0, lane);
        }
        int slot = 0;
        uint32_t par = 1;   // a fresh mbarrier passes a wait on the phase "before the first": all slots start empty
        for (int i = 0; i < n; ++i) {
            float* st = ring + slot * STAGE;
            const int t0 = (n - 1 - i) * RTC, tl = min(RTC, k.T - t0);
            mbar_wait(empty + slot, par);
            if (tl == RTC) {
                fetch_full<P, RSV, TR, RTC>(st + L::OFF_V, k.v.p, io_v, rows, t0, lane);
                if (HAS_REW) fetch_full<P, RSV, TR, RTC>(st + L::OFF_R, k.rew.p, io_r, rows, t0, lane);
                if (NEED_RHO) {
                    fetch_full<WR, RSR, TR, RTC>(st + L::OFF_RHO, k.rho.p, io_rho, rows, t0, lane);
                    if (!alias) fetch_full<WR, RSR, TR, RTC>(st + L::OFF_RHO + TR * RSR, k.c.p, io_c, rows, t0, lane);
                }
            } else {
                fetch_tile<P, RSV>(st + L::OFF_V, k.v.p, k.v.sb, b0, rows, t0, tl, lane);
                if (HAS_REW) fetch_tile<P, RSV>(st + L::OFF_R, k.rew.p, k.rew.sb, b0, rows, t0, tl, lane);
                if (NEED_RHO) {
                    fetch_tile<WR, RSR>(st + L::OFF_RHO, k.rho.p, k.rho.sb, b0, rows, t0, tl, lane);
                    if (!alias) fetch_tile<WR, RSR>(st + L::OFF_RHO + TR * RSR, k.c.p, k.c.sb, b0, rows, t0, tl, lane);
                }
            }
            mbar_arrive_on_cp_async(full + slot);
            if (++slot == ns) {
                slot = 0;
                par ^= 1;
            }
        }
        cp_async_wait<0>();   // nothing of this thread is in flight when the CTA retires
    } else if (warp == 1) {
        staged_pre_warp<MT, MA, P, HAS_REW, RHO_PP, RTC>(k, ring, smem + ST_BAR_ONLY, alias, full, ready, ns, n,
                                                         b0, rows, lane);
    } else if (warp == 2) {
        staged_scan_warp<MT, MA, P, HAS_REW, RHO_PP, RTC>(k, ring, ready, done, ns, n, b0, rows, lane);
    } else {
        staged_post_warp<MT, MA, P, HAS_REW, RHO_PP, RTC>(k, ring, smem + ST_BAR_ONLY + 64, done, empty, ns, n,
                                                          b0, rows, lane);
    }
}
