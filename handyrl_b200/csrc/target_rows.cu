// Target recursion for the layouts compute_loss actually produces (sm_100a): shared-memory pipelines fed by cp.async.
//
// Fast paths of b200rl_target_scan / _fused / _heads (handyrl/train.py:218-258): P in {1,2}, values / rewards
// contiguous (B,T,P), rhos / cs either (B,T,1) broadcast over players (turn-based training) or (B,T,P),
// T % 4 == 0, 16-byte aligned.  Everything else goes to the generic kernel in target_scan.cu.
//
// Rows of <= 64 steps (k_target_rows, k_target_rows2): one WARP (= one CTA, so the hardware hands tiles to SMs as
// they free up) owns a tile of 32 batch rows and walks it backward in chunks of 16 timesteps.  Chunks are fetched
// with 16-byte cp.async (coalesced along the rows, no register staging) into a warp-private ring of 2-4 stages, so
// the loads of the next chunks are in flight while one is scanned.  Each lane scans ONE row — P independent
// sequences, which gives the dependent fp32 chains some ILP — reading its row with conflict-free LDS.128 (row
// stride = RTC*P+4 floats, an odd number of 16-byte units) and writing targets / advantages in place; the warp
// then streams the tile out with coalesced 128-bit stores.  No __syncthreads, no CTA-wide phases.
// k_target_rows2 runs the value head and the return head of one loss in one launch.
//
// Longer rows (k_target_rows_ws): three specialised warps per tile — loader, compute (one sequence per lane),
// storer — around an mbarrier ring; see the comment at the kernel.
//
// Arithmetic is the same Rec<> code as the generic kernel: bit-identical to torch fp32.
#include <cstdlib>
#include <type_traits>

#include "target_common.cuh"

namespace {

#ifndef B200RL_SHORT_T
#define B200RL_SHORT_T 64
#endif
#ifndef B200RL_NSTAGE
#define B200RL_NSTAGE 2
#endif
// Tile shapes (batch rows per tile x timesteps per chunk).  What HBM delivers for this access pattern is set by the
// contiguous bytes fetched per row (tools/probe_rows.cu: 128 B -> 4.9 TB/s, 256 B -> 5.8 TB/s, 1 KB -> 6.0 of a
// 6.2 TB/s streaming copy), so long rows fetch 256-512 B per row and chunk (16 rows x 32-64 steps at P = 2) and
// short rows, where a whole 32-row tile is one contiguous block anyway, use 32 x 16.

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
    asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
// wait until at most n (1..3, warp-uniform) commit groups are pending
__device__ __forceinline__ void cp_async_wait_n(int n) {
    if (n == 1) cp_async_wait<1>();
    else if (n == 2) cp_async_wait<2>();
    else cp_async_wait<3>();
}

// copy rows x (tl*W/4) 16-byte units of a contiguous (B, T, W) view into a [32][RS] tile
template <int W, int RS>
__device__ __forceinline__ void fetch_tile(float* tile, const float* __restrict__ g, int64_t sb, int b0, int rows,
                                           int t0, int tl, int lane) {
    const int upr = (tl * W) >> 2;  // 16-byte units per row
    const int total = rows * upr;
    const float* base = g + (int64_t)b0 * sb + (int64_t)t0 * W;
    for (int e = lane; e < total; e += 32) {
        const int r = e / upr, q = e - r * upr;
        cp_async16(tile + r * RS + q * 4, base + (int64_t)r * sb + q * 4);
    }
}

template <int W, int RS>
__device__ __forceinline__ void flush_tile(const float* tile, float* __restrict__ out, int64_t sb, int b0, int rows,
                                           int t0, int tl, int lane) {
    const int upr = (tl * W) >> 2;
    const int total = rows * upr;
    float* base = out + (int64_t)b0 * sb + (int64_t)t0 * W;
    for (int e = lane; e < total; e += 32) {
        const int r = e / upr, q = e - r * upr;
        const float4 v = *reinterpret_cast<const float4*>(tile + r * RS + q * 4);
        st_stream_f4(reinterpret_cast<float4*>(base + (int64_t)r * sb + q * 4), v);
    }
}

// Full-chunk I/O with per-lane addressing hoisted out of the chunk loop: a (32 rows x RTC*W floats) tile is
// U = RTC*W/4 sixteen-byte units per row, so lane l always moves unit q = l % U of rows r0 + i*RPI
// (r0 = l / U, RPI = 32 / U): global and shared addresses advance by constants, no division or 64-bit multiply.
template <int W, int TR, int RTC>
struct LaneIO {
    static constexpr int U = RTC * W / 4, RPI = 32 / U, NI = TR / RPI;
    static_assert(U <= 32 && 32 % U == 0 && TR % RPI == 0, "chunk width must divide a warp");
    int64_t off, step;  // element offset of (row b0+r0, unit q) at t = 0; RPI rows in elements
    __device__ __forceinline__ void init(int64_t sb, int b0, int lane) {
        off = (int64_t)(b0 + lane / U) * sb + (lane % U) * 4;
        step = (int64_t)RPI * sb;
    }
};

__device__ __forceinline__ void cp_async16_s(uint32_t smem_addr, const void* gmem) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_addr), "l"(gmem) : "memory");
}

// Running pointers (one 64-bit add per row) and a predicate-free path for full tiles: the copy instructions share
// the warp's issue slots with the recursion, so their address arithmetic is kept to two instructions per 16 bytes.
template <int W, int RS, int TR, int RTC>
__device__ __forceinline__ void fetch_full(float* tile, const float* __restrict__ g, const LaneIO<W, TR, RTC>& io,
                                           int rows, int t0, int lane) {
    using IO = LaneIO<W, TR, RTC>;
    constexpr int U = IO::U, RPI = IO::RPI, NI = IO::NI;
    const int r0 = lane / U;
    const float* p = g + io.off + t0 * W;
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(tile + r0 * RS + (lane % U) * 4);
    if (rows == TR) {
#pragma unroll
        for (int i = 0; i < NI; ++i) {
            cp_async16_s(s + i * RPI * RS * 4, p);
            p += io.step;
        }
    } else {
#pragma unroll
        for (int i = 0; i < NI; ++i) {
            if (r0 + i * RPI < rows) cp_async16_s(s + i * RPI * RS * 4, p);
            p += io.step;
        }
    }
}

template <int W, int RS, int TR, int RTC>
__device__ __forceinline__ void flush_full(const float* tile, float* __restrict__ out, const LaneIO<W, TR, RTC>& io,
                                           int rows, int t0, int lane) {
    using IO = LaneIO<W, TR, RTC>;
    constexpr int U = IO::U, RPI = IO::RPI, NI = IO::NI;
    const int r0 = lane / U;
    float* p = out + io.off + t0 * W;
    const float* s = tile + r0 * RS + (lane % U) * 4;
    if (rows == TR) {
#pragma unroll
        for (int i = 0; i < NI; ++i) {
            st_stream_f4(reinterpret_cast<float4*>(p), *reinterpret_cast<const float4*>(s + i * RPI * RS));
            p += io.step;
        }
    } else {
#pragma unroll
        for (int i = 0; i < NI; ++i) {
            if (r0 + i * RPI < rows)
                st_stream_f4(reinterpret_cast<float4*>(p), *reinterpret_cast<const float4*>(s + i * RPI * RS));
            p += io.step;
        }
    }
}

// Four timesteps (descending) of the P sequences of one row.  FIRST: the group contains t = T-1 at j = 3.
// JHI..JLO: the timesteps of the group to process (the row-per-lane pipeline below works in halves).
template <int MT, int MA, int P, bool HAS_REW, bool NEED_RHO, bool RHO_PP, bool FIRST, int JHI = 3, int JLO = 0>
__device__ __forceinline__ void scan_group(const float (&v)[4 * P], const float (&rw)[4 * P],
                                           const float (&rho)[4 * (RHO_PP ? P : 1)],
                                           const float (&cc)[4 * (RHO_PP ? P : 1)], Rec<MT> (&rec_t)[P],
                                           Rec<MA> (&rec_a)[P], const float (&R)[P], const ScanParams& k,
                                           float (&tg)[4 * P], float (&ad)[4 * P]) {
#pragma unroll
    for (int j = JHI; j >= JLO; --j) {
#pragma unroll
        for (int p = 0; p < P; ++p) {
            const int i = j * P + p;
            // `rewards is None` -> the reference adds the literal 0 (losses.py:23,35,44)
            const float r_ = HAS_REW ? rw[i] : 0.f;
            const float rh = NEED_RHO ? rho[RHO_PP ? i : j] : 0.f;
            const float c_ = NEED_RHO ? cc[RHO_PP ? i : j] : 0.f;
            float t1, a1, t2, a2;
            if (FIRST && j == 3) {
                rec_t[p].first(v[i], r_, rh, c_, R[p], k, t1, a1);
                if (MT != MA) {
                    rec_a[p].first(v[i], r_, rh, c_, R[p], k, t2, a2);
                    a1 = a2;
                }
            } else {
                rec_t[p].next(v[i], r_, rh, c_, k, t1, a1);
                if (MT != MA) {
                    rec_a[p].next(v[i], r_, rh, c_, k, t2, a2);
                    a1 = a2;
                }
            }
            tg[i] = t1;
            ad[i] = a1;
        }
    }
}

// One chunk of `tl` timesteps (a multiple of 4, descending) of ONE batch row held by this lane, all P players: the
// row's slivers sit in shared memory at pv (values -> advantages), pr (rewards -> targets), prho / pc.  Every
// LDS.128 / STS.128 carries 4 useful floats (2 timesteps x 2 players at P = 2), and the P independent sequences of
// the row interleave in the fp32 pipe.  Software pipeline over the 4-timestep groups with two register sets
// (ping-pong): the LDS.128 of group q-1 are issued before the dependent chain of group q, and no register is copied
// between iterations.  `first_chunk`: the chunk starts with the bootstrap step t = T-1, whose group is peeled off.
template <int MT, int MA, int P, bool HAS_REW, bool NEED_RHO, bool RHO_PP>
__device__ __forceinline__ void scan_chunk_rowlane(float* pv, float* pr, const float* prho, const float* pc,
                                                   bool c_alias, bool rew_rt, int tl, bool first_chunk,
                                                   Rec<MT> (&rec_t)[P], Rec<MA> (&rec_a)[P], const float (&R)[P],
                                                   const ScanParams& k) {
    constexpr int WR = RHO_PP ? P : 1;
    struct Grp { float4 v[P], r[P], rho[WR], c[WR]; };
    auto load_group = [&](Grp& g, int q) {
#pragma unroll
        for (int u = 0; u < P; ++u) {
            g.v[u] = *reinterpret_cast<const float4*>(pv + q * 4 * P + 4 * u);
            if (HAS_REW) {
                if (rew_rt) g.r[u] = *reinterpret_cast<const float4*>(pr + q * 4 * P + 4 * u);
                else g.r[u] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
        if (NEED_RHO) {
#pragma unroll
            for (int u = 0; u < WR; ++u) {
                g.rho[u] = *reinterpret_cast<const float4*>(prho + q * 4 * WR + 4 * u);
                if (!c_alias) g.c[u] = *reinterpret_cast<const float4*>(pc + q * 4 * WR + 4 * u);
            }
        }
    };
    auto compute_group = [&](const Grp& g, int q, auto first_tag) {
        constexpr bool FIRST = decltype(first_tag)::value;
        float v[4 * P], rw[4 * P], rho[4 * WR], cc[4 * WR];
#pragma unroll
        for (int u = 0; u < P; ++u) {
            v[4 * u] = g.v[u].x; v[4 * u + 1] = g.v[u].y; v[4 * u + 2] = g.v[u].z; v[4 * u + 3] = g.v[u].w;
            if (HAS_REW) {
                rw[4 * u] = g.r[u].x; rw[4 * u + 1] = g.r[u].y; rw[4 * u + 2] = g.r[u].z; rw[4 * u + 3] = g.r[u].w;
            }
        }
        if (NEED_RHO) {
#pragma unroll
            for (int u = 0; u < WR; ++u) {
                rho[4 * u] = g.rho[u].x; rho[4 * u + 1] = g.rho[u].y; rho[4 * u + 2] = g.rho[u].z; rho[4 * u + 3] = g.rho[u].w;
                const float4 c4 = c_alias ? g.rho[u] : g.c[u];
                cc[4 * u] = c4.x; cc[4 * u + 1] = c4.y; cc[4 * u + 2] = c4.z; cc[4 * u + 3] = c4.w;
            }
        }
        float tg[4 * P], ad[4 * P];
        scan_group<MT, MA, P, HAS_REW, NEED_RHO, RHO_PP, FIRST>(v, rw, rho, cc, rec_t, rec_a, R, k, tg, ad);
#pragma unroll
        for (int u = 0; u < P; ++u) {
            *reinterpret_cast<float4*>(pv + q * 4 * P + 4 * u) =
                make_float4(ad[4 * u], ad[4 * u + 1], ad[4 * u + 2], ad[4 * u + 3]);
            *reinterpret_cast<float4*>(pr + q * 4 * P + 4 * u) =
                make_float4(tg[4 * u], tg[4 * u + 1], tg[4 * u + 2], tg[4 * u + 3]);
        }
    };
    Grp ga, gb;
    int q = (tl >> 2) - 1;
    load_group(ga, q);
    if (first_chunk) {   // warp-uniform
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

// Driver of the pinned-order scans below.  Q >= 1 groups of four timesteps, processed from group Q-1 down to 0 with two
// input (ia, ib) and two result (ra, rb) register sets in ping-pong:
//     body(cur, res, nxt, prev, q, first, store_prev, load_next) =
//         steps 3,2 of group q | store(prev, q+1) | load(nxt, q-1) | steps 1,0 of group q
// The loop is unrolled by two so every set is addressed statically; first / store_prev / load_next are tag types.
template <class In, class Out, class Load, class Store, class Body>
__device__ __forceinline__ void pinned_order_groups(int Q, bool first_chunk, In& ia, In& ib, Out& ra, Out& rb, Load load,
                                                    Store store, Body body) {
    using Y = std::true_type;
    using N = std::false_type;
    load(ia, Q - 1);
    if (Q == 1) {
        if (first_chunk) body(ia, ra, ib, rb, 0, Y{}, N{}, N{});
        else body(ia, ra, ib, rb, 0, N{}, N{}, N{});
        store(ra, 0);
        return;
    }
    if (first_chunk) body(ia, ra, ib, rb, Q - 1, Y{}, N{}, Y{});
    else body(ia, ra, ib, rb, Q - 1, N{}, N{}, Y{});
    int q = Q - 2;
    for (; q >= 2; q -= 2) {
        body(ib, rb, ia, ra, q, N{}, Y{}, Y{});
        body(ia, ra, ib, rb, q - 1, N{}, Y{}, Y{});
    }
    if (q == 1) {
        body(ib, rb, ia, ra, 1, N{}, Y{}, Y{});
        body(ia, ra, ib, rb, 0, N{}, Y{}, N{});
        store(ra, 0);
    } else {
        body(ib, rb, ia, ra, 0, N{}, Y{}, N{});
        store(rb, 0);
    }
}

__device__ __forceinline__ float lds_volatile(const float* p) {
    float r;
    asm volatile("ld.volatile.shared.f32 %0, [%1];" : "=f"(r) : "r"((uint32_t)__cvta_generic_to_shared(p)));
    return r;
}
__device__ __forceinline__ void sts_volatile(float* p, float v) {
    asm volatile("st.volatile.shared.f32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(p)), "f"(v));
}
__device__ __forceinline__ void lds4_volatile(const float* p, float* d) {
    asm volatile("ld.volatile.shared.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(d[0]), "=f"(d[1]), "=f"(d[2]), "=f"(d[3])
                 : "r"((uint32_t)__cvta_generic_to_shared(p)));
}
__device__ __forceinline__ void sts4_volatile(float* p, const float* d) {
    asm volatile("st.volatile.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"((uint32_t)__cvta_generic_to_shared(p)), "f"(d[0]),
                 "f"(d[1]), "f"(d[2]), "f"(d[3]));
}

// One chunk of `tl` timesteps (a multiple of 4) of ONE (row, player) sequence held by this lane, in pinned order (see
// scan_chunk_rowlane_v2 below for why).  pv / pr point at this sequence's first element in the row-major [t][p] slivers
// (stride P floats per timestep); prho / pc at its rho / c (stride WR).
template <int MT, int MA, int P, bool HAS_REW, bool NEED_RHO, bool RHO_PP>
__device__ __forceinline__ void scan_chunk_seqlane_v2(float* pv, float* pr, const float* prho, const float* pc,
                                                      bool c_alias, int tl, bool first_chunk, Rec<MT>& rec_t,
                                                      Rec<MA>& rec_a, float R, const ScanParams& k) {
    constexpr int WR = RHO_PP ? P : 1;
    struct In { float v[4], r[4], rho[4], c[4]; };
    struct Out { float t[4], a[4]; };
    auto load = [&](In& g, int q) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            g.v[j] = lds_volatile(pv + (q * 4 + j) * P);
            if (HAS_REW) g.r[j] = lds_volatile(pr + (q * 4 + j) * P);
        }
        if (NEED_RHO) {
            if (WR == 1) {   // rho / c broadcast over players: 4 steps are one 16-byte unit
                lds4_volatile(prho + q * 4, g.rho);
                if (!c_alias) lds4_volatile(pc + q * 4, g.c);
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    g.rho[j] = lds_volatile(prho + (q * 4 + j) * WR);
                    if (!c_alias) g.c[j] = lds_volatile(pc + (q * 4 + j) * WR);
                }
            }
        }
    };
    auto store = [&](const Out& o, int q) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            sts_volatile(pv + (q * 4 + j) * P, o.a[j]);
            sts_volatile(pr + (q * 4 + j) * P, o.t[j]);
        }
    };
    auto steps = [&](const In& g, Out& o, auto first_tag, auto hi_tag) {
        constexpr bool FIRST = decltype(first_tag)::value;
        constexpr int JHI = decltype(hi_tag)::value ? 3 : 1;
#pragma unroll
        for (int j = JHI; j >= JHI - 1; --j) {
            const float r_ = HAS_REW ? g.r[j] : 0.f;
            const float rh = NEED_RHO ? g.rho[j] : 0.f;
            const float c_ = NEED_RHO ? (c_alias ? g.rho[j] : g.c[j]) : 0.f;
            float t1, a1, t2, a2;
            if (FIRST && j == 3) {
                rec_t.first(g.v[j], r_, rh, c_, R, k, t1, a1);
                if (MT != MA) {
                    rec_a.first(g.v[j], r_, rh, c_, R, k, t2, a2);
                    a1 = a2;
                }
            } else {
                rec_t.next(g.v[j], r_, rh, c_, k, t1, a1);
                if (MT != MA) {
                    rec_a.next(g.v[j], r_, rh, c_, k, t2, a2);
                    a1 = a2;
                }
            }
            o.t[j] = t1;
            o.a[j] = a1;
        }
    };
    auto body = [&](In& cur, Out& res, In& nxt, const Out& prev, int q, auto first_tag, auto store_prev, auto load_next) {
        steps(cur, res, first_tag, std::true_type{});
        if (decltype(store_prev)::value) store(prev, q + 1);
        if (decltype(load_next)::value) load(nxt, q - 1);
        steps(cur, res, std::false_type{}, std::false_type{});
    };
    In ia, ib;
    Out ra, rb;
    pinned_order_groups(tl >> 2, first_chunk, ia, ib, ra, rb, load, store, body);
}

#ifndef B200RL_ROWLANE_V2
#define B200RL_ROWLANE_V2 1
#endif
// scan_chunk_rowlane for the warp-specialised kernel at P = 2, with the order of the shared-memory instructions pinned
// (volatile accesses keep their program order through ptxas).  Why (profiles/r2_target_ws_rowlane.md, section 5): in the
// version above ptxas issues the LDS.128 batch of the next group next to the first uses of the current one and gives both
// the same scoreboard, so a consumer of completed loads also waits ~29 cycles for the fresh ones; and it overwrites the
// source registers of an STS.128 a few instructions after the store, which then waits for the store to read them.  Here a
// group of four timesteps is computed in two halves and the memory instructions sit between them:
//     steps 3,2 of group q | STS results of group q+1 | LDS inputs of group q-1 | steps 1,0 of group q
// with two input and two result register sets, so loads are consumed half a group after they were issued (and only one
// batch is ever outstanding) and result registers are rewritten half a group after their store.
template <int MT, int MA, bool HAS_REW, bool NEED_RHO, bool RHO_PP>
__device__ __forceinline__ void scan_chunk_rowlane_v2(float* pv, float* pr, const float* prho, const float* pc,
                                                      bool c_alias, int tl, bool first_chunk, Rec<MT> (&rec_t)[2],
                                                      Rec<MA> (&rec_a)[2], const float (&R)[2], const ScanParams& k) {
    constexpr int P = 2;
    constexpr int WR = RHO_PP ? P : 1;
    struct In { float v[4 * P], r[4 * P], rho[4 * WR], c[4 * WR]; };
    struct Out { float t[4 * P], a[4 * P]; };
    auto lds4 = [](const float* p, float* d) { lds4_volatile(p, d); };
    auto sts4 = [](float* p, const float* d) { sts4_volatile(p, d); };
    auto load = [&](In& g, int q) {
#pragma unroll
        for (int u = 0; u < P; ++u) {
            lds4(pv + q * 4 * P + 4 * u, g.v + 4 * u);
            if (HAS_REW) lds4(pr + q * 4 * P + 4 * u, g.r + 4 * u);
        }
        if (NEED_RHO) {
#pragma unroll
            for (int u = 0; u < WR; ++u) {
                lds4(prho + q * 4 * WR + 4 * u, g.rho + 4 * u);
                if (!c_alias) lds4(pc + q * 4 * WR + 4 * u, g.c + 4 * u);
            }
        }
    };
    auto store = [&](const Out& o, int q) {
#pragma unroll
        for (int u = 0; u < P; ++u) {
            sts4(pv + q * 4 * P + 4 * u, o.a + 4 * u);
            sts4(pr + q * 4 * P + 4 * u, o.t + 4 * u);
        }
    };
    // one group: cur -> res; `prev` = results of group q + 1 still in registers, `nxt` = inputs of group q - 1
    auto body = [&](In& cur, Out& res, In& nxt, const Out& prev, int q, auto first_tag, auto store_prev, auto load_next) {
        constexpr bool FIRST = decltype(first_tag)::value;
        if (NEED_RHO && c_alias) {
#pragma unroll
            for (int i = 0; i < 4 * WR; ++i) cur.c[i] = cur.rho[i];
        }
        scan_group<MT, MA, P, HAS_REW, NEED_RHO, RHO_PP, FIRST, 3, 2>(cur.v, cur.r, cur.rho, cur.c, rec_t, rec_a, R, k,
                                                                        res.t, res.a);
        if (decltype(store_prev)::value) store(prev, q + 1);
        if (decltype(load_next)::value) load(nxt, q - 1);
        scan_group<MT, MA, P, HAS_REW, NEED_RHO, RHO_PP, false, 1, 0>(cur.v, cur.r, cur.rho, cur.c, rec_t, rec_a, R, k,
                                                                        res.t, res.a);
    };
    In ia, ib;
    Out ra, rb;
    if (!HAS_REW) {
#pragma unroll
        for (int i = 0; i < 4 * P; ++i) ia.r[i] = ib.r[i] = 0.f;
    }
    pinned_order_groups(tl >> 2, first_chunk, ia, ib, ra, rb, load, store, body);
}

// ns = ring depth (2..4 stages): 2 when the grid fills the SMs (more resident warps hide the latency better than
// deeper rings), up to 4 when there are few tiles and each warp has to cover the DRAM latency on its own.
// DUAL: the tile belongs to one of two heads launched together (k_target_rows2); whether that head has rewards
// is then a run-time property of its ScanParams (a head without rewards scans with r = 0, like HAS_REW = false).
template <int MT, int MA, int P, bool HAS_REW, bool NEED_RHO, bool RHO_PP, int TR, int RTC, bool DUAL>
__device__ __forceinline__ void target_rows_tiles(const ScanParams& k, int tile0, int tile_step, int ntiles, int c_alias,
                                                  int ns) {
    extern __shared__ __align__(16) float smem[];
    const bool rew_rt = !DUAL || k.rew.p != nullptr;
    constexpr int RSV = RTC * P + 4;                 // row stride of the value / reward tiles (floats)
    constexpr int WR = RHO_PP ? P : 1;               // floats per timestep of rho / c
    constexpr int RSR = RTC * WR + 4;
    // c_alias: cs is the same tensor as rhos (the reference clips both at 1.0, train.py:222,230-231): one tile
    const int STAGE = TR * (2 * RSV + (NEED_RHO ? (c_alias ? RSR : 2 * RSR) : 0));
    const int lane = threadIdx.x;
    const int64_t row_out = (int64_t)k.T * P;

    for (int tile = tile0; tile < ntiles; tile += tile_step) {
        const int b0 = tile * TR;
        const int rows = min(TR, k.B - b0);
        const bool active = lane < rows;
        float R[P];
#pragma unroll
        for (int p = 0; p < P; ++p)
            R[p] = active ? k.boot[(int64_t)(b0 + lane) * k.boot_sb + (int64_t)p * k.boot_sp] : 0.f;
        Rec<MT> rec_t[P];
        Rec<MA> rec_a[P];

        const int nchunks = (k.T + RTC - 1) / RTC;
        LaneIO<P, TR, RTC> io_v, io_r, io_o;
        LaneIO<WR, TR, RTC> io_rho, io_c;
        io_v.init(k.v.sb, b0, lane);
        io_o.init(row_out, b0, lane);
        if (HAS_REW && rew_rt) io_r.init(k.rew.sb, b0, lane);
        if (NEED_RHO) {
            io_rho.init(k.rho.sb, b0, lane);
            io_c.init(k.c.sb, b0, lane);
        }
        auto fetch = [&](int ch) {
            if (ch >= 0) {
                float* st = smem + ((nchunks - 1 - ch) % ns) * STAGE;
                const int t0 = ch * RTC, tl = min(RTC, k.T - t0);
                if (tl == RTC) {
                    fetch_full<P, RSV, TR, RTC>(st, k.v.p, io_v, rows, t0, lane);
                    if (HAS_REW && rew_rt) fetch_full<P, RSV, TR, RTC>(st + TR * RSV, k.rew.p, io_r, rows, t0, lane);
                    if (NEED_RHO) {
                        fetch_full<WR, RSR, TR, RTC>(st + 2 * TR * RSV, k.rho.p, io_rho, rows, t0, lane);
                        if (!c_alias) fetch_full<WR, RSR, TR, RTC>(st + 2 * TR * RSV + TR * RSR, k.c.p, io_c, rows, t0, lane);
                    }
                } else {
                    fetch_tile<P, RSV>(st, k.v.p, k.v.sb, b0, rows, t0, tl, lane);
                    if (HAS_REW && rew_rt) fetch_tile<P, RSV>(st + TR * RSV, k.rew.p, k.rew.sb, b0, rows, t0, tl, lane);
                    if (NEED_RHO) {
                        fetch_tile<WR, RSR>(st + 2 * TR * RSV, k.rho.p, k.rho.sb, b0, rows, t0, tl, lane);
                        if (!c_alias) fetch_tile<WR, RSR>(st + 2 * TR * RSV + TR * RSR, k.c.p, k.c.sb, b0, rows, t0, tl, lane);
                    }
                }
            }
            cp_async_commit();  // one group per slot, possibly empty, keeps wait_group counting uniform
        };
        for (int j = 0; j < ns - 1; ++j) fetch(nchunks - 1 - j);

        for (int ch = nchunks - 1; ch >= 0; --ch) {
            fetch(ch - (ns - 1));
            cp_async_wait_n(ns - 1);
            __syncwarp();
            float* st = smem + ((nchunks - 1 - ch) % ns) * STAGE;
            const int t0 = ch * RTC, tl = min(RTC, k.T - t0);
            if (active) {
                float* pv = st + lane * RSV;
                float* pr = st + TR * RSV + lane * RSV;
                const float* prho = st + 2 * TR * RSV + lane * RSR;
                const float* pc = c_alias ? prho : prho + TR * RSR;
                scan_chunk_rowlane<MT, MA, P, HAS_REW, NEED_RHO, RHO_PP>(pv, pr, prho, pc, c_alias != 0, rew_rt, tl,
                                                                         ch == nchunks - 1, rec_t, rec_a, R, k);
            }
            __syncwarp();
            if (tl == RTC) {
                flush_full<P, RSV, TR, RTC>(st + TR * RSV, k.out_t, io_o, rows, t0, lane);
                flush_full<P, RSV, TR, RTC>(st, k.out_a, io_o, rows, t0, lane);
            } else {
                flush_tile<P, RSV>(st + TR * RSV, k.out_t, row_out, b0, rows, t0, tl, lane);
                flush_tile<P, RSV>(st, k.out_a, row_out, b0, rows, t0, tl, lane);
            }
            __syncwarp();
        }
    }
}

template <int MT, int MA, int P, bool HAS_REW, bool NEED_RHO, bool RHO_PP, int TR, int RTC>
__global__ void __launch_bounds__(32) k_target_rows(const ScanParams k, int ntiles, int c_alias, int ns) {
    target_rows_tiles<MT, MA, P, HAS_REW, NEED_RHO, RHO_PP, TR, RTC, false>(k, blockIdx.x, gridDim.x, ntiles, c_alias, ns);
}

// Two heads of one loss (value head: outcome bootstrap, no rewards, gamma 1; return head: rewards, gamma) in ONE
// launch: CTA 2j is tile j of head 0, CTA 2j+1 tile j of head 1.  At B = 65,536, T = 32 each head is a 13-16 us
// kernel; two launches cost their ramp-up and drain twice (bench.py: 0.042 ms for 0.029 ms of kernel time).
struct ScanParams2 {
    ScanParams h[2];
};
template <int MT, int MA, int P, bool NEED_RHO, bool RHO_PP, int TR, int RTC>
__global__ void __launch_bounds__(32) k_target_rows2(const ScanParams2 kk, int ntiles, int c_alias, int ns) {
    target_rows_tiles<MT, MA, P, true, NEED_RHO, RHO_PP, TR, RTC, true>(kk.h[blockIdx.x & 1], blockIdx.x >> 1, ntiles,
                                                                        ntiles, c_alias, ns);
}

// ---------------------------------------------------------------------------------------------------------------
// Warp-specialised variant for FEW tiles (small B, long rows): the recursion is then bounded by the sequential
// depth of one warp, not by memory (profiles/r1_target_rows_smallB.md: a lone warp issues the ~30 instructions of
// a V-Trace timestep for two players at 0.3 IPC = 48 ns per timestep).  Here a CTA is three warps:
//   warp 0 (loader)   cp.async (16-byte, L2-only) of the next chunks into free ring slots; it never waits for its
//                     own copies - cp.async.mbarrier.arrive.noinc makes each lane's copies arrive on the slot's
//                     `full` mbarrier when they land;
//   warp 1 (compute)  lane = (row, player): ONE sequence per lane, so a timestep is half the instructions, and
//                     no copy instruction sits between two steps of the dependent fp32 chain;
//   warp 2 (storer)   streams finished slots out with 128-bit stores and hands them back (`empty`).
// Three mbarriers per slot (full: bytes landed, done: results written, empty: slot read out).  Two earlier
// versions are on record in profiles/: one I/O warp doing both directions (460 instructions per chunk, the
// bottleneck) and per-row TMA bulk copies (cp.async.bulk, 256 B each: the copy engine takes ~30 cycles per
// operation and SM, which caps such small copies at ~2 TB/s for the chip).
// Same tiles, same Rec<> arithmetic, same bits as k_target_rows.
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* b) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
// arrive (without raising the pending count: it is part of the barrier's init count) once all cp.async issued so
// far by this thread have landed
__device__ __forceinline__ void mbar_arrive_on_cp_async(uint64_t* b) {
    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(b)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
    uint32_t ok;
    do {
        asm volatile(
            "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
            : "=r"(ok)
            : "r"(smem_u32(b)), "r"(parity)
            : "memory");
    } while (!ok);
}

constexpr int WS_MAX_STAGES = 8;
constexpr size_t WS_SMEM_CAP = 227 * 1024;             // opt-in dynamic shared memory per CTA on sm_100
constexpr int WS_BAR_FLOATS = 3 * WS_MAX_STAGES * 2;   // 3 x 8 mbarriers of 8 bytes, in floats

// ROWLANE: the compute warp's lane owns a whole ROW (all P players) instead of one (row, player) sequence, tiles are
// 32 rows.  Why (profiles/r2_target_ws_rowlane.md): with one sequence per lane the compute warp, which is the critical
// path at small B, spends 44 cycles per timestep on 12.5 instructions - four of them scalar LDS / STS that cost 3-4
// issue cycles each, the rest stalled on the 4-deep dependent fp32 chain.  Row per lane, every LDS.128 / STS.128
// carries 2 timesteps x 2 players (4x fewer shared-memory instructions per sequence) and the second player's chain
// fills the issue slots the first one leaves empty.
template <int MT, int MA, int P, bool HAS_REW, bool NEED_RHO, bool RHO_PP, int RTC, bool ROWLANE>
__global__ void __launch_bounds__(96) k_target_rows_ws(const ScanParams k, int c_alias, int ns) {
    extern __shared__ __align__(16) float smem[];
    constexpr int TR = ROWLANE ? 32 : 32 / P;        // rows per tile: one lane per row / per (row, player)
    constexpr int RSV = RTC * P + 4;
    constexpr int WR = RHO_PP ? P : 1;
    constexpr int RSR = RTC * WR + 4;
    const int STAGE = TR * (2 * RSV + (NEED_RHO ? (c_alias ? RSR : 2 * RSR) : 0));
    uint64_t* full = reinterpret_cast<uint64_t*>(smem);
    uint64_t* done = full + WS_MAX_STAGES;
    uint64_t* empty = done + WS_MAX_STAGES;
    float* ring = smem + WS_BAR_FLOATS;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int b0 = blockIdx.x * TR;
    const int rows = min(TR, k.B - b0);
    const int n = (k.T + RTC - 1) / RTC;             // chunks, processed in order i = 0..n-1 <-> ch = n-1-i
    const int64_t row_out = (int64_t)k.T * P;

    if (threadIdx.x == 0) {
        for (int s = 0; s < ns; ++s) {
            mbar_init(full + s, 32);
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
                fetch_full<P, RSV, TR, RTC>(st, k.v.p, io_v, rows, t0, lane);
                if (HAS_REW) fetch_full<P, RSV, TR, RTC>(st + TR * RSV, k.rew.p, io_r, rows, t0, lane);
                if (NEED_RHO) {
                    fetch_full<WR, RSR, TR, RTC>(st + 2 * TR * RSV, k.rho.p, io_rho, rows, t0, lane);
                    if (!c_alias) fetch_full<WR, RSR, TR, RTC>(st + 2 * TR * RSV + TR * RSR, k.c.p, io_c, rows, t0, lane);
                }
            } else {
                fetch_tile<P, RSV>(st, k.v.p, k.v.sb, b0, rows, t0, tl, lane);
                if (HAS_REW) fetch_tile<P, RSV>(st + TR * RSV, k.rew.p, k.rew.sb, b0, rows, t0, tl, lane);
                if (NEED_RHO) {
                    fetch_tile<WR, RSR>(st + 2 * TR * RSV, k.rho.p, k.rho.sb, b0, rows, t0, tl, lane);
                    if (!c_alias) fetch_tile<WR, RSR>(st + 2 * TR * RSV + TR * RSR, k.c.p, k.c.sb, b0, rows, t0, tl, lane);
                }
            }
            mbar_arrive_on_cp_async(full + slot);
            if (++slot == ns) {
                slot = 0;
                par ^= 1;
            }
        }
        cp_async_wait<0>();   // nothing of this thread is in flight when the CTA retires
    } else if (warp == 2) {
        // ---------------- storer ----------------
        LaneIO<P, TR, RTC> io_o;
        io_o.init(row_out, b0, lane);
        int slot = 0;
        uint32_t par = 0;
        for (int i = 0; i < n; ++i) {
            const float* st = ring + slot * STAGE;
            const int t0 = (n - 1 - i) * RTC, tl = min(RTC, k.T - t0);
            mbar_wait(done + slot, par);
            if (tl == RTC) {
                flush_full<P, RSV, TR, RTC>(st + TR * RSV, k.out_t, io_o, rows, t0, lane);
                flush_full<P, RSV, TR, RTC>(st, k.out_a, io_o, rows, t0, lane);
            } else {
                flush_tile<P, RSV>(st + TR * RSV, k.out_t, row_out, b0, rows, t0, tl, lane);
                flush_tile<P, RSV>(st, k.out_a, row_out, b0, rows, t0, tl, lane);
            }
            mbar_arrive(empty + slot);   // in program order after the stores, which consumed this lane's LDS data
            if (++slot == ns) {
                slot = 0;
                par ^= 1;
            }
        }
    } else if (ROWLANE) {
        // ---------------- compute warp: lane = row, P sequences per lane ----------------
        const bool active = lane < rows;
        float R[P];
        Rec<MT> rec_t[P];
        Rec<MA> rec_a[P];
#pragma unroll
        for (int p = 0; p < P; ++p)
            R[p] = active ? k.boot[(int64_t)(b0 + lane) * k.boot_sb + (int64_t)p * k.boot_sp] : 0.f;
        int slot = 0;
        uint32_t par = 0;
        for (int i = 0; i < n; ++i) {
            float* st = ring + slot * STAGE;
            const int tl = min(RTC, k.T - (n - 1 - i) * RTC);
            mbar_wait(full + slot, par);
            if (active) {
                float* pv = st + lane * RSV;
                float* pr = st + TR * RSV + lane * RSV;
                const float* prho = st + 2 * TR * RSV + lane * RSR;
                // pinned-order version where it measured faster (1024 x 4,096, same box): with V-Trace in either role
                // (0.63 -> 0.66 value head, 0.67 -> 0.75 return head, fused pair 0.53 -> 0.58 / 0.64 -> 0.68) and TD with
                // rewards (0.81 -> 0.89); TD / UPGO without rewards and UPGO with rewards keep the plain version
                // (0.76 / 0.69 / 0.72 against 0.73 / 0.63 / 0.70)
                constexpr bool V2 = B200RL_ROWLANE_V2 != 0 && P == 2 &&
                                    (NEED_RHO || (HAS_REW && MT == B200RL_TARGET_TD && MA == B200RL_TARGET_TD));
                if constexpr (V2)
                    scan_chunk_rowlane_v2<MT, MA, HAS_REW, NEED_RHO, RHO_PP>(pv, pr, prho, prho + TR * RSR, c_alias != 0,
                                                                             tl, i == 0, rec_t, rec_a, R, k);
                else
                    scan_chunk_rowlane<MT, MA, P, HAS_REW, NEED_RHO, RHO_PP>(pv, pr, prho, prho + TR * RSR, c_alias != 0, true,
                                                                             tl, i == 0, rec_t, rec_a, R, k);
            }
            mbar_arrive(done + slot);
            if (++slot == ns) {
                slot = 0;
                par ^= 1;
            }
        }
    } else {
        // ---------------- compute warp: lane = (row, player) ----------------
        const int r = lane / P, p = lane % P;
        const bool active = r < rows;
        const float R = active ? k.boot[(int64_t)(b0 + r) * k.boot_sb + (int64_t)p * k.boot_sp] : 0.f;
        Rec<MT> rec_t;
        Rec<MA> rec_a;
        int slot = 0;
        uint32_t par = 0;
        for (int i = 0; i < n; ++i) {
            const int ch = n - 1 - i;
            float* st = ring + slot * STAGE;
            const int tl = min(RTC, k.T - ch * RTC);
            mbar_wait(full + slot, par);
            if (active) {
                float* pv = st + r * RSV + p;
                float* pr = st + TR * RSV + r * RSV + p;
                const float* prho = st + 2 * TR * RSV + r * RSR + (RHO_PP ? p : 0);
                const float* pc = prho + TR * RSR;
                scan_chunk_seqlane_v2<MT, MA, P, HAS_REW, NEED_RHO, RHO_PP>(pv, pr, prho, pc, c_alias != 0, tl, i == 0,
                                                                            rec_t, rec_a, R, k);
            }
            mbar_arrive(done + slot);
            if (++slot == ns) {
                slot = 0;
                par ^= 1;
            }
        }
    }
}

}  // namespace

namespace {

// One-time setup per DEVICE and kernel instantiation (the opt-in shared-memory attribute is per device, and a process
// may drive more than one); returns the SM count of the current device.  Races are benign: the setup is idempotent.
struct DevOnce {
    unsigned long long done = 0;
    int sms[64] = {};
};
template <class F>
int dev_setup(DevOnce& o, F&& set_attr) {
    int dev = 0;
    cudaGetDevice(&dev);
    dev &= 63;
    if (!((o.done >> dev) & 1ull)) {
        cudaDeviceGetAttribute(&o.sms[dev], cudaDevAttrMultiProcessorCount, dev);
        set_attr();
        o.done |= 1ull << dev;
    }
    return o.sms[dev];
}

constexpr int NSTAGE = B200RL_NSTAGE;  // ring depth when the grid fills the machine

template <int MT, int MA, int P, bool HAS_REW, bool NEED_RHO, bool RHO_PP, int TR, int RTC>
int launch_shape(const ScanParams& k, cudaStream_t st) {
    constexpr int RSV = RTC * P + 4;
    constexpr int RSR = RTC * (RHO_PP ? P : 1) + 4;
    const bool alias = NEED_RHO && k.c.p == k.rho.p && k.c.sb == k.rho.sb && k.c.st == k.rho.st && k.c.sp == k.rho.sp;
    const int stage = TR * (2 * RSV + (NEED_RHO ? (alias ? RSR : 2 * RSR) : 0));
    auto kern = k_target_rows<MT, MA, P, HAS_REW, NEED_RHO, RHO_PP, TR, RTC>;
    static int per_sm_tab[2][3] = {};  // resident warps per SM of this instantiation by [c aliases rho][ns - 2]
    static DevOnce once;
    const int sms = dev_setup(once, [&] {
        constexpr size_t SMEM_MAX = (size_t)4 * TR * (2 * RSV + 2 * RSR) * sizeof(float);
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_MAX);
    });
    const int ntiles = (k.B + TR - 1) / TR;
    // deepen the ring while all tiles still fit on the SMs at once (small B: latency-bound per warp)
    const int want = (ntiles + sms - 1) / sms;
    int ns = NSTAGE;
    while (ns < 4 && (size_t)want * (ns + 1) * stage * sizeof(float) <= (size_t)200 * 1024) ++ns;
    const size_t SMEM = (size_t)ns * stage * sizeof(float);
    int& per_sm = per_sm_tab[alias ? 1 : 0][ns - 2];
    if (per_sm == 0) {
        int n = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kern, 32, SMEM) != cudaSuccess || n < 1) n = 1;
        per_sm = n;
    }
    // One CTA (= one warp) per tile, dispatched by the hardware as slots free up: SMs differ by up to 2x in the
    // HBM bandwidth they can draw (tools/probe_write4.cu), so a static tile stride runs at the pace of the slowest
    // SM (6.3 TB/s for a pure write stream) while dynamic dispatch reaches 7.3.  B200RL_TARGET_PERSISTENT=1
    // restores the persistent grid-stride launch for A/B measurements.
    const bool persistent = b200rl_tuning(B200RL_TUNE_TARGET_PERSISTENT, nullptr);
    const int grid = persistent ? min(ntiles, sms * per_sm) : ntiles;
    kern<<<grid, 32, SMEM, st>>>(k, ntiles, alias ? 1 : 0, ns);
    return b200rl_check_launch("k_target_rows");
}

// Two heads in one launch (short rows only: for longer rows one head already fills the machine for > 70 us).
template <int MT, int MA, int P, bool NEED_RHO, bool RHO_PP>
int launch_dual(const ScanParams& k0, const ScanParams& k1, cudaStream_t st) {
    constexpr int TR = 32, RTC = 16;
    constexpr int RSV = RTC * P + 4;
    constexpr int RSR = RTC * (RHO_PP ? P : 1) + 4;
    auto is_alias = [](const ScanParams& k) {
        return NEED_RHO && k.c.p == k.rho.p && k.c.sb == k.rho.sb && k.c.st == k.rho.st && k.c.sp == k.rho.sp;
    };
    const bool alias = is_alias(k0);
    if (alias != is_alias(k1)) return -1;
    const int stage = TR * (2 * RSV + (NEED_RHO ? (alias ? RSR : 2 * RSR) : 0));
    auto kern = k_target_rows2<MT, MA, P, NEED_RHO, RHO_PP, TR, RTC>;
    static DevOnce once;
    const int sms = dev_setup(once, [&] {
        constexpr size_t SMEM_MAX = (size_t)4 * TR * (2 * RSV + 2 * RSR) * sizeof(float);
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_MAX);
    });
    const int ntiles = (k0.B + TR - 1) / TR;
    const int want = (2 * ntiles + sms - 1) / sms;
    int ns = NSTAGE;
    while (ns < 4 && (size_t)want * (ns + 1) * stage * sizeof(float) <= (size_t)200 * 1024) ++ns;
    ScanParams2 kk;
    kk.h[0] = k0;
    kk.h[1] = k1;
    kern<<<2 * ntiles, 32, (size_t)ns * stage * sizeof(float), st>>>(kk, ntiles, alias ? 1 : 0, ns);
    return b200rl_check_launch("k_target_rows2");
}

// Warp-specialised launch (k_target_rows_ws): one CTA of three warps per tile of 32/P rows, ring as deep as the
// shared memory allows for the number of CTAs an SM has to hold.
template <int MT, int MA, int P, bool HAS_REW, bool NEED_RHO, bool RHO_PP, int RTC, bool ROWLANE>
int launch_ws(const ScanParams& k, cudaStream_t st) {
    constexpr int TR = ROWLANE ? 32 : 32 / P;
    constexpr int RSV = RTC * P + 4;
    constexpr int RSR = RTC * (RHO_PP ? P : 1) + 4;
    const bool alias = NEED_RHO && k.c.p == k.rho.p && k.c.sb == k.rho.sb && k.c.st == k.rho.st && k.c.sp == k.rho.sp;
    const size_t stage = sizeof(float) * TR * (2 * RSV + (NEED_RHO ? (alias ? RSR : 2 * RSR) : 0));
    auto kern = k_target_rows_ws<MT, MA, P, HAS_REW, NEED_RHO, RHO_PP, RTC, ROWLANE>;
    constexpr size_t BAR = WS_BAR_FLOATS * sizeof(float);
    static DevOnce once;
    const int sms = dev_setup(once, [&] {
        constexpr size_t SMEM_ALL = BAR + (size_t)WS_MAX_STAGES * sizeof(float) * TR * (2 * RSV + 2 * RSR);
        constexpr size_t SMEM_MAX = SMEM_ALL < WS_SMEM_CAP ? SMEM_ALL : WS_SMEM_CAP;
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_MAX);
    });
    const int ntiles = (k.B + TR - 1) / TR;
    const int want = (ntiles + sms - 1) / sms;   // CTAs an SM holds when the whole grid is resident
    long long ns_knob = 0;
    const int ns_env = b200rl_tuning(B200RL_TUNE_TARGET_WS_STAGES, &ns_knob) ? (int)ns_knob : 0;
    int ns = 2;
    while (ns < WS_MAX_STAGES && (size_t)want * (BAR + (ns + 1) * stage + 1024) <= (size_t)220 * 1024) ++ns;
    if (ns_env >= 2 && ns_env <= WS_MAX_STAGES) ns = ns_env;
    while (ns > 2 && BAR + ns * stage > WS_SMEM_CAP) --ns;
    kern<<<ntiles, 96, BAR + ns * stage, st>>>(k, alias ? 1 : 0, ns);
    return b200rl_check_launch("k_target_rows_ws");
}

// Chunk width of the warp-specialised kernel: 256 B per row and fetch (32 timesteps at P = 2) or 512 B (64 timesteps).
// Wide chunks halve the chunk hand-offs per row and pay at both ends: from B200RL_TARGET_WS_WIDE_ROWS batch rows up
// (measured at B = 65,536: +3-9 % at T = 200, +2-4 % at T = 1024) and for few tiles - B <= 2,048: +4-13 % on all 24
// cells of T in {200, 512, 1024}; the head without rewards at B = 4,096: +0-5 % on 11 of 12 cells (with rewards it is
// mixed, -3 to +2 %).  In between they lose: B = 8,192 value head up to -14 %, B = 16,384 and T = 200 -15 %.
// Knob target_ws_row = 256 / 512 forces either (b200rl_set_tuning, so that a test can cover both in one process).
inline int ws_row_bytes(int B, bool has_rew) {
    long long v = 0;
    if (b200rl_tuning(B200RL_TUNE_TARGET_WS_ROW, &v)) return v >= 512 ? 512 : 256;
    if (B <= 2048 || (B < 6144 && !has_rew)) return 512;
    const long long wide = b200rl_tuning(B200RL_TUNE_TARGET_WS_WIDE_ROWS, &v) ? v : 32768;
    return B >= wide ? 512 : 256;
}

// Which lane mapping the warp-specialised kernel uses by default (measured: profiles/r2_target_ws_rowlane.md, section 7;
// both mappings run their pinned-order scans).  Row per lane (32-row tiles) pays where the compute warp is the critical
// path, i.e. few tiles and long rows.  Fraction of the HBM peak, sequence per lane | row per lane, return / value head:
//   1024 x 4,096: TD 0.81 / 0.71 | 0.89 / 0.76, UPGO 0.75 / 0.62 | 0.72 / 0.70, V-Trace 0.78 / 0.70 | 0.76 / 0.66,
//                 V-Trace + UPGO 0.74 / 0.65 | 0.69 / 0.57     (same pattern at T = 512 and 2048)
//   1024 x 8,192: every mode 0.73-0.77 | 0.77-0.81           (same at T = 512 and 2048)
// At T = 200, at B >= 16,384 and at B <= 2,048 row per lane is neutral or slower.
inline bool ws_rowlane_default(int B, int T, bool need_rho, bool upgo_with_rewards) {
    if (T < 512 || B <= 2048 || B > 8192) return false;
    if (B >= 6144) return true;
    return !need_rho && !upgo_with_rewards;
}

// long rows: the warp-specialised kernel (16-row tiles at P = 2); rows of <= 64 steps: one warp per 32 x 16 tile
template <int MT, int MA, int P, bool HAS_REW, bool NEED_RHO, bool RHO_PP>
int launch_rows(const ScanParams& k, cudaStream_t st) {
    if (k.T > B200RL_SHORT_T) {
        // knob target_ws_rowlane = 0 / 1: lane = (row, player) with 32/P-row tiles / lane = row with 32-row tiles
        long long rl = 0;
        const bool rowlane = P > 1 && (b200rl_tuning(B200RL_TUNE_TARGET_WS_ROWLANE, &rl) ? rl != 0 : ws_rowlane_default(k.B, k.T, NEED_RHO,
                                                      HAS_REW && (MT == B200RL_TARGET_UPGO || MA == B200RL_TARGET_UPGO)));
        if (rowlane) {
            if (ws_row_bytes(k.B, HAS_REW) == 512) return launch_ws<MT, MA, P, HAS_REW, NEED_RHO, RHO_PP, 128 / P, true>(k, st);
            return launch_ws<MT, MA, P, HAS_REW, NEED_RHO, RHO_PP, 64 / P, true>(k, st);
        }
        if (ws_row_bytes(k.B, HAS_REW) == 512) return launch_ws<MT, MA, P, HAS_REW, NEED_RHO, RHO_PP, 128 / P, false>(k, st);
        return launch_ws<MT, MA, P, HAS_REW, NEED_RHO, RHO_PP, 64 / P, false>(k, st);
    }
    return launch_shape<MT, MA, P, HAS_REW, NEED_RHO, RHO_PP, 32, 16>(k, st);
}

template <int MT, int MA, int P>
int launch_rows_p(const ScanParams& k, bool has_rew, bool rho_pp, cudaStream_t st) {
    constexpr bool NR = (MT == B200RL_TARGET_VTRACE || MA == B200RL_TARGET_VTRACE);
    if (NR && P > 1 && rho_pp) {
        return has_rew ? launch_rows<MT, MA, P, true, NR, (NR && P > 1)>(k, st)
                       : launch_rows<MT, MA, P, false, NR, (NR && P > 1)>(k, st);
    }
    return has_rew ? launch_rows<MT, MA, P, true, NR, false>(k, st) : launch_rows<MT, MA, P, false, NR, false>(k, st);
}

bool rows_ok(const View3& x, int W, int T) {  // contiguous (B, T, W), 16-byte aligned rows
    if (x.p == nullptr) return false;
    if (((uintptr_t)x.p % 16) != 0 || (x.sb % 4) != 0) return false;
    if (T > 1 && x.st != W) return false;
    if (W > 1 && x.sp != 1) return false;
    return true;
}

}  // namespace

int b200rl_launch_target_rows(int mt, int ma, const ScanParams& k, bool has_rew, cudaStream_t st) {
    const int P = k.P;
    if (P != 1 && P != 2) return -1;
    if ((k.T % 4) != 0 || k.B < 1) return -1;
    if (!rows_ok(k.v, P, k.T)) return -1;
    if (has_rew && !rows_ok(k.rew, P, k.T)) return -1;
    const bool need_rho = mt == B200RL_TARGET_VTRACE || ma == B200RL_TARGET_VTRACE;
    bool rho_pp = false;
    if (need_rho) {
        if (k.rho.px != k.c.px) return -1;
        rho_pp = P > 1 && k.rho.px == P;
        const int W = rho_pp ? P : 1;
        if (!rows_ok(k.rho, W, k.T) || !rows_ok(k.c, W, k.T)) return -1;
        if (!rho_pp && P > 1 && (k.rho.sp != 0 || k.c.sp != 0)) return -1;
    }
#define CASE(A, B_)                                                                      \
    if (mt == A && ma == B_)                                                             \
        return P == 1 ? launch_rows_p<A, B_, 1>(k, has_rew, rho_pp, st) : launch_rows_p<A, B_, 2>(k, has_rew, rho_pp, st);
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
    return -1;
}

// Both heads in one launch; -1 when the pair is not eligible (the caller then launches the heads one by one).
// A head without rewards has k.rew.p == nullptr.
int b200rl_launch_target_rows2(int mt, int ma, const ScanParams& k0, const ScanParams& k1, cudaStream_t st) {
    const int P = k0.P;
    if (P != 1 && P != 2) return -1;
    if (k0.B != k1.B || k0.T != k1.T || k1.P != P) return -1;
    if ((k0.T % 4) != 0 || k0.T > B200RL_SHORT_T || k0.B < 1) return -1;
    const bool need_rho = mt == B200RL_TARGET_VTRACE || ma == B200RL_TARGET_VTRACE;
    bool rho_pp = false;
    for (const ScanParams* k : {&k0, &k1}) {
        if (!rows_ok(k->v, P, k->T)) return -1;
        if (k->rew.p != nullptr && !rows_ok(k->rew, P, k->T)) return -1;
        if (need_rho) {
            if (k->rho.px != k->c.px) return -1;
            const bool pp = P > 1 && k->rho.px == P;
            if (k != &k0 && pp != rho_pp) return -1;
            rho_pp = pp;
            const int W = pp ? P : 1;
            if (!rows_ok(k->rho, W, k->T) || !rows_ok(k->c, W, k->T)) return -1;
            if (!pp && P > 1 && (k->rho.sp != 0 || k->c.sp != 0)) return -1;
        }
    }
#define CASE(A, B_)                                                                                         \
    if (mt == A && ma == B_) {                                                                              \
        constexpr bool NR = (A == B200RL_TARGET_VTRACE || B_ == B200RL_TARGET_VTRACE);                      \
        if (P == 1) return launch_dual<A, B_, 1, NR, false>(k0, k1, st);                                    \
        if (NR && rho_pp) return launch_dual<A, B_, 2, NR, NR>(k0, k1, st);                                 \
        return launch_dual<A, B_, 2, NR, false>(k0, k1, st);                                                \
    }
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
    return -1;
}
