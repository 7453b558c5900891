// Batched game engine kernels: one game per thread, 32 games per warp tile.
//
// Replaces, for N concurrent games, the per-game Python calls of the reference
// `Environment` interface (handyrl/environment.py:43-146) and the body of the
// self-play loop that drives it (handyrl/generation.py:31-68).
#include <cstdlib>

#include <atomic>

#include "common.cuh"
#include "games.cuh"
#include "emit.cuh"
#include "../../include/b200rl.h"

namespace {

constexpr int BLK = 64;            // threads per CTA = games per CTA: 65,536 games -> 1024 CTAs, 6.9 per SM
constexpr int WARPS = BLK / 32;

// Shared staging: per warp the flat bit strings of its 32 games (emit.cuh), per CTA the
// nibble -> float4 tables of the two value pairs that occur (0/1 and 1e32/0).
template <class G>
struct WarpStage {
    uint32_t obs[flat_words(32, G::OBITS)];
    uint32_t legal[flat_words(32, G::A)];
    uint32_t scalar[flat_words(32, G::SBITS > 0 ? G::SBITS : 1)];
};

template <class G>
struct CtaStage {
    float4 tbl_obs[16];    // observation planes / scalars: bit -> 1.0f, else 0.0f
    float4 tbl_amask[16];  // action_mask: legal bit -> 0.0f, else 1e32f (generation.py:50-51)
    WarpStage<G> w[WARPS];
    __device__ __forceinline__ void init() {
        nibble_table_init(tbl_obs, 0.f, 1.f, threadIdx.x);
        nibble_table_init(tbl_amask, 1e32f, 0.f, threadIdx.x);
        __syncthreads();
    }
};

template <class G>
__global__ void __launch_bounds__(BLK) k_reset(uint32_t* state, int64_t N, const uint8_t* __restrict__ mask) {
    int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x;
    if (i >= N) return;
    if (mask && !mask[i]) return;
    typename G::State s;
    G::reset(s);
    G::store(state, N, i, s);
}

template <class G>
__global__ void __launch_bounds__(BLK) k_step(uint32_t* state, int64_t N, const int32_t* __restrict__ actions,
                                              const int32_t* __restrict__ sel, float* __restrict__ out_reward,
                                              uint8_t* __restrict__ out_terminal) {
    int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x;
    if (i >= N) return;
    typename G::State s = G::load(state, N, i);
    const int a = actions[i];
    if (a >= 0) {
        G::step(s, a, sel ? sel[i] : 0);
        G::store(state, N, i, s);
    }
    if (out_reward) {
        const float r = G::reward(s);
        out_reward[2 * i] = r;
        out_reward[2 * i + 1] = r;
    }
    if (out_terminal) out_terminal[i] = G::terminal(s) ? 1 : 0;
}

template <class G>
__global__ void __launch_bounds__(BLK) k_legal_list(const uint32_t* state, int64_t N, int32_t* __restrict__ out_ids,
                                                    int32_t* __restrict__ out_count) {
    int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x;
    if (i >= N) return;
    typename G::State s = G::load(state, N, i);
    int ids[G::LMAX];
    int n = G::legal_list(s, ids);
    for (int k = 0; k < G::LMAX; ++k) out_ids[i * G::LMAX + k] = k < n ? ids[k] : -1;
    out_count[i] = n;
}

template <class G>
__global__ void __launch_bounds__(BLK) k_outcome(const uint32_t* state, int64_t N, float* __restrict__ out) {
    int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x;
    if (i >= N) return;
    typename G::State s = G::load(state, N, i);
    float o0, o1;
    G::outcome(s, o0, o1);
    out[2 * i] = o0;
    out[2 * i + 1] = o1;
}

template <class G>
__global__ void __launch_bounds__(BLK) k_info(const uint32_t* state, int64_t N, int32_t* __restrict__ out) {
    int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x;
    if (i >= N) return;
    typename G::State s = G::load(state, N, i);
    float o0, o1;
    G::outcome(s, o0, o1);
    out[4 * i + 0] = G::turn(s);
    out[4 * i + 1] = G::terminal(s) ? 1 : 0;
    out[4 * i + 2] = G::steps(s);
    out[4 * i + 3] = G::terminal(s) ? (o0 > 0 ? 1 : (o1 > 0 ? 2 : 3)) : 0;
}

constexpr int RO_CH = 4096, RO_NBUF = 3;   // staging of the bulk-store emit path of k_rollout<G, true>

// Pack this lane's bits into the warp's flat strings, then let the warp stream the tile out.
template <class G>
__device__ __forceinline__ void emit_views(CtaStage<G>& cta, int warp, const typename G::State& s, bool valid,
                                           int view, int64_t g0, int ng, int lane, float* out_main,
                                           float* out_scalar, float* out_amask, uint32_t* out_bits,
                                           TmaStage<RO_CH, RO_NBUF>* tma = nullptr) {
    WarpStage<G>& st = cta.w[warp];
    constexpr int SB = G::SBITS > 0 ? G::SBITS : 1;
    const bool want_obs = out_main || (G::SBITS > 0 && out_scalar);
    const bool want_legal = out_amask || out_bits;
    uint32_t ob[G::OW], sc[1] = {0}, lm[G::LW];
#pragma unroll
    for (int k = 0; k < G::OW; ++k) ob[k] = 0;
#pragma unroll
    for (int k = 0; k < G::LW; ++k) lm[k] = 0;
    if constexpr (G::OBITS < 32 && G::A < 32 && G::SBITS == 0) {
        // TicTacToe: the whole tile is one register per lane, packed and streamed with shuffles (emit.cuh)
        if (!tma) {
            if (valid) {
                if (want_obs) G::obs_bits(s, view, ob, sc[0]);
                if (want_legal) G::legal_bits(s, lm);
                if (out_bits) out_bits[g0 + lane] = lm[0];
            }
            if (out_main) emit_short_rows_f32<G::OBITS>(ob[0], cta.tbl_obs, out_main + g0 * G::OBITS, ng, lane, 0.f, 1.f);
            if (out_amask) emit_short_rows_f32<G::A>(lm[0], cta.tbl_amask, out_amask + g0 * G::A, ng, lane, 1e32f, 0.f);
            return;
        }
    }
    if (valid) {
        if (want_obs) G::obs_bits(s, view, ob, sc[0]);
        if (want_legal) G::legal_bits(s, lm);
        const bool last = lane == ng - 1;
        if (G::OBITS < 32 && out_main) flat_clear_row<G::OBITS>(st.obs, lane, last);
        if (G::SBITS > 0 && out_scalar) flat_clear_row<SB>(st.scalar, lane, last);
        if (G::A < 32 && out_amask) flat_clear_row<G::A>(st.legal, lane, last);
    }
    if (G::OBITS < 32 || G::SBITS > 0 || G::A < 32) __syncwarp();
    // rows of >= 32 bits: shuffle-merged packing (all lanes participate); shorter rows: atomicOr packing
    if (G::OBITS >= 32) {
        if (out_main) flat_pack_warp<(G::OBITS >= 32 ? G::OBITS : 32), G::OW>(st.obs, lane, ng, ob);
    } else if (valid && out_main) {
        flat_put<G::OBITS, G::OW>(st.obs, lane, ob);
    }
    if (G::A >= 32) {
        if (out_amask) flat_pack_warp<(G::A >= 32 ? G::A : 32), G::LW>(st.legal, lane, ng, lm);
    } else if (valid && out_amask) {
        flat_put<G::A, G::LW>(st.legal, lane, lm);
    }
    if (valid) {
        if (G::SBITS > 0 && out_scalar) flat_put<SB, 1>(st.scalar, lane, sc);
        if (out_bits) {
#pragma unroll
            for (int k = 0; k < G::LW; ++k) out_bits[(g0 + lane) * G::LW + k] = lm[k];
        }
    }
    __syncwarp();
    if (tma) {   // staged bulk stores (emit.cuh); the flat strings are free again when this returns
        if (out_main) emit_flat_f32_tma<G::OBITS>(st.obs, out_main + g0 * G::OBITS, ng, lane, 0.f, 1.f, *tma);
        if (G::SBITS > 0 && out_scalar) emit_flat_f32_tma<SB>(st.scalar, out_scalar + g0 * G::SBITS, ng, lane, 0.f, 1.f, *tma);
        if (out_amask) emit_flat_f32_tma<G::A>(st.legal, out_amask + g0 * G::A, ng, lane, 1e32f, 0.f, *tma);
        __syncwarp();
        return;
    }
    if (out_main) emit_flat_f32<G::OBITS>(st.obs, cta.tbl_obs, out_main + g0 * G::OBITS, ng, lane, 0.f, 1.f);
    if (G::SBITS > 0 && out_scalar)
        emit_flat_f32<SB>(st.scalar, cta.tbl_obs, out_scalar + g0 * G::SBITS, ng, lane, 0.f, 1.f);
    // action_mask = 1e32 on illegal, 0 on legal (generation.py:50-51)
    if (out_amask) emit_flat_f32<G::A>(st.legal, cta.tbl_amask, out_amask + g0 * G::A, ng, lane, 1e32f, 0.f);
    __syncwarp();
}

template <class G>
__global__ void __launch_bounds__(BLK) k_observe(const uint32_t* state, int64_t N, int view, const int32_t* views,
                                                 float* out_main, float* out_scalar, float* out_amask,
                                                 uint32_t* out_bits) {
    __shared__ CtaStage<G> stage;
    stage.init();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t g0 = (blockIdx.x * (int64_t)WARPS + warp) * 32;
    if (g0 >= N) return;
    const int ng = (int)min((int64_t)32, N - g0);
    const bool valid = lane < ng;
    typename G::State s;
    G::reset(s);
    if (valid) s = G::load(state, N, g0 + lane);
    const int v = (views && valid) ? views[g0 + lane] : view;
    emit_views<G>(stage, warp, s, valid, v, g0, ng, lane, out_main, out_scalar, out_amask, out_bits);
}

// Fused ply: apply the chosen actions, report reward / terminal / outcome, optionally
// restart finished games, then encode the next observation (turn player's view) and
// legal mask — the body of generation.py:31-68 for N games in one launch.
template <class G>
__global__ void __launch_bounds__(BLK)
k_step_observe(uint32_t* state, int64_t N, const int32_t* __restrict__ actions, const int32_t* __restrict__ sel,
               int auto_reset, float* out_main, float* out_scalar, float* out_amask, float* __restrict__ out_reward,
               uint8_t* __restrict__ out_terminal, float* __restrict__ out_outcome, int32_t* __restrict__ out_turn) {
    __shared__ CtaStage<G> stage;
    stage.init();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t g0 = (blockIdx.x * (int64_t)WARPS + warp) * 32;
    if (g0 >= N) return;
    const int ng = (int)min((int64_t)32, N - g0);
    const bool valid = lane < ng;
    const int64_t i = g0 + lane;
    typename G::State s;
    G::reset(s);
    if (valid) {
        s = G::load(state, N, i);
        const int a = actions ? actions[i] : -1;
        if (a >= 0) G::step(s, a, sel ? sel[i] : 0);
        const bool term = G::terminal(s);
        if (out_reward) {
            const float r = G::reward(s);
            out_reward[2 * i] = r;
            out_reward[2 * i + 1] = r;
        }
        if (out_terminal) out_terminal[i] = term ? 1 : 0;
        if (out_outcome) {
            float o0, o1;
            G::outcome(s, o0, o1);
            out_outcome[2 * i] = o0;
            out_outcome[2 * i + 1] = o1;
        }
        if (term && auto_reset) G::reset(s);
        if (a >= 0 || (term && auto_reset)) G::store(state, N, i, s);
        if (out_turn) out_turn[i] = G::turn(s);
    }
    emit_views<G>(stage, warp, s, valid, B200RL_VIEW_TURN, g0, ng, lane, out_main, out_scalar, out_amask, nullptr);
}

// position of the k-th (0-based) set bit of x, k < popc(x): branch-free binary search on popcounts (the
// clear-lowest-bit loop it replaces diverged inside the warp and was 8 % of the TicTacToe rollout's instructions)
template <int BITS>
__device__ __forceinline__ int select_bit(uint32_t x, int k) {
    int pos = 0;
#pragma unroll
    for (int h = 16; h >= 1; h >>= 1) {
        if (h < BITS) {
            const int c = __popc(x & ((1u << h) - 1u));
            const bool up = k >= c;
            k -= up ? c : 0;
            pos += up ? h : 0;
            x >>= up ? h : 0;
        }
    }
    return pos;
}

// k-th (0-based) set bit of an LW-word mask of NBITS valid bits
template <int LW, int NBITS>
__device__ __forceinline__ int kth_set_bit(const uint32_t (&m)[LW], int k) {
    if (LW == 1) return select_bit<(NBITS <= 16 ? 16 : 32)>(m[0], k);
    // find the word, then the bit inside it
    int base = 0;
    uint32_t x = m[0];
    bool found = false;
#pragma unroll
    for (int w = 0; w < LW; ++w) {
        const int c = __popc(m[w]);
        if (!found) {
            if (k < c) {
                found = true;
                x = m[w];
                base = 32 * w;
            } else {
                k -= c;
            }
        }
    }
    return found ? base + select_bit<32>(x, k) : -1;
}

struct RolloutParams {
    uint32_t* state;
    int64_t N;
    int64_t n_plies;
    uint64_t seed, game_id_base, it0;
    float* out_main;    // [waves][N][OBITS] or null
    float* out_scalar;  // [waves][N][SBITS] or null
    float* out_amask;   // [waves][N][A] or null
    int32_t* out_action;  // [waves][N] or null
    int64_t waves;
    uint4* traj;        // [N][D] 32-byte compact records (slot-major ring) or null
    int64_t D;
    int32_t* ep_table;  // [E][4] = slot, steps | win << 16, end_it lo, end_it hi
    int64_t E;
    unsigned long long* counters;  // [0] episodes, [1] env steps, [2] BLACK wins, [3] WHITE wins, [4] draws
    unsigned int* sched;   // [0] ticket counter, [SCHED_HDR + g] chunks completed for game group g (zeroed per launch)
    int64_t chunk;         // plies per ticket
};
constexpr int SCHED_HDR = 32;

__device__ __forceinline__ unsigned int ld_acquire_u32(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_u32(unsigned int* p, unsigned int v) {
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

constexpr int EP_BUF = 64;   // finished-episode rows buffered per warp before one append to the table

// append n buffered rows to the episode-table ring: one atomic on the head, then coalesced row stores
__device__ __forceinline__ void ep_flush(const RolloutParams& p, const int4* buf, int n, int lane) {
    __syncwarp();
    unsigned long long base = 0;
    if (lane == 0) base = atomicAdd(&p.counters[5], (unsigned long long)n);
    base = __shfl_sync(0xffffffffu, base, 0);
    for (int r = lane; r < n; r += 32)
        *reinterpret_cast<int4*>(p.ep_table + (int64_t)ring_mod(base + (unsigned long long)r, (uint64_t)p.E) * 4) = buf[r];
    __syncwarp();
}

// Persistent self-play with the epoch-0 policy (uniform over legal actions ==
// RandomModel's zero logits, handyrl/model.py:58-64): every ply emits the reference-format moment data
// (observation, action_mask, action) and/or a 32-byte compact trajectory record.
//
// Work is handed out as TICKETS (group of 32 games, chunk of `chunk` plies) from a global counter: ticket
// t = chunk c of group g = t % ngroups.  Within a ticket the 32 game states live in registers; between the
// chunks of a group they go through global memory (64 B per game) and a per-group progress word
// (release/acquire).  Why: the SMs of a B200 differ by up to 2x in the HBM bandwidth they can draw
// (tools/probe_write4.cu), so a static games->warp assignment finishes at the pace of the slowest SM (a pure
// write stream: 6.3 TB/s static, 7.2 TB/s with dynamic assignment); with tickets the fast SMs take more chunks.
// A ticket only ever waits for a ticket that a RUNNING warp claimed earlier, so there is no deadlock whatever
// the number of resident CTAs.  Results depend on (game, iteration) only, not on which warp ran a ticket.
// TMA: the moment tensors leave through staged bulk copies (cp.async.bulk, emit.cuh) instead of STG.128; dynamic
// shared memory then holds WARPS x RO_NBUF x RO_CH bytes of staging.  Besides moving the stores off the LSU pipe
// this takes them out of the scope of the ticket fence (__threadfence orders the thread's own stores - the 64-byte
// states other tickets wait for -, not the copy engine's writes, which nobody reads before the kernel ends).
template <class G, bool TMA>
__global__ void __launch_bounds__(BLK, TMA ? 5 : 8) k_rollout(const RolloutParams p) {
    __shared__ CtaStage<G> stage;
    extern __shared__ __align__(128) unsigned char ro_dyn_smem[];
    TmaStage<RO_CH, RO_NBUF> tma_stage;
    tma_stage.buf = reinterpret_cast<float4*>(ro_dyn_smem) + (size_t)(threadIdx.x >> 5) * RO_NBUF * (RO_CH / 16);
    tma_stage.c = 0;
    // write combining of the trajectory records: 4 plies x 32 B per game = one 128-byte line of the slot-major
    // ring, staged as [ply & 3][half][lane] 16-byte units (conflict-free) and flushed every 4th ply
    __shared__ uint4 rec_stage[WARPS][4][2][32];
    stage.init();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t ngroups = (p.N + 31) / 32;
    const int64_t nchunks = (p.n_plies + p.chunk - 1) / p.chunk;
    const int64_t ntickets = ngroups * nchunks;
    unsigned int n_ep = 0, n_w0 = 0, n_w1 = 0, n_dr = 0;
    unsigned long long n_steps = 0;
    uint32_t rec_fill = 0;   // bit q: rec_stage[..][q] holds the record of the current 4-ply group
    __shared__ int4 ep_buf[WARPS][EP_BUF];
    int ep_n = 0;            // rows waiting in ep_buf[warp] (warp-uniform)

    int64_t pend_grp = -1;       // group whose progress word still has to be published (warp-uniform)
    unsigned int pend_val = 0;
    auto publish = [&]() {
        if (pend_grp >= 0) {
            __threadfence();     // every lane: its state stores before the flag
            __syncwarp();
            if (lane == 0) st_release_u32(p.sched + SCHED_HDR + pend_grp, pend_val);
            pend_grp = -1;
        }
    };
    for (;;) {
        unsigned int tk = 0;
        if (lane == 0) tk = atomicAdd(&p.sched[0], 1u);
        tk = __shfl_sync(0xffffffffu, tk, 0);
        if ((int64_t)tk >= ntickets) break;
        const int64_t ch = (int64_t)tk / ngroups, grp = (int64_t)tk - ch * ngroups;
        const int64_t g0 = grp * 32;
        const int ng = (int)min((int64_t)32, p.N - g0);
        const bool valid = lane < ng;
        const int64_t i = g0 + lane;
        if (ch > 0) {   // the previous chunk of this group must have stored its states
            const unsigned int* flag = p.sched + SCHED_HDR + grp;
            bool ready = true;
            if (lane == 0) ready = ld_acquire_u32(flag) >= (unsigned int)ch;
            ready = __shfl_sync(0xffffffffu, ready ? 1 : 0, 0) != 0;
            if (!ready) {
                publish();   // never wait while holding back a publication of our own (it may be what others need)
                if (lane == 0) {
                    unsigned int spins = 0;
                    while (ld_acquire_u32(flag) < (unsigned int)ch) {
                        __nanosleep(100);
                        if (++spins > (1u << 24)) break;   // ~2 s: never hang the GPU on a scheduling bug
                    }
                }
                __syncwarp();
            }
            (void)ld_acquire_u32(flag);   // every lane acquires for itself (a fence here would also wait for OUR stores)
        }
        typename G::State s;
        G::reset(s);
        if (valid) s = G::load_cg(p.state, p.N, i);
        // Publication of the PREVIOUS ticket, deferred to here: its fence has to wait until that ticket's ~256 KB of
        // streaming stores are performed; issued right after them it was 13 % of the kernel's stall samples, issued
        // behind the ticket take and this state load most of that drain has already happened.
        publish();
        const uint64_t stream_id = p.game_id_base + (uint64_t)i;
        const int64_t k_begin = ch * p.chunk, k_end = min(k_begin + p.chunk, p.n_plies);

        // Wave of the current ply and the element offsets of its output tiles.  The general form derives them from
        // the iteration number every ply (a remainder and three 64-bit products: noise for Geister, whose warp writes
        // 60 KB per ply); games with rows shorter than a word (TicTacToe: 5.6 KB per warp and ply, bound by
        // instruction issue at 2,048 warps - profiles/r2_ttt_rollout.md) advance them by one wave per ply instead.
        constexpr bool LIGHT = G::OBITS < 32 && G::SBITS == 0;
        int64_t wv = 0, wv_main = 0, wv_amask = 0, wv_act = 0;
        if constexpr (LIGHT) {
            wv = p.waves > 0 ? (int64_t)ring_mod(p.it0 + (uint64_t)k_begin, (uint64_t)p.waves) : 0;
            wv_main = wv * p.N * G::OBITS;
            wv_amask = wv * p.N * G::A;
            wv_act = wv * p.N;
        }
        for (int64_t k = k_begin; k < k_end; ++k) {
            const uint64_t it = p.it0 + (uint64_t)k;
            bool first = false;
            if (valid && G::terminal(s)) G::reset(s);  // auto-reset: a finished slot starts its next episode
            if (valid) first = G::steps(s) == 0;
            // policy: uniform over the legal actions
            uint32_t lm[G::LW];
#pragma unroll
            for (int w = 0; w < G::LW; ++w) lm[w] = 0;
            int action = -1, selp = 0;
            if (valid) {
                G::legal_bits(s, lm);
                int cnt = 0;
#pragma unroll
                for (int w = 0; w < G::LW; ++w) cnt += __popc(lm[w]);
                const uint32_t r = b200rl_rand32(p.seed, stream_id, it);
                const int kk = (int)__umulhi(r, (uint32_t)cnt);
                action = kth_set_bit<G::LW, G::A>(lm, kk);
                selp = (int)((r >> 7) & 1);  // ParallelTicTacToe: which player's move lands
            }
            if constexpr (!LIGHT) wv = p.waves > 0 ? (int64_t)ring_mod(it, (uint64_t)p.waves) : 0;
            if (p.out_main || p.out_amask || p.out_scalar) {
                emit_views<G>(stage, warp, s, valid, B200RL_VIEW_TURN, g0, ng, lane,
                              p.out_main ? p.out_main + (LIGHT ? wv_main : wv * p.N * G::OBITS) : nullptr,
                              p.out_scalar ? p.out_scalar + wv * p.N * G::SBITS : nullptr,
                              p.out_amask ? p.out_amask + (LIGHT ? wv_amask : wv * p.N * G::A) : nullptr, nullptr,
                              TMA ? &tma_stage : nullptr);
            }
            bool term = false;
            uint32_t wincode = 0;
            if (valid) {
                if (p.out_action) p.out_action[(LIGHT ? wv_act : wv * p.N) + i] = action;
                uint32_t rec[8];
                if (p.traj) G::pack_record(s, rec);
                G::step(s, action, selp);
                term = G::terminal(s);
                float o0 = 0.f, o1 = 0.f;
                if (term) G::outcome(s, o0, o1);
                wincode = term ? (o0 > 0 ? 1u : (o1 > 0 ? 2u : 3u)) : 0u;
                if (p.traj) {
                    rec[5] = ((uint32_t)action & 0xFFFF) | ((uint32_t)selp << 16) | ((first ? 1u : 0u) << 24) |
                             ((term ? 1u : 0u) << 25) | (wincode << 26);
                    rec[6] = (uint32_t)G::steps(s);
                    rec[7] = 0;
                    const int q = (int)(it & 3);
                    rec_stage[warp][q][0][lane] = make_uint4(rec[0], rec[1], rec[2], rec[3]);
                    rec_stage[warp][q][1][lane] = make_uint4(rec[4], rec[5], rec[6], rec[7]);
                }
                if (term) {
                    n_ep += 1;
                    n_w0 += wincode == 1;
                    n_w1 += wincode == 2;
                    n_dr += wincode == 3;
                }
            }
            if (p.ep_table) {
                // episode-table append: finished games' rows are collected in a per-warp shared buffer and appended
                // EP_BUF at a time (one atomic on the table head + coalesced 16-byte row stores).  TicTacToe finishes
                // ~4 of a warp's 32 games every ply; waiting for one returning global atomic per ply was 12 % of that
                // rollout's stall samples, and before that 3 G same-address atomics per second were its bound.
                const unsigned int fin = __ballot_sync(0xffffffffu, term);
                if (fin) {
                    const int nf = __popc(fin);
                    if (ep_n + nf > EP_BUF) {
                        ep_flush(p, ep_buf[warp], ep_n, lane);
                        ep_n = 0;
                    }
                    if (term) {
                        int4 row;
                        row.x = (int32_t)i;
                        row.y = (int32_t)((uint32_t)G::steps(s) | (wincode << 16));
                        row.z = (int32_t)(uint32_t)it;
                        row.w = (int32_t)(uint32_t)(it >> 32);
                        ep_buf[warp][ep_n + __popc(fin & ((1u << lane) - 1u))] = row;
                    }
                    ep_n += nf;
                }
            }
            if (p.traj) {
                // flush the 4-ply group, transposed: lanes 8q..8q+7 carry ply q of 8 consecutive games, so every
                // game's (up to) 128-byte line leaves the SM as one request of full 32-byte sectors
                rec_fill |= 1u << (int)(it & 3);
                if ((it & 3) == 3 || k == k_end - 1) {
                    __syncwarp();
                    const int q = lane >> 3;
                    if (rec_fill & (1u << q)) {
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const int gq = 8 * j + (lane & 7);
                            if (gq < ng) {
                                const uint4 lo = rec_stage[warp][q][0][gq], hi = rec_stage[warp][q][1][gq];
                                const uint32_t r8[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
                                st_record(p.traj, p.D, p.N, (it & ~(uint64_t)3) + (uint64_t)q, g0 + gq, r8);
                            }
                        }
                    }
                    rec_fill = 0;
                    __syncwarp();
                }
            }
            if constexpr (LIGHT) {
                if (++wv >= p.waves) {
                    wv = wv_main = wv_amask = wv_act = 0;
                } else {
                    wv_main += p.N * G::OBITS;
                    wv_amask += p.N * G::A;
                    wv_act += p.N;
                }
            }
        }
        if (p.ep_table && ep_n) {
            ep_flush(p, ep_buf[warp], ep_n, lane);
            ep_n = 0;
        }
        if (valid) G::store(p.state, p.N, i, s);
        n_steps += (unsigned long long)ng * (unsigned long long)(k_end - k_begin);
        if (nchunks > 1) {   // to publish: states of this group are valid up to chunk ch
            pend_grp = grp;
            pend_val = (unsigned int)(ch + 1);
        }
    }
    publish();
    if (TMA) tma_stage.drain(lane);   // the copy engine still reads the staging buffers of the last chunks
    if (p.counters) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            n_ep += __shfl_xor_sync(0xffffffffu, n_ep, o);
            n_w0 += __shfl_xor_sync(0xffffffffu, n_w0, o);
            n_w1 += __shfl_xor_sync(0xffffffffu, n_w1, o);
            n_dr += __shfl_xor_sync(0xffffffffu, n_dr, o);
        }
        if (lane == 0) {
            atomicAdd(&p.counters[0], (unsigned long long)n_ep);
            atomicAdd(&p.counters[1], n_steps);
            atomicAdd(&p.counters[2], (unsigned long long)n_w0);
            atomicAdd(&p.counters[3], (unsigned long long)n_w1);
            atomicAdd(&p.counters[4], (unsigned long long)n_dr);
        }
    }
}

inline int grid_for(int64_t N) { return (int)((N + BLK - 1) / BLK); }

// persistent grid: as many CTAs as are resident at once (8 per SM for Geister at 128 registers, more for TicTacToe)
template <class G>
void launch_rollout(const RolloutParams& p, int64_t want, int dev, int sms, cudaStream_t stream) {
    // Staged bulk stores are OFF by default here: measured on the bench workload they are 1 % SLOWER than STG.128
    // (0.641 -> 0.648 ms, profiles/r2_tma_emit_ab.md) although the isolated emit probe gains 4-6 %
    // (profiles/r2_probe_write_tma.txt) - the 168-register variant holds 10 warps per SM instead of 16 and the
    // rollout is not bound by its store instructions.  Knob emit_tma = 1 selects them (tests cover both).
    long long knob = 0;
    const bool has_out = p.out_main || p.out_scalar || p.out_amask;
    const bool tma = has_out && b200rl_tuning(B200RL_TUNE_EMIT_TMA, &knob) && knob != 0;
    constexpr int DYN = WARPS * RO_NBUF * RO_CH;
    static std::atomic<int> per_sm_dev[64][2];   // per device: the occupancy query is a property of (kernel, device)
    std::atomic<int>& slot = per_sm_dev[dev & 63][tma ? 1 : 0];
    int per_sm = slot.load(std::memory_order_relaxed);
    if (per_sm == 0) {
        int n = 0;
        cudaError_t e;
        if (tma) {
            cudaFuncSetAttribute(k_rollout<G, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, DYN);
            e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_rollout<G, true>, BLK, DYN);
        } else {
            e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_rollout<G, false>, BLK, 0);
        }
        if (e != cudaSuccess || n < 1) n = tma ? 5 : 8;
        per_sm = n;
        slot.store(n, std::memory_order_relaxed);
    }
    const int grid = (int)min(want, (int64_t)sms * per_sm);
    if (tma) k_rollout<G, true><<<grid, BLK, DYN, stream>>>(p);
    else k_rollout<G, false><<<grid, BLK, 0, stream>>>(p);
}

}  // namespace

#define DISPATCH_GAME(game, EXPR)                                                       \
    switch (game) {                                                                     \
        case B200RL_GAME_TICTACTOE: { using G = TicTacToeGame<false>; EXPR; } break;     \
        case B200RL_GAME_PARALLEL_TICTACTOE: { using G = TicTacToeGame<true>; EXPR; } break; \
        case B200RL_GAME_GEISTER: { using G = GeisterGame<false>; EXPR; } break;         \
        case B200RL_GAME_CIGEISTER: { using G = GeisterGame<true>; EXPR; } break;        \
        default:                                                                        \
            b200rl_set_error("unknown game id %d", game);                               \
            return B200RL_ERR_ARG;                                                      \
    }

#define REQ_STATE(state, N)                                                                       \
    B200RL_REQUIRE((state) != nullptr && (N) >= 0, "env: null state or negative N");             \
    B200RL_REQUIRE(((uintptr_t)(state) % 16) == 0, "env: state must be 16-byte aligned");         \
    if ((N) == 0) return B200RL_OK;

extern "C" int64_t b200rl_env_state_bytes(int game) {
    switch (game) {
        case B200RL_GAME_TICTACTOE:
        case B200RL_GAME_PARALLEL_TICTACTOE: return 4;
        case B200RL_GAME_GEISTER:
        case B200RL_GAME_CIGEISTER: return 64;
    }
    b200rl_set_error("unknown game id %d", game);
    return -1;
}

extern "C" int b200rl_env_dims(int game, int64_t* out) {
    B200RL_REQUIRE(out != nullptr, "env_dims: null out");
    DISPATCH_GAME(game, {
        out[0] = G::A; out[1] = G::OBITS; out[2] = G::SBITS; out[3] = G::LMAX; out[4] = G::LW;
        out[5] = G::PLAYERS; out[6] = G::MAX_STEPS; out[7] = G::HAS_REWARD ? 1 : 0;
    });
    return B200RL_OK;
}

extern "C" int b200rl_env_reset(int game, void* state, int64_t N, const uint8_t* slot_mask, void* stream) {
    REQ_STATE(state, N);
    DISPATCH_GAME(game, (k_reset<G><<<grid_for(N), BLK, 0, (cudaStream_t)stream>>>((uint32_t*)state, N, slot_mask)));
    return b200rl_check_launch("k_reset");
}

extern "C" int b200rl_env_step(int game, void* state, int64_t N, const int32_t* actions, const int32_t* selected,
                               float* out_reward, uint8_t* out_terminal, void* stream) {
    REQ_STATE(state, N);
    B200RL_REQUIRE(actions != nullptr, "env_step: null actions");
    DISPATCH_GAME(game, (k_step<G><<<grid_for(N), BLK, 0, (cudaStream_t)stream>>>((uint32_t*)state, N, actions,
                                                                                  selected, out_reward, out_terminal)));
    return b200rl_check_launch("k_step");
}

extern "C" int b200rl_env_legal_list(int game, const void* state, int64_t N, int32_t* out_ids, int32_t* out_count,
                                     void* stream) {
    REQ_STATE(state, N);
    B200RL_REQUIRE(out_ids && out_count, "env_legal_list: null output");
    DISPATCH_GAME(game, (k_legal_list<G><<<grid_for(N), BLK, 0, (cudaStream_t)stream>>>((const uint32_t*)state, N,
                                                                                        out_ids, out_count)));
    return b200rl_check_launch("k_legal_list");
}

extern "C" int b200rl_env_outcome(int game, const void* state, int64_t N, float* out, void* stream) {
    REQ_STATE(state, N);
    B200RL_REQUIRE(out != nullptr, "env_outcome: null output");
    DISPATCH_GAME(game, (k_outcome<G><<<grid_for(N), BLK, 0, (cudaStream_t)stream>>>((const uint32_t*)state, N, out)));
    return b200rl_check_launch("k_outcome");
}

extern "C" int b200rl_env_info(int game, const void* state, int64_t N, int32_t* out, void* stream) {
    REQ_STATE(state, N);
    B200RL_REQUIRE(out != nullptr, "env_info: null output");
    DISPATCH_GAME(game, (k_info<G><<<grid_for(N), BLK, 0, (cudaStream_t)stream>>>((const uint32_t*)state, N, out)));
    return b200rl_check_launch("k_info");
}

static bool aligned16(const void* p) { return ((uintptr_t)p % 16) == 0; }

extern "C" int b200rl_env_observe(int game, const void* state, int64_t N, int view, const int32_t* views,
                                  float* out_main, float* out_scalar, float* out_amask, uint32_t* out_bits,
                                  void* stream) {
    REQ_STATE(state, N);
    B200RL_REQUIRE(view >= -1 && view <= 2, "env_observe: view must be -1 (None), 0, 1 or 2 (turn player)");
    B200RL_REQUIRE(aligned16(out_main) && aligned16(out_scalar) && aligned16(out_amask),
                   "env_observe: outputs must be 16-byte aligned");
    DISPATCH_GAME(game, (k_observe<G><<<grid_for(N), BLK, 0, (cudaStream_t)stream>>>(
                            (const uint32_t*)state, N, view, views, out_main, out_scalar, out_amask, out_bits)));
    return b200rl_check_launch("k_observe");
}

extern "C" int b200rl_env_step_observe(int game, void* state, int64_t N, const int32_t* actions,
                                       const int32_t* selected, int auto_reset, float* out_main, float* out_scalar,
                                       float* out_amask, float* out_reward, uint8_t* out_terminal,
                                       float* out_outcome, int32_t* out_turn, void* stream) {
    REQ_STATE(state, N);
    B200RL_REQUIRE(aligned16(out_main) && aligned16(out_scalar) && aligned16(out_amask),
                   "env_step_observe: outputs must be 16-byte aligned");
    DISPATCH_GAME(game, (k_step_observe<G><<<grid_for(N), BLK, 0, (cudaStream_t)stream>>>(
                            (uint32_t*)state, N, actions, selected, auto_reset, out_main, out_scalar, out_amask,
                            out_reward, out_terminal, out_outcome, out_turn)));
    return b200rl_check_launch("k_step_observe");
}

// ticket scratch of one rollout launch: header + one progress word per group of 32 games
extern "C" int64_t b200rl_env_rollout_scratch_bytes(int64_t N) {
    if (N < 0) return -1;
    return (int64_t)(SCHED_HDR + (N + 31) / 32) * (int64_t)sizeof(unsigned int);
}

extern "C" int b200rl_env_rollout(int game, void* state, int64_t N, int64_t n_plies, uint64_t seed,
                                  uint64_t game_id_base, uint64_t it0, float* out_main, float* out_scalar,
                                  float* out_amask, int32_t* out_action, int64_t waves, void* traj, int64_t D,
                                  int32_t* ep_table, int64_t E, uint64_t* counters, void* scratch,
                                  int64_t scratch_bytes, void* stream) {
    REQ_STATE(state, N);
    B200RL_REQUIRE(n_plies >= 0, "env_rollout: negative n_plies");
    B200RL_REQUIRE(scratch != nullptr && ((uintptr_t)scratch % 16) == 0 &&
                       scratch_bytes >= b200rl_env_rollout_scratch_bytes(N),
                   "env_rollout: the caller-owned ticket scratch must hold b200rl_env_rollout_scratch_bytes(N) bytes, "
                   "16-byte aligned");
    B200RL_REQUIRE(aligned16(out_main) && aligned16(out_scalar) && aligned16(out_amask) && ((uintptr_t)traj % 32) == 0,
                   "env_rollout: outputs must be 16-byte aligned (trajectory ring: 32)");
    B200RL_REQUIRE(!(out_main || out_scalar || out_amask || out_action) || waves >= 1, "env_rollout: waves must be >= 1");
    B200RL_REQUIRE(!traj || D >= 1, "env_rollout: trajectory ring depth D must be >= 1");
    B200RL_REQUIRE(!ep_table || (E >= 1 && counters), "env_rollout: episode table needs E >= 1 and counters");
    B200RL_REQUIRE(((uintptr_t)ep_table % 16) == 0, "env_rollout: episode table must be 16-byte aligned");
    if (waves > 1 && (out_main || out_scalar || out_amask)) {
        // wave w of a moment tensor starts at w * N * row floats and is written with 128-bit stores: every wave base
        // must be 16-byte aligned (Geister: N even; TicTacToe, 27 + 9 floats per game: N a multiple of 4)
        int64_t dims[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (b200rl_env_dims(game, dims) != B200RL_OK) return B200RL_ERR_ARG;
        const int64_t A = dims[0], obits = dims[1], sbits = dims[2];
        B200RL_REQUIRE((!out_main || (N * obits) % 4 == 0) && (!out_scalar || (N * sbits) % 4 == 0) &&
                           (!out_amask || (N * A) % 4 == 0),
                       "env_rollout: with several waves every wave base must be 16-byte aligned: N * floats-per-game "
                       "(observation %lld, scalar %lld, action_mask %lld) must be a multiple of 4; N = %lld is not",
                       (long long)obits, (long long)sbits, (long long)A, (long long)N);
    }
    RolloutParams p;
    p.state = (uint32_t*)state; p.N = N; p.n_plies = n_plies; p.seed = seed; p.game_id_base = game_id_base;
    p.it0 = it0; p.out_main = out_main; p.out_scalar = out_scalar; p.out_amask = out_amask;
    p.out_action = out_action; p.waves = waves; p.traj = (uint4*)traj; p.D = D; p.ep_table = ep_table; p.E = E;
    p.counters = (unsigned long long*)counters;
    if (N <= 0 || n_plies <= 0) return B200RL_OK;
    // tickets: 4-ply chunks (one 128-byte record line) once there are enough groups and plies to balance
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int64_t ngroups = (N + 31) / 32;
    long long chunk_knob = 0;   // knob rollout_chunk (experiments): 0 = one chunk per group (static schedule)
    int64_t chunk;
    if (b200rl_tuning(B200RL_TUNE_ROLLOUT_CHUNK, &chunk_knob)) {
        chunk = chunk_knob;
    } else {
        // a ticket should move ~256 KB (Geister with full moments: 4 plies x 32 games x 1972 B) so that the ticket
        // overhead (atomic, 64 B/game state round trip, fence) stays ~1 %; always a multiple of the 4-ply record line
        int64_t dims[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (b200rl_env_dims(game, dims) != B200RL_OK) return B200RL_ERR_ARG;
        const int64_t A = dims[0], obits = dims[1], sbits = dims[2];
        const int64_t per_step = (out_main ? 4 * obits : 0) + (out_scalar ? 4 * sbits : 0) + (out_amask ? 4 * A : 0) +
                                 (out_action ? 4 : 0) + (traj ? 32 : 0) + 8;
        chunk = ((262144 / (32 * per_step)) + 3) / 4 * 4;
        if (chunk < 4) chunk = 4;
    }
    if (chunk <= 0 || n_plies < 2 * chunk || ngroups < 2 * sms) chunk = n_plies;
    const int64_t nchunks = (n_plies + chunk - 1) / chunk;
    B200RL_REQUIRE(ngroups * nchunks < (int64_t)0xFFFF0000u, "env_rollout: too many tickets (N * n_plies)");
    unsigned int* sched = (unsigned int*)scratch;
    cudaMemsetAsync(sched, 0, (size_t)(SCHED_HDR + ngroups) * sizeof(unsigned int), (cudaStream_t)stream);
    p.sched = sched;
    p.chunk = chunk;
    const int64_t want = (ngroups + WARPS - 1) / WARPS;
    DISPATCH_GAME(game, launch_rollout<G>(p, want, dev, sms, (cudaStream_t)stream));
    return b200rl_check_launch("k_rollout");
}

// ---------------------------------------------------------------------------------------------------------------
// b200rl_env_rollout_static: b200rl_env_rollout for callers that cannot allocate (CUDA graph capture, tight serving
// loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The ticket scratch of k_rollout lives in a file-scope __device__ array of a stated maximum size instead
// of a caller buffer: every GPU that loads the library holds these B200RL_ROLLOUT_STATIC_MAX_GAMES / 32 + 32 words
// (32,896 bytes) until the process exits.
//
// In HOST code `g_rollout_static_sched` is only a stand-in for the symbol, not a device address: its device address is
// asked for once per device, with that device current (cudaGetSymbolAddress), and that pointer is what the kernel gets.
// All calls on a device share the array.  It keeps the previous call's contents; the launcher clears the words a
// launch uses on the caller's stream (cudaMemsetAsync) before the kernel, so calls on ONE stream are safe back to
// back, while calls that overlap on the GPU (different streams, concurrent graphs) race on the ticket counters.
// ---------------------------------------------------------------------------------------------------------------
static_assert(B200RL_ROLLOUT_STATIC_MAX_GAMES % 32 == 0, "whole groups of 32 games");
static __device__ __align__(256) unsigned int g_rollout_static_sched[SCHED_HDR + B200RL_ROLLOUT_STATIC_MAX_GAMES / 32];

extern "C" int b200rl_env_rollout_static(int game, void* state, int64_t N, int64_t n_plies, uint64_t seed,
                                         uint64_t game_id_base, uint64_t it0, float* out_main, float* out_scalar,
                                         float* out_amask, int32_t* out_action, int64_t waves, void* traj, int64_t D,
                                         int32_t* ep_table, int64_t E, uint64_t* counters, void* stream) {
    B200RL_REQUIRE(N <= (int64_t)B200RL_ROLLOUT_STATIC_MAX_GAMES,
                   "env_rollout_static: N = %lld games exceeds the static scratch (B200RL_ROLLOUT_STATIC_MAX_GAMES = %d); "
                   "use b200rl_env_rollout with a caller-owned scratch buffer",
                   (long long)N, (int)B200RL_ROLLOUT_STATIC_MAX_GAMES);
    static std::atomic<unsigned int*> sched_of_dev[64];   // device address of the array, per device
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    B200RL_REQUIRE(e == cudaSuccess && dev >= 0 && dev < 64, "env_rollout_static: cudaGetDevice: %s (device %d)",
                   cudaGetErrorString(e), dev);
    unsigned int* sched = sched_of_dev[dev].load(std::memory_order_acquire);
    if (sched == nullptr) {   // set-up, once per device, with that device current (a race here is benign: same answer)
        e = cudaGetSymbolAddress((void**)&sched, g_rollout_static_sched);
        B200RL_REQUIRE(e == cudaSuccess && sched != nullptr, "env_rollout_static: cudaGetSymbolAddress: %s",
                       cudaGetErrorString(e));
        sched_of_dev[dev].store(sched, std::memory_order_release);
    }
    return b200rl_env_rollout(game, state, N, n_plies, seed, game_id_base, it0, out_main, out_scalar, out_amask,
                              out_action, waves, traj, D, ep_table, E, counters, sched,
                              (int64_t)sizeof(unsigned int) * (SCHED_HDR + B200RL_ROLLOUT_STATIC_MAX_GAMES / 32), stream);
}
