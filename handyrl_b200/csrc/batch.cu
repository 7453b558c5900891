// Episode-to-batch assembly (handyrl/train.py:33-133 make_batch) from the device
// trajectory ring, plus the recording step that fills the ring.
//
// The ring holds one 32-byte record per (ply, game slot): the compact pre-move
// state, the action(s) and episode flags.  make_batch re-encodes observation,
// policy and action_mask from the record instead of copying 2.8 kB of fp32 per
// (b, t): HBM reads are 32 B per cell, writes are the batch tensors themselves.
#include <atomic>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>
#include <vector>

#include "common.cuh"
#include "games.cuh"
#include "emit.cuh"
#include "../../include/b200rl.h"

namespace {

constexpr int BLK = 128;
constexpr int WARPS = BLK / 32;

// Net outputs recorded beside the 32-byte record of a ply (generation.py:40-57: moment['value'], moment['policy']).
// Both side rings are slot-major like the record ring, so the plies of a training window are contiguous:
//   logit ring [N][D][PL][LS]  the logits of the LEGAL actions only, in ascending action-id order (everything else
//                              of moment['policy'] is logit - 1e32 = -1e32 in fp32): LS = LMAX rounded up to 4 floats,
//                              288 B per Geister ply instead of 856 B
//   value ring [N][D][2]       moment['value'][player] (0 where the player did not observe)
template <class G>
struct LogitSlots { static constexpr int value = (G::LMAX + 3) & ~3; };

template <class G>
__device__ __forceinline__ void record_net_outputs(const typename G::State& pre, int64_t i, int64_t D, uint64_t it,
                                                   const float* __restrict__ logits, const float* __restrict__ values,
                                                   float* logit_ring, float* value_ring) {
    constexpr int PL = G::SIMULTANEOUS ? 2 : 1;
    constexpr int LS = LogitSlots<G>::value;
    const int64_t cell = i * D + (int64_t)ring_mod(it, (uint64_t)D);
    if (logit_ring && logits) {
        uint32_t lm[G::LW];
        G::legal_bits(pre, lm);
#pragma unroll
        for (int q = 0; q < PL; ++q) {
            const float* src = logits + (i * PL + q) * G::A;
            float* dst = logit_ring + (cell * PL + q) * LS;
            int r = 0;
#pragma unroll
            for (int w = 0; w < G::LW; ++w)
                for (uint32_t m = lm[w]; m; m &= m - 1) dst[r++] = src[32 * w + __ffs((int)m) - 1];
        }
    }
    if (value_ring && values) {
        *reinterpret_cast<float2*>(value_ring + cell * 2) = make_float2(values[2 * i], values[2 * i + 1]);
    }
}

// ---------------------------------------------------------------------------
// step + record: one ply driven by caller-chosen actions (net-in-the-loop self-play
// or tape replay), writing the same trajectory records as k_rollout.
// ---------------------------------------------------------------------------
template <class G>
__global__ void __launch_bounds__(BLK)
k_step_record(uint32_t* state, int64_t N, const int32_t* __restrict__ actions, const int32_t* __restrict__ sel,
              const int32_t* __restrict__ both, int auto_reset, uint4* traj, int64_t D, uint64_t it,
              int32_t* ep_table, int64_t E, unsigned long long* counters, uint8_t* __restrict__ out_terminal,
              const float* __restrict__ logits, const float* __restrict__ values, float* logit_ring,
              float* value_ring) {
    const int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x;
    if (i >= N) return;
    typename G::State s = G::load(state, N, i);
    if (auto_reset && G::terminal(s)) G::reset(s);
    const int a = actions[i];
    if (a < 0) {  // no ply for this slot
        if (out_terminal) out_terminal[i] = G::terminal(s) ? 1 : 0;
        G::store(state, N, i, s);
        return;
    }
    const bool first = G::steps(s) == 0;
    const int selp = sel ? sel[i] : 0;
    uint32_t rec[8];
    G::pack_record(s, rec);
    record_net_outputs<G>(s, i, D, it, logits, values, logit_ring, value_ring);
    G::step(s, a, selp);
    const bool term = G::terminal(s);
    float o0 = 0.f, o1 = 0.f;
    if (term) G::outcome(s, o0, o1);
    const uint32_t wincode = term ? (o0 > 0 ? 1u : (o1 > 0 ? 2u : 3u)) : 0u;
    rec[5] = ((uint32_t)a & 0xFFFF) | ((uint32_t)(selp & 1) << 16) | ((first ? 1u : 0u) << 24) |
             ((term ? 1u : 0u) << 25) | (wincode << 26);
    rec[6] = (uint32_t)G::steps(s);
    rec[7] = both ? (((uint32_t)both[2 * i] & 0xFF) | (((uint32_t)both[2 * i + 1] & 0xFF) << 8)) : 0u;
    st_record(traj, D, N, it, i, rec);
    if (term && counters) {
        atomicAdd(&counters[0], 1ull);
        atomicAdd(&counters[1 + wincode], 1ull);
        if (ep_table) {
            unsigned long long idx = atomicAdd(&counters[5], 1ull);
            int32_t* e = ep_table + (int64_t)ring_mod(idx, (uint64_t)E) * 4;
            e[0] = (int32_t)i;
            e[1] = (int32_t)((uint32_t)G::steps(s) | (wincode << 16));
            e[2] = (int32_t)(uint32_t)it;
            e[3] = (int32_t)(uint32_t)(it >> 32);
        }
    }
    if (out_terminal) out_terminal[i] = term ? 1 : 0;
    G::store(state, N, i, s);
}

// ---------------------------------------------------------------------------
// sample + step + record: the acting half of one self-play ply with a real policy
// (handyrl/generation.py:48-61).  For every game: legal_actions in the reference's list
// order, p = logits[legal] (the 1e32 mask only removes the others), weights = softmax(p)
// as util.py:61-63 (fp32 exp(p - max) / sum), then random.choices' rule: cumulative weights
// accumulated as fp32 (they are numpy float32 scalars), first index whose cumulative weight exceeds u * total.  The chosen
// action is applied and recorded exactly like k_step_record.
// ---------------------------------------------------------------------------
template <class G>
__device__ __forceinline__ int sample_from_logits(const typename G::State& s, const float* __restrict__ lg, double u) {
    int ids[G::LMAX];
    const int n = G::legal_list(s, ids);
    if (n <= 0) return -1;
    float m = -INFINITY;
    for (int k = 0; k < n; ++k) m = fmaxf(m, lg[ids[k]]);
    // x.sum() of util.py:63 is numpy's pairwise sum: sequential below 8 terms, else 8 interleaved partial sums
    // combined as a tree and the n % 8 tail added one by one (n <= LMAX < 128: no recursion)
    auto ex = [&](int k) { return expf(__fsub_rn(lg[ids[k]], m)); };
    float sum = 0.f;
    if (n < 8) {
        for (int k = 0; k < n; ++k) sum = __fadd_rn(sum, ex(k));
    } else {
        float r[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) r[j] = ex(j);
        int i = 8;
        for (; i < n - (n % 8); i += 8) {
#pragma unroll
            for (int j = 0; j < 8; ++j) r[j] = __fadd_rn(r[j], ex(i + j));
        }
        sum = __fadd_rn(__fadd_rn(__fadd_rn(r[0], r[1]), __fadd_rn(r[2], r[3])),
                        __fadd_rn(__fadd_rn(r[4], r[5]), __fadd_rn(r[6], r[7])));
        for (; i < n; ++i) sum = __fadd_rn(sum, ex(i));
    }
    // random.choices: cum_weights = list(accumulate(weights)) over numpy float32 scalars = an fp32 running sum;
    // total = cum_weights[-1]; population[bisect(cum_weights, random() * total, 0, n - 1)]
    float total = 0.f;
    for (int k = 0; k < n; ++k) total = __fadd_rn(total, __fdiv_rn(ex(k), sum));
    const double x = u * (double)total;
    float cum = 0.f;
    int pick = n - 1;
    for (int k = 0; k < n - 1; ++k) {
        cum = __fadd_rn(cum, __fdiv_rn(ex(k), sum));
        if (x < (double)cum) { pick = k; break; }   // bisect_right(cum_weights, x, 0, n - 1)
    }
    return ids[pick];
}

template <class G>
__global__ void __launch_bounds__(BLK)
k_sample_record(uint32_t* state, int64_t N, const float* __restrict__ logits, const double* __restrict__ u_in,
                uint64_t seed, uint64_t game_id_base, uint4* traj, int64_t D, uint64_t it, int32_t* ep_table,
                int64_t E, unsigned long long* counters, int32_t* __restrict__ out_action,
                uint8_t* __restrict__ out_terminal, const float* __restrict__ values, float* logit_ring,
                float* value_ring) {
    const int64_t i = blockIdx.x * (int64_t)BLK + threadIdx.x;
    if (i >= N) return;
    typename G::State s = G::load(state, N, i);
    if (G::terminal(s)) G::reset(s);  // generation.py:27: a finished slot starts its next episode
    constexpr int PL = G::SIMULTANEOUS ? 2 : 1;  // players acting per ply
    const uint32_t r = b200rl_rand32(seed, game_id_base + (uint64_t)i, it);
    int acts[2] = {-1, -1};
#pragma unroll
    for (int q = 0; q < PL; ++q) {
        double u;
        if (u_in) u = u_in[i * PL + q];
        else {  // 53 random bits like random.random()
            const uint64_t hi = b200rl_rand32(seed ^ 0xA5A5A5A5ull, game_id_base + (uint64_t)i, it * 2 + q);
            const uint64_t lo = b200rl_rand32(seed ^ 0x5A5A5A5Aull, game_id_base + (uint64_t)i, it * 2 + q);
            u = (double)(((hi << 32) | lo) >> 11) * (1.0 / 9007199254740992.0);
        }
        acts[q] = sample_from_logits<G>(s, logits + (i * PL + q) * G::A, u);
    }
    const int selp = G::SIMULTANEOUS ? (int)((r >> 7) & 1) : 0;   // parallel_tictactoe.py:22
    const int a = G::SIMULTANEOUS ? acts[selp] : acts[0];
    const bool first = G::steps(s) == 0;
    uint32_t rec[8];
    G::pack_record(s, rec);
    if (traj) record_net_outputs<G>(s, i, D, it, logits, values, logit_ring, value_ring);
    G::step(s, a, selp);
    const bool term = G::terminal(s);
    float o0 = 0.f, o1 = 0.f;
    if (term) G::outcome(s, o0, o1);
    const uint32_t wincode = term ? (o0 > 0 ? 1u : (o1 > 0 ? 2u : 3u)) : 0u;
    rec[5] = ((uint32_t)a & 0xFFFF) | ((uint32_t)(selp & 1) << 16) | ((first ? 1u : 0u) << 24) |
             ((term ? 1u : 0u) << 25) | (wincode << 26);
    rec[6] = (uint32_t)G::steps(s);
    rec[7] = G::SIMULTANEOUS ? (((uint32_t)acts[0] & 0xFF) | (((uint32_t)acts[1] & 0xFF) << 8)) : 0u;
    if (traj) st_record(traj, D, N, it, i, rec);
    if (counters) {
        atomicAdd(&counters[1], 1ull);
        if (term) {
            atomicAdd(&counters[0], 1ull);
            atomicAdd(&counters[1 + wincode], 1ull);
            if (ep_table) {
                unsigned long long idx = atomicAdd(&counters[5], 1ull);
                int32_t* e = ep_table + (int64_t)ring_mod(idx, (uint64_t)E) * 4;
                e[0] = (int32_t)i;
                e[1] = (int32_t)((uint32_t)G::steps(s) | (wincode << 16));
                e[2] = (int32_t)(uint32_t)it;
                e[3] = (int32_t)(uint32_t)(it >> 32);
            }
        }
    }
    if (out_action) {
#pragma unroll
        for (int q = 0; q < PL; ++q) out_action[i * PL + q] = acts[q];
    }
    if (out_terminal) out_terminal[i] = term ? 1 : 0;
    G::store(state, N, i, s);
}

// ---------------------------------------------------------------------------
// make_batch
// ---------------------------------------------------------------------------
enum { MB_TURN = 0, MB_ALL = 1, MB_SOLO = 2 };

struct BatchParams {
    const uint4* traj;
    int64_t D, N;
    const float* values;  // [N][D][2] value estimates recorded with the ring, or null (RandomModel: 0)
    const float* logits;  // [N][D][PL][LS] logits of the legal actions, or null (RandomModel: zero logits)
    const int32_t* win;   // [B][8] = slot, it_lo, it_hi, len, st, total, wincode, solo player
    const float* ret_table;  // discounted return of a ply with m plies left, m = 0..ret_len-1
    int ret_len;
    int64_t B;
    int T, mode, obs_all;
    float *obs_main, *obs_scalar, *policy, *amask;
    long long* action;
    float *value, *reward, *ret, *outcome, *emask, *tmask, *omask, *progress;
};

// per-warp staging for 32 cells x up to 2 observed players: flat bit strings (emit.cuh)
template <class G>
struct BatchStage {
    uint32_t obs[flat_words(64, G::OBITS)];
    uint32_t legal[flat_words(64, G::A)];     // 1 = legal                (action_mask: 0 / 1e32)
    uint32_t illegal[flat_words(64, G::A)];   // 1 = illegal on a live row (policy: -1e32 / 0)
    uint32_t scalar[flat_words(64, G::SBITS > 0 ? G::SBITS : 1)];
};

template <class G>
struct BatchCta {
    float4 tbl_obs[16], tbl_amask[16], tbl_policy[16];
    BatchStage<G> w[WARPS];
};

// Po (1 or 2) consecutive rows per lane: the tail of a lane's last row goes to the next lane by shuffle, the
// tail of its first row (Po == 2) is carried in registers.  bits[q] of absent rows must be zero.
template <int NB, int RW>
__device__ __forceinline__ void pack_rows(uint32_t* flat, int lane, int Po, int n_rows, bool valid,
                                          const uint32_t (&bits)[2][RW]) {
    const uint32_t t0 = flat_row_tail<NB, RW>(lane * Po, bits[0]);
    const uint32_t t1 = Po == 2 ? flat_row_tail<NB, RW>(lane * Po + 1, bits[1]) : 0u;
    const uint32_t carry = __shfl_up_sync(0xffffffffu, Po == 2 ? t1 : t0, 1);
    if (!valid) return;
    const int r0 = lane * Po;
    flat_row_store<NB, RW>(flat, r0, bits[0], lane > 0 ? carry : 0u, r0 == n_rows - 1, t0);
    if (Po == 2) flat_row_store<NB, RW>(flat, r0 + 1, bits[1], t0, r0 + 1 == n_rows - 1, t1);
}

// TMA = the big tensors leave through staged bulk copies (emit_flat_f32_tma) instead of STG.128 (emit_flat_f32);
// dynamic shared memory then holds WARPS x MB_NBUF x MB_CH bytes of staging.
constexpr int MB_CH = 4096, MB_NBUF = 3;

template <class G, bool TMA>
__global__ void __launch_bounds__(BLK) k_make_batch(const BatchParams p) {
    __shared__ BatchCta<G> cta;
    extern __shared__ __align__(128) unsigned char mb_dyn_smem[];
    nibble_table_init(cta.tbl_obs, 0.f, 1.f, threadIdx.x);
    nibble_table_init(cta.tbl_amask, 1e32f, 0.f, threadIdx.x);
    nibble_table_init(cta.tbl_policy, 0.f, -1e32f, threadIdx.x);
    __syncthreads();
    BatchStage<G>& st = cta.w[threadIdx.x >> 5];
    const int lane = threadIdx.x & 31;
    const int64_t ncell = p.B * p.T;
    const int64_t c0 = (blockIdx.x * (int64_t)WARPS + (threadIdx.x >> 5)) * 32;
    if (c0 >= ncell) return;
    const int nc = (int)min((int64_t)32, ncell - c0);
    const bool valid = lane < nc;
    const int64_t c = c0 + lane;
    const int Po = p.mode == MB_ALL ? 2 : 1;   // players carrying observation / policy
    const int Pv = p.mode == MB_SOLO ? 1 : 2;  // players carrying value / reward / return

    // ---- per-cell integer work ----
    uint32_t ob[2][G::OW], sc[2] = {0, 0}, lg[2][G::LW], il[2][G::LW];
#pragma unroll
    for (int q = 0; q < 2; ++q) {
#pragma unroll
        for (int k = 0; k < G::OW; ++k) ob[q][k] = 0;
#pragma unroll
        for (int k = 0; k < G::LW; ++k) lg[q][k] = il[q][k] = 0;
    }
    if (valid) {
        const int64_t b = c / p.T;
        const int t = (int)(c - b * p.T);
        const int32_t* w = p.win + b * 8;
        const int len = w[3], stt = w[4], total = w[5], wincode = w[6], solo = w[7];
        const float oc0 = wincode == 1 ? 1.f : (wincode == 2 ? -1.f : 0.f);
        const float oc1 = wincode == 1 ? -1.f : (wincode == 2 ? 1.f : 0.f);
        const bool live = t < len;
        long long act[2] = {0, 0};
        float val[2] = {oc0, oc1}, rew[2] = {0.f, 0.f}, rtn[2] = {0.f, 0.f}, tm[2] = {0.f, 0.f}, om[2] = {0.f, 0.f};
        if (live) {
            const uint64_t it = ((uint64_t)(uint32_t)w[1] | ((uint64_t)(uint32_t)w[2] << 32)) + (uint64_t)t;
            const int64_t slot = w[0];
            uint32_t rec[8];
            ld_record(p.traj, p.D, p.N, it, slot, rec);
            typename G::State s = G::unpack_record(rec);
            const int tp = G::turn(s);  // m['turn'][0]
            const bool parallel = G::SIMULTANEOUS;
            const int a_turn = (int)(rec[5] & 0xFFFF);
            int pa[2] = {a_turn, a_turn};
            if (parallel) { pa[0] = (int)(rec[7] & 0xFF); pa[1] = (int)((rec[7] >> 8) & 0xFF); }
            uint32_t lm[G::LW];
            G::legal_bits(s, lm);
            for (int q = 0; q < Po; ++q) {
                // which player does output slot q describe?
                const int pl = p.mode == MB_TURN ? tp : (p.mode == MB_SOLO ? solo : q);
                const bool is_turn = parallel || pl == tp;
                // observation exists for turn players, and for everyone when args['observation'];
                // otherwise it is the zero template (train.py:60,70)
                const bool has_obs = is_turn || p.obs_all;
                if (has_obs) G::obs_bits(s, pl, ob[q], sc[q]);
                if (is_turn) {
#pragma unroll
                    for (int k = 0; k < G::LW; ++k) {
                        lg[q][k] = lm[k];
                        il[q][k] = ~lm[k];
                    }
                    il[q][G::LW - 1] &= (G::A % 32) ? ((1u << (G::A % 32)) - 1u) : 0xFFFFFFFFu;
                    act[q] = parallel ? pa[pl] : a_turn;
                }
            }
            for (int q = 0; q < Pv; ++q) {
                const int pl = p.mode == MB_SOLO ? solo : q;
                const bool is_turn = parallel || pl == tp;
                const bool observed = is_turn || p.obs_all;
                float v = 0.f;
                if (observed && p.values) v = p.values[(slot * p.D + (int64_t)ring_mod(it, (uint64_t)p.D)) * 2 + pl];
                val[q] = v;
                rew[q] = G::HAS_REWARD ? G::reward(s) : 0.f;
                const int left = total - (stt + t);
                rtn[q] = (G::HAS_REWARD && left >= 0 && left < p.ret_len) ? p.ret_table[left] : 0.f;
                tm[q] = is_turn ? 1.f : 0.f;
                om[q] = observed ? 1.f : 0.f;
            }
        } else if (p.mode == MB_SOLO) {
            val[0] = solo == 0 ? oc0 : oc1;
        }
        // ---- small per-cell tensors ----
        for (int q = 0; q < Po; ++q) p.action[c * Po + q] = act[q];
        if (Pv == 2) {   // both players of a cell in one 8-byte store: a warp writes 256 contiguous bytes per tensor
            *reinterpret_cast<float2*>(p.value + c * 2) = make_float2(val[0], val[1]);
            *reinterpret_cast<float2*>(p.reward + c * 2) = make_float2(rew[0], rew[1]);
            *reinterpret_cast<float2*>(p.ret + c * 2) = make_float2(rtn[0], rtn[1]);
            *reinterpret_cast<float2*>(p.tmask + c * 2) = make_float2(tm[0], tm[1]);
            *reinterpret_cast<float2*>(p.omask + c * 2) = make_float2(om[0], om[1]);
        } else {
            p.value[c] = val[0];
            p.reward[c] = rew[0];
            p.ret[c] = rtn[0];
            p.tmask[c] = tm[0];
            p.omask[c] = om[0];
        }
        p.emask[c] = live ? 1.f : 0.f;
        p.progress[c] = live ? __fdiv_rn((float)(stt + t), (float)total) : 1.f;
        if (t == 0) {
            if (p.mode == MB_SOLO) p.outcome[b] = solo == 0 ? oc0 : oc1;
            else { p.outcome[b * 2] = oc0; p.outcome[b * 2 + 1] = oc1; }
        }
    }
    // ---- pack bits into the warp's flat strings, then stream the big tensors ----
    constexpr int SB = G::SBITS > 0 ? G::SBITS : 1;
    const int ng = nc * Po;
    const bool rnd_policy = p.logits == nullptr;
    // rows of >= 32 bits (Geister): shuffle-merged packing, every lane takes part (invalid lanes carry zero bits);
    // shorter rows (TicTacToe, the 18 scalar bits): atomicOr packing after a clearing pass
    constexpr bool BIG_O = G::OBITS >= 32, BIG_A = G::A >= 32;
    if constexpr (G::OBITS < 32 && G::A < 16 && G::SBITS == 0 && !TMA) {
        // TicTacToe.  The Po rows of a lane are neighbours in the flat string, so a lane owns ONE row of Po x NB bits:
        // mask and policy tiles (9 or 18 bits per lane) and the one-player observation tile (27) stay in one register
        // per lane and are packed and streamed with shuffles (emit_short_rows_f32); two players' 54 observation bits
        // take the shuffle-merged path of rows >= 32 bits.  No clearing pass, no atomicOr.
        if (Po == 1) {
            emit_short_rows_f32<G::OBITS>(ob[0][0], cta.tbl_obs, p.obs_main + c0 * G::OBITS, nc, lane, 0.f, 1.f);
            emit_short_rows_f32<G::A>(lg[0][0], cta.tbl_amask, p.amask + c0 * G::A, nc, lane, 1e32f, 0.f);
            if (rnd_policy)
                emit_short_rows_f32<G::A>(il[0][0], cta.tbl_policy, p.policy + c0 * G::A, nc, lane, 0.f, -1e32f);
        } else {
            const uint32_t o2[2] = {ob[0][0] | (ob[1][0] << G::OBITS), ob[1][0] >> (32 - G::OBITS)};
            flat_pack_warp<2 * G::OBITS, 2>(st.obs, lane, nc, o2);
            __syncwarp();
            emit_flat_f32<2 * G::OBITS>(st.obs, cta.tbl_obs, p.obs_main + c0 * 2 * G::OBITS, nc, lane, 0.f, 1.f);
            emit_short_rows_f32<2 * G::A>(lg[0][0] | (lg[1][0] << G::A), cta.tbl_amask, p.amask + c0 * 2 * G::A, nc,
                                          lane, 1e32f, 0.f);
            if (rnd_policy)
                emit_short_rows_f32<2 * G::A>(il[0][0] | (il[1][0] << G::A), cta.tbl_policy, p.policy + c0 * 2 * G::A,
                                              nc, lane, 0.f, -1e32f);
        }
        return;
    }
    if (valid) {
        for (int q = 0; q < Po; ++q) {
            const int row = lane * Po + q;
            const bool last = row == ng - 1;
            if (!BIG_O) flat_clear_row<G::OBITS>(st.obs, row, last);
            if (G::SBITS > 0) flat_clear_row<SB>(st.scalar, row, last);
            if (!BIG_A) {
                flat_clear_row<G::A>(st.legal, row, last);
                if (rnd_policy) flat_clear_row<G::A>(st.illegal, row, last);
            }
        }
    }
    if (!BIG_O || !BIG_A || G::SBITS > 0) __syncwarp();
    if (BIG_O) pack_rows<(BIG_O ? G::OBITS : 32), G::OW>(st.obs, lane, Po, ng, valid, ob);
    if (BIG_A) {
        pack_rows<(BIG_A ? G::A : 32), G::LW>(st.legal, lane, Po, ng, valid, lg);
        if (rnd_policy) pack_rows<(BIG_A ? G::A : 32), G::LW>(st.illegal, lane, Po, ng, valid, il);
    }
    if (valid) {
        for (int q = 0; q < Po; ++q) {
            const int row = lane * Po + q;
            if (!BIG_O) flat_put<G::OBITS, G::OW>(st.obs, row, ob[q]);
            if (G::SBITS > 0) {
                const uint32_t s1[1] = {sc[q]};
                flat_put<SB, 1>(st.scalar, row, s1);
            }
            if (!BIG_A) {
                flat_put<G::A, G::LW>(st.legal, row, lg[q]);
                if (rnd_policy) flat_put<G::A, G::LW>(st.illegal, row, il[q]);
            }
        }
    }
    __syncwarp();
    if (TMA) {
        TmaStage<MB_CH, MB_NBUF> ts;
        ts.buf = reinterpret_cast<float4*>(mb_dyn_smem) + (size_t)(threadIdx.x >> 5) * MB_NBUF * (MB_CH / 16);
        ts.c = 0;
        emit_flat_f32_tma<G::OBITS>(st.obs, p.obs_main + c0 * Po * G::OBITS, ng, lane, 0.f, 1.f, ts);
        if (G::SBITS > 0) emit_flat_f32_tma<SB>(st.scalar, p.obs_scalar + c0 * Po * G::SBITS, ng, lane, 0.f, 1.f, ts);
        emit_flat_f32_tma<G::A>(st.legal, p.amask + c0 * Po * G::A, ng, lane, 1e32f, 0.f, ts);
        if (rnd_policy) emit_flat_f32_tma<G::A>(st.illegal, p.policy + c0 * Po * G::A, ng, lane, 0.f, -1e32f, ts);
        ts.drain(lane);
        return;
    }
    emit_flat_f32<G::OBITS>(st.obs, cta.tbl_obs, p.obs_main + c0 * Po * G::OBITS, ng, lane, 0.f, 1.f);
    if (G::SBITS > 0) emit_flat_f32<SB>(st.scalar, cta.tbl_obs, p.obs_scalar + c0 * Po * G::SBITS, ng, lane, 0.f, 1.f);
    emit_flat_f32<G::A>(st.legal, cta.tbl_amask, p.amask + c0 * Po * G::A, ng, lane, 1e32f, 0.f);
    if (rnd_policy) {
        // RandomModel: policy = 0 - action_mask on live turn rows, 0 elsewhere (train.py:61,70,94)
        emit_flat_f32<G::A>(st.illegal, cta.tbl_policy, p.policy + c0 * Po * G::A, ng, lane, 0.f, -1e32f);
    }
}

// policy = logits - action_mask for recorded behaviour logits (generation.py:52): one warp per
// (cell, player) row; rows without any legal action (padding, non-turn players) stay zero.  The ring holds the
// legal logits in ascending action-id order: entry a of a legal action is its rank among the legal bits, found
// with a ballot per 32 actions; an illegal entry is logit - 1e32 = -1e32 for any logit below 4.8e24.
template <class G>
__global__ void __launch_bounds__(BLK) k_policy_from_logits(const BatchParams p) {
    constexpr int PL = G::SIMULTANEOUS ? 2 : 1;
    constexpr int LS = LogitSlots<G>::value;
    const int Po = p.mode == MB_ALL ? 2 : 1;
    const int lane = threadIdx.x & 31;
    const int64_t rows = p.B * p.T * Po;
    const int64_t r = blockIdx.x * (int64_t)WARPS + (threadIdx.x >> 5);
    if (r >= rows) return;
    const int64_t c = r / Po;
    const int q = (int)(r - c * Po);
    const int64_t b = c / p.T;
    const int t = (int)(c - b * p.T);
    const int32_t* w = p.win + b * 8;
    // the player this row describes; only simultaneous games record one logit row per player
    const int pl = PL == 1 ? 0 : (p.mode == MB_TURN ? 0 : (p.mode == MB_SOLO ? (w[7] & 1) : q));
    const uint64_t it = ((uint64_t)(uint32_t)w[1] | ((uint64_t)(uint32_t)w[2] << 32)) + (uint64_t)t;
    const float* lg = p.logits + (((int64_t)w[0] * p.D + (int64_t)ring_mod(it, (uint64_t)p.D)) * PL + pl) * LS;
    const bool in_window = t < w[3];
    int base = 0;
    float out[(G::A + 31) / 32];
    bool any = false;
#pragma unroll
    for (int j = 0; j < (G::A + 31) / 32; ++j) {
        const int a = j * 32 + lane;
        const bool legal = a < G::A && p.amask[r * G::A + a] == 0.f;
        const uint32_t bits = __ballot_sync(0xffffffffu, legal);
        const int rank = base + __popc(bits & ((1u << lane) - 1u));
        out[j] = (legal && in_window) ? lg[rank] : -1e32f;
        base += __popc(bits);
        any |= bits != 0u;
    }
    const bool live = in_window && any;
#pragma unroll
    for (int j = 0; j < (G::A + 31) / 32; ++j) {
        const int a = j * 32 + lane;
        if (a < G::A) p.policy[r * G::A + a] = live ? out[j] : 0.f;
    }
}

// Ragged -> padded stack for host-decoded reference episodes: out[b][t][:] = src[off[b]+t][:]
// for t < off[b+1]-off[b], else fill_rows[b][:] (value <- outcome, train.py:96) or the constant.
__global__ void __launch_bounds__(256)
k_pad_stack(const uint32_t* __restrict__ src, const int64_t* __restrict__ off, const uint32_t* __restrict__ fill_rows,
            uint32_t fill, uint32_t* __restrict__ out, int64_t B, int64_t T, int64_t row) {
    const int64_t n = B * T * row;
    for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < n; e += (int64_t)gridDim.x * blockDim.x) {
        const int64_t j = e % row;
        const int64_t bt = e / row;
        const int64_t t = bt % T, b = bt / T;
        const int64_t len = off[b + 1] - off[b];
        uint32_t v;
        if (t < len) v = src[(off[b] + t) * row + j];
        else v = fill_rows ? fill_rows[b * row + j] : fill;
        out[e] = v;
    }
}

// Batcher.select_episode (handyrl/train.py:284-302) for B windows at once, over the finished-episode
// table the rollout kernels fill.  The replay buffer `self.episodes` is the last
// L = min(finished, E, maximum_episodes) rows of the table in finishing order (index 0 = oldest);
// a row whose first ply has already been overwritten in the trajectory ring is skipped like a rejected
// draw.  Draws come from the counter RNG keyed (seed, window b, draw0 + 4*try + j).
// One window; shared by the device kernel and the host entry point so both give identical rows.
__host__ __device__ inline void select_one_window(const int32_t* ep_table, int64_t E, unsigned long long cursor,
                                                  int64_t maximum_episodes, int forward_steps, uint64_t seed,
                                                  uint64_t draw0, uint64_t it_now, int64_t D, int64_t b, int32_t* w) {
    // the reference trims self.episodes to maximum_episodes (train.py:502-503)
    int64_t L = (int64_t)(cursor < (unsigned long long)E ? cursor : (unsigned long long)E);
    if (L > maximum_episodes) L = maximum_episodes;
    if (L <= 0) {  // nothing finished yet: an empty window (all rows padded)
        w[0] = 0; w[1] = 0; w[2] = 0; w[3] = 0; w[4] = 0; w[5] = 1; w[6] = 3; w[7] = 0;
        return;
    }
    const int64_t n_draw = L < maximum_episodes ? L : maximum_episodes;
    int32_t slot = 0, steps = 1, win = 3;
    uint64_t end_it = 0, c = draw0;
    const bool e_pow2 = (E & (E - 1)) == 0;
    for (int tr = 0; tr < 64; ++tr) {
        c = draw0 + 2ull * (uint64_t)tr;
        const uint64_t x = b200rl_rand64(seed, (uint64_t)b, c);   // one hash per try: index draw | acceptance draw
        const uint32_t r0 = (uint32_t)(x >> 32), r1 = (uint32_t)x;
        int64_t ep_idx = (int64_t)(((unsigned long long)r0 * (unsigned long long)n_draw) >> 32);  // train.py:286
        const double accept = 1.0 - (double)(L - 1 - ep_idx) / (double)maximum_episodes;         // train.py:287
        const bool last_try = tr == 63;
        if (last_try) ep_idx = L - 1;
        // train.py:288 first: a try rejected by the recency law never needs its table row (one random 16-byte
        // read less per rejection; the outcome is the same as testing `alive && accepted`)
        if (!last_try && !((double)r1 * (1.0 / 4294967296.0) < accept)) continue;
        const unsigned long long pos = cursor - (unsigned long long)L + (unsigned long long)ep_idx;
        const int32_t* e = ep_table + (int64_t)(e_pow2 ? (pos & (unsigned long long)(E - 1)) : (pos % (unsigned long long)E)) * 4;
        slot = e[0];
        steps = e[1] & 0xFFFF;
        win = (e[1] >> 16) & 3;
        end_it = (uint64_t)(uint32_t)e[2] | ((uint64_t)(uint32_t)e[3] << 32);
        const bool alive = it_now - (end_it + 1 - (uint64_t)steps) <= (uint64_t)D;
        if (last_try || alive) break;
    }
    // the window-start and solo-player draws of the accepted try (rejected tries never needed theirs)
    const uint64_t y = b200rl_rand64(seed, (uint64_t)b, c + 1);
    const uint32_t r2 = (uint32_t)(y >> 32), r3 = (uint32_t)y;
    const int cand = 1 + (steps - forward_steps > 0 ? steps - forward_steps : 0);         // train.py:291
    const int st = (int)(((unsigned long long)r2 * (unsigned long long)cand) >> 32);       // train.py:292
    const int ed = st + forward_steps < steps ? st + forward_steps : steps;               // train.py:293
    const uint64_t first = end_it + 1 - (uint64_t)steps + (uint64_t)st;
    w[0] = slot; w[1] = (int32_t)(uint32_t)first; w[2] = (int32_t)(uint32_t)(first >> 32);
    w[3] = ed - st; w[4] = st; w[5] = steps; w[6] = win; w[7] = (int32_t)(r3 >> 31);  // solo player, train.py:57-58
}

__global__ void __launch_bounds__(128)
k_select_windows(const int32_t* __restrict__ ep_table, int64_t E, const unsigned long long* __restrict__ counters,
                 int64_t maximum_episodes, int forward_steps, uint64_t seed, uint64_t draw0, uint64_t it_now,
                 int64_t D, int32_t* __restrict__ out, int64_t B) {
    const int64_t b = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
    if (b >= B) return;
    select_one_window(ep_table, E, counters[5], maximum_episodes, forward_steps, seed, draw0, it_now, D, b, out + b * 8);
}

inline int grid_for(int64_t n) { return (int)((n + BLK - 1) / BLK); }

template <class G>
int launch_make_batch(const BatchParams& p, cudaStream_t stream) {
    B200RL_REQUIRE(G::SBITS == 0 || p.obs_scalar, "make_batch: this game needs obs_scalar");
    const int64_t ncell = p.B * p.T;
    const int Po = p.mode == MB_ALL ? 2 : 1;
    // staged bulk stores for the games whose tiles are large (Geister: 32 KB of observation per warp tile); the
    // 3 KB TicTacToe tiles stay on STG.128.  Measured on the bench workload: 0.943 -> 0.939 ms (two A/B pairs,
    // profiles/r2_tma_emit_ab.md) - within 1 %, although the isolated emit probe gains 4-6 %: what holds make_batch
    // at 6.4 TB/s is the record gather inside the write stream, not the store path.  Knob emit_tma = 0 / 1 forces
    // either (A/B measurements, tests).
    long long knob = 0;
    const bool tma = b200rl_tuning(B200RL_TUNE_EMIT_TMA, &knob) ? knob != 0 : G::OBITS >= 128;
    if (tma) {
        constexpr int DYN = WARPS * MB_NBUF * MB_CH;
        static std::atomic<int> attr_set[64];
        int dev = 0;
        cudaGetDevice(&dev);
        if (attr_set[dev & 63].load(std::memory_order_relaxed) == 0) {
            cudaFuncSetAttribute(k_make_batch<G, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, DYN);
            attr_set[dev & 63].store(1, std::memory_order_relaxed);
        }
        k_make_batch<G, true><<<grid_for(ncell), BLK, DYN, stream>>>(p);
    } else {
        k_make_batch<G, false><<<grid_for(ncell), BLK, 0, stream>>>(p);
    }
    if (p.logits) {
        const int64_t rows = ncell * Po;
        k_policy_from_logits<G><<<(int)((rows + WARPS - 1) / WARPS), BLK, 0, stream>>>(p);
    }
    return b200rl_check_launch("k_make_batch");
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

extern "C" int64_t b200rl_env_logit_slots(int game) {
    DISPATCH_GAME(game, return LogitSlots<G>::value);
    return -1;
}

extern "C" int b200rl_env_step_record(int game, void* state, int64_t N, const int32_t* actions,
                                      const int32_t* selected, const int32_t* both, int auto_reset, void* traj,
                                      int64_t D, uint64_t it, int32_t* ep_table, int64_t E, uint64_t* counters,
                                      uint8_t* out_terminal, const float* logits, const float* values,
                                      float* logit_ring, float* value_ring, void* stream) {
    B200RL_REQUIRE(!value_ring || ((uintptr_t)value_ring % 8) == 0, "env_step_record: value ring must be 8-byte aligned");
    B200RL_REQUIRE(state && actions && traj && D >= 1, "env_step_record: null state/actions/traj or D < 1");
    B200RL_REQUIRE(((uintptr_t)state % 16) == 0 && ((uintptr_t)traj % 32) == 0, "env_step_record: 16-byte alignment");
    B200RL_REQUIRE(!ep_table || (E >= 1 && counters), "env_step_record: episode table needs E >= 1 and counters");
    if (N <= 0) return B200RL_OK;
    DISPATCH_GAME(game, (k_step_record<G><<<grid_for(N), BLK, 0, (cudaStream_t)stream>>>(
                            (uint32_t*)state, N, actions, selected, both, auto_reset, (uint4*)traj, D, it, ep_table,
                            E, (unsigned long long*)counters, out_terminal, logits, values, logit_ring, value_ring)));
    return b200rl_check_launch("k_step_record");
}

extern "C" int b200rl_env_sample_record(int game, void* state, int64_t N, const float* logits, const double* u,
                                        uint64_t seed, uint64_t game_id_base, void* traj, int64_t D, uint64_t it,
                                        int32_t* ep_table, int64_t E, uint64_t* counters, int32_t* out_action,
                                        uint8_t* out_terminal, const float* values, float* logit_ring,
                                        float* value_ring, void* stream) {
    B200RL_REQUIRE(!value_ring || ((uintptr_t)value_ring % 8) == 0, "env_sample_record: value ring must be 8-byte aligned");
    B200RL_REQUIRE(state && logits, "env_sample_record: null state/logits");
    B200RL_REQUIRE(((uintptr_t)state % 16) == 0 && ((uintptr_t)traj % 32) == 0, "env_sample_record: 16-byte alignment");
    B200RL_REQUIRE(!traj || D >= 1, "env_sample_record: trajectory ring depth D must be >= 1");
    B200RL_REQUIRE(!ep_table || (E >= 1 && counters), "env_sample_record: episode table needs E >= 1 and counters");
    if (N <= 0) return B200RL_OK;
    DISPATCH_GAME(game, (k_sample_record<G><<<grid_for(N), BLK, 0, (cudaStream_t)stream>>>(
                            (uint32_t*)state, N, logits, u, seed, game_id_base, (uint4*)traj, D, it, ep_table, E,
                            (unsigned long long*)counters, out_action, out_terminal, values, logit_ring, value_ring)));
    return b200rl_check_launch("k_sample_record");
}

extern "C" int b200rl_make_batch(int game, const void* traj, int64_t D, int64_t N, const float* values,
                                 const float* logits, const int32_t* windows, int64_t B, int64_t T, int mode,
                                 int obs_all, const float* ret_table, int64_t ret_len, float* obs_main,
                                 float* obs_scalar, float* policy, float* amask, int64_t* action, float* value,
                                 float* reward, float* ret, float* outcome, float* emask, float* tmask, float* omask,
                                 float* progress, void* stream) {
    B200RL_REQUIRE(traj && windows && D >= 1 && N >= 1, "make_batch: null trajectory ring / windows");
    B200RL_REQUIRE(mode >= 0 && mode <= 2, "make_batch: mode must be 0 (turn), 1 (all players) or 2 (solo)");
    B200RL_REQUIRE(T >= 1 && B >= 0, "make_batch: bad B/T");
    B200RL_REQUIRE(obs_main && policy && amask && action && value && reward && ret && outcome && emask && tmask &&
                       omask && progress, "make_batch: null output");
    B200RL_REQUIRE(ret_table && ret_len >= 1, "make_batch: null return table");
    B200RL_REQUIRE(((uintptr_t)obs_main % 16) == 0 && ((uintptr_t)policy % 16) == 0 && ((uintptr_t)amask % 16) == 0 &&
                       ((uintptr_t)obs_scalar % 16) == 0, "make_batch: outputs must be 16-byte aligned");
    B200RL_REQUIRE(((uintptr_t)value % 8) == 0 && ((uintptr_t)reward % 8) == 0 && ((uintptr_t)ret % 8) == 0 &&
                       ((uintptr_t)tmask % 8) == 0 && ((uintptr_t)omask % 8) == 0 && ((uintptr_t)windows % 16) == 0,
                   "make_batch: value/reward/return/masks must be 8-byte aligned, windows 16-byte aligned");
    if (B == 0) return B200RL_OK;
    BatchParams p;
    p.traj = (const uint4*)traj; p.D = D; p.N = N; p.values = values; p.logits = logits; p.win = windows;
    p.ret_table = ret_table; p.ret_len = (int)ret_len; p.B = B; p.T = (int)T; p.mode = mode; p.obs_all = obs_all;
    p.obs_main = obs_main; p.obs_scalar = obs_scalar; p.policy = policy; p.amask = amask;
    p.action = (long long*)action; p.value = value; p.reward = reward; p.ret = ret; p.outcome = outcome;
    p.emask = emask; p.tmask = tmask; p.omask = omask; p.progress = progress;
    DISPATCH_GAME(game, return launch_make_batch<G>(p, (cudaStream_t)stream));
    return B200RL_OK;
}

extern "C" int b200rl_select_windows(const int32_t* ep_table, int64_t E, const uint64_t* counters,
                                     int64_t maximum_episodes, int64_t forward_steps, uint64_t seed, uint64_t draw0,
                                     uint64_t it_now, int64_t D, int32_t* out_windows, int64_t B, void* stream) {
    B200RL_REQUIRE(ep_table && counters && out_windows, "select_windows: null pointer");
    B200RL_REQUIRE(E >= 1 && D >= 1 && maximum_episodes >= 1 && forward_steps >= 1 && B >= 0,
                   "select_windows: bad E/D/maximum_episodes/forward_steps/B");
    if (B == 0) return B200RL_OK;
    k_select_windows<<<(int)((B + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
        ep_table, E, (const unsigned long long*)counters, maximum_episodes, (int)forward_steps, seed, draw0, it_now, D,
        out_windows, B);
    return b200rl_check_launch("k_select_windows");
}


// Persistent host worker pool for b200rl_select_windows_host: creating and joining std::threads on every call cost
// more than the selection itself (8 threads: 0.35 ms per call for 0.19 ms of work).
namespace {
class HostPool {
public:
    ~HostPool() {
        {
            std::lock_guard<std::mutex> lk(m_);
            stop_ = true;
        }
        cv_work_.notify_all();
        for (auto& t : th_) t.join();
    }
    // runs f(0..n-1) on n pool threads and `mine` on the calling thread, returns when all are done; one caller at
    // a time
    void run(int n, const std::function<void(int)>& f, const std::function<void()>& mine) {
        std::lock_guard<std::mutex> serial(run_mu_);
        std::unique_lock<std::mutex> lk(m_);
        while ((int)th_.size() < n) {
            const int id = (int)th_.size();
            th_.emplace_back([this, id] { worker(id); });
        }
        job_ = &f;
        n_active_ = n;
        pending_ = n;
        ++generation_;
        cv_work_.notify_all();
        lk.unlock();
        mine();
        lk.lock();
        cv_done_.wait(lk, [this] { return pending_ == 0; });
        job_ = nullptr;
    }

private:
    void worker(int id) {
        uint64_t seen = 0;
        std::unique_lock<std::mutex> lk(m_);
        for (;;) {
            cv_work_.wait(lk, [&] { return stop_ || generation_ != seen; });
            if (stop_) return;
            seen = generation_;
            if (id < n_active_) {
                const std::function<void(int)>* job = job_;
                lk.unlock();
                (*job)(id);
                lk.lock();
                if (--pending_ == 0) cv_done_.notify_one();
            }
        }
    }
    std::mutex m_, run_mu_;
    std::condition_variable cv_work_, cv_done_;
    std::vector<std::thread> th_;
    const std::function<void(int)>* job_ = nullptr;
    int n_active_ = 0, pending_ = 0;
    uint64_t generation_ = 0;
    bool stop_ = false;
};
HostPool& host_pool() {
    static HostPool pool;
    return pool;
}
}  // namespace

extern "C" int b200rl_select_windows_host(const int32_t* ep_table_host, int64_t E, uint64_t cursor,
                                          int64_t maximum_episodes, int64_t forward_steps, uint64_t seed,
                                          uint64_t draw0, uint64_t it_now, int64_t D, int32_t* out_windows_host,
                                          int64_t B, int n_threads) {
    B200RL_REQUIRE(ep_table_host && out_windows_host, "select_windows_host: null pointer");
    B200RL_REQUIRE(E >= 1 && D >= 1 && maximum_episodes >= 1 && forward_steps >= 1 && B >= 0,
                   "select_windows_host: bad E/D/maximum_episodes/forward_steps/B");
    if (n_threads < 1) n_threads = 1;
    if (n_threads > 64) n_threads = 64;
    // The rows a call draws from are the newest min(cursor, E, maximum_episodes) x 16 B (1.6 MB at the reference's
    // maximum_episodes): they sit in the host's L2/L3, and a try rejected by the recency law never touches its row, so
    // r1's software prefetch (which recomputed the draws of the next windows) cost more hashing than it saved.
    auto work = [&](int64_t lo, int64_t hi) {
        for (int64_t b = lo; b < hi; ++b)
            select_one_window(ep_table_host, E, cursor, maximum_episodes, (int)forward_steps, seed, draw0, it_now, D, b,
                              out_windows_host + b * 8);
    };
    if (n_threads == 1 || B < 4096) {
        work(0, B);
        return B200RL_OK;
    }
    // Blocks of 1024 windows handed out from a shared counter (a thread that loses its core for a while — eight
    // ranks share the box's cores — no longer holds up the call), and the calling thread works too.
    constexpr int64_t BLOCK = 1024;
    std::atomic<int64_t> next{0};
    auto drain = [&] {
        for (;;) {
            const int64_t lo = next.fetch_add(BLOCK, std::memory_order_relaxed);
            if (lo >= B) break;
            work(lo, lo + BLOCK < B ? lo + BLOCK : B);
        }
    };
    host_pool().run(n_threads - 1, [&](int) { drain(); }, drain);
    return B200RL_OK;
}

extern "C" int b200rl_pad_stack(const void* src, const int64_t* offsets, const void* fill_rows, uint32_t fill_bits,
                                void* out, int64_t B, int64_t T, int64_t row_words, void* stream) {
    B200RL_REQUIRE(offsets && out && B >= 0 && T >= 1 && row_words >= 1, "pad_stack: bad arguments");
    const int64_t n = B * T * row_words;
    if (n == 0) return B200RL_OK;
    int64_t blocks = (n + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    k_pad_stack<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>((const uint32_t*)src, offsets,
                                                               (const uint32_t*)fill_rows, fill_bits, (uint32_t*)out,
                                                               B, T, row_words);
    return b200rl_check_launch("k_pad_stack");
}
