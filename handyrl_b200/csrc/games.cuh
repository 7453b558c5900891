// Per-game rules as register-resident bitboard state machines (one game per thread).
//
// Each `Game` type exposes the same static interface so that the kernels in
// env.cu (reset / step / legal / observe / fused step+observe / rollout) are
// written once:
//   State                      compact per-game state
//   A, LW                      action-space size, legal-mask words (ceil(A/32))
//   OBITS, OW                  observation bits of the main tensor, words
//   SBITS                      bits of the scalar side tensor (0 = none)
//   LMAX                       longest legal list
//   load/store                 SoA state planes in HBM
//   reset, step, terminal, turn, outcome, reward
//   legal_bits, legal_list, obs_bits
//
// Reference semantics: handyrl/envs/tictactoe.py:72-172,
// handyrl/envs/parallel_tictactoe.py:13-58, handyrl/envs/geister.py:170-538,
// handyrl/envs/ci_geister.py:518-573.
#pragma once
#include "common.cuh"

#include "../../include/b200rl.h"  // game ids and view selectors

// ---------------------------------------------------------------------------
// TicTacToe / ParallelTicTacToe
// ---------------------------------------------------------------------------
// state word: bits 0-8 'O' (BLACK, +1) cells, 9-17 'X' (WHITE, -1) cells,
// 18-21 number of applied moves (len(record)), 22-23 winner (0 none, 1 BLACK, 2 WHITE).
// cell index a = x*3 + y, the action id (tictactoe.py:106).
template <bool PARALLEL>
struct TicTacToeGame {
    struct State {
        uint32_t w;
    };
    static constexpr int A = 9, LW = 1, OBITS = 27, OW = 1, SBITS = 0, LMAX = 9, STATE_U32 = 1, PLAYERS = 2;
    static constexpr bool HAS_REWARD = false;
    static constexpr bool SIMULTANEOUS = PARALLEL;  // both players act every ply
    static constexpr int MAX_STEPS = 9;

    __device__ __forceinline__ static State load(const uint32_t* __restrict__ st, int64_t N, int64_t i) {
        State s;
        s.w = st[i];
        return s;
    }
    // L2-coherent load: the state may have been written by another SM earlier in the same kernel (rollout tickets)
    __device__ __forceinline__ static State load_cg(const uint32_t* st, int64_t N, int64_t i) {
        State s;
        s.w = __ldcg(st + i);
        return s;
    }
    __device__ __forceinline__ static void store(uint32_t* __restrict__ st, int64_t N, int64_t i, const State& s) {
        st[i] = s.w;
    }
    __device__ __forceinline__ static void reset(State& s) { s.w = 0; }
    __device__ __forceinline__ static int plies(const State& s) { return (s.w >> 18) & 15; }
    __device__ __forceinline__ static int win(const State& s) { return (s.w >> 22) & 3; }
    __device__ __forceinline__ static bool terminal(const State& s) {  // tictactoe.py:136-138
        return win(s) != 0 || plies(s) == 9;
    }
    // tictactoe.py:133-134; ParallelTicTacToe.turn() is not a player (parallel_tictactoe.py:54-55)
    __device__ __forceinline__ static int turn(const State& s) { return PARALLEL ? 0 : (plies(s) & 1); }
    __device__ __forceinline__ static int steps(const State& s) { return plies(s); }

    // tictactoe.py:103-119 / parallel_tictactoe.py:26-38.  `sel` = selected player (parallel only).
    __device__ __forceinline__ static void step(State& s, int action, int sel) {
        if (action < 0 || action >= 9) return;
        const int mover = PARALLEL ? (sel & 1) : (plies(s) & 1);  // 0 = BLACK 'O', 1 = WHITE 'X'
        uint32_t o = s.w & 0x1ff, x = (s.w >> 9) & 0x1ff;
        const uint32_t bit = 1u << action;
        uint32_t mine = mover ? x : o, other = mover ? o : x;
        mine |= bit;
        other &= ~bit;  // board[x, y] = color overwrites the cell
        // lines through the played cell (row, column, diagonals when on them)
        const int cx = action / 3, cy = action - cx * 3;
        const uint32_t row = 0x7u << (cx * 3), col = 0x49u << cy;
        bool w = (mine & row) == row || (mine & col) == col;
        if (cx == cy) w = w || (mine & 0x111u) == 0x111u;
        if (cx == 2 - cy) w = w || (mine & 0x54u) == 0x54u;
        uint32_t wn = win(s);
        if (w) wn = mover ? 2 : 1;
        o = mover ? other : mine;
        x = mover ? mine : other;
        s.w = o | (x << 9) | ((uint32_t)(plies(s) + 1) << 18) | (wn << 22);
    }
    __device__ __forceinline__ static void outcome(const State& s, float& o0, float& o1) {  // tictactoe.py:140-147
        const int wn = win(s);
        o0 = wn == 1 ? 1.f : (wn == 2 ? -1.f : 0.f);
        o1 = -o0;
        if (wn == 0) o0 = o1 = 0.f;
    }
    __device__ __forceinline__ static float reward(const State&) { return 0.f; }  // no reward() override -> {}
    __device__ __forceinline__ static void legal_bits(const State& s, uint32_t (&m)[LW]) {  // tictactoe.py:149-151
        m[0] = ~((s.w & 0x1ff) | ((s.w >> 9) & 0x1ff)) & 0x1ff;
    }
    __device__ __forceinline__ static int legal_list(const State& s, int* out) {
        uint32_t m[LW];
        legal_bits(s, m);
        int n = 0;
        for (int a = 0; a < 9; ++a)
            if ((m[0] >> a) & 1) out[n++] = a;
        return n;
    }
    // compact trajectory record: r[0..4] hold the state needed to re-encode obs / legal mask
    __device__ __forceinline__ static void pack_record(const State& s, uint32_t (&r)[8]) {
        r[0] = s.w; r[1] = r[2] = r[3] = r[4] = 0;
    }
    __device__ __forceinline__ static State unpack_record(const uint32_t (&r)[8]) {
        State s;
        s.w = r[0];
        return s;
    }
    // tictactoe.py:163-172.  view: -1 = None, 0/1 = that player, 2 = turn player.
    __device__ __forceinline__ static void obs_bits(const State& s, int view, uint32_t (&ob)[OW], uint32_t& scalar) {
        const uint32_t o = s.w & 0x1ff, x = (s.w >> 9) & 0x1ff;
        bool turn_view;
        if (PARALLEL)
            turn_view = (view == B200RL_VIEW_NONE);  // player == turn() is never true
        else
            turn_view = (view == B200RL_VIEW_NONE) || (view == B200RL_VIEW_TURN) || (view == (plies(s) & 1));
        // self.color: BLACK on even plies; ParallelTicTacToe never flips it
        const bool color_black = PARALLEL ? true : ((plies(s) & 1) == 0);
        const bool c_black = turn_view ? color_black : !color_black;
        const uint32_t mine = c_black ? o : x, opp = c_black ? x : o;
        ob[0] = (turn_view ? 0x1ffu : 0u) | (mine << 9) | (opp << 18);
        scalar = 0;
    }
};

// ---------------------------------------------------------------------------
// Geister / CIGeister
// ---------------------------------------------------------------------------
// cell index i = x*6 + y (x = letter A-F, y = digit 1-6), geister.py:279.
__constant__ uint8_t c_geister_oseq[70] = {
    15,  23,  39,  71,  135, 27,  43,  75,  139, 51,  83,  147, 99,  163, 195, 29,  45,  77,
    141, 53,  85,  149, 101, 165, 197, 57,  89,  153, 105, 169, 201, 113, 177, 209, 225, 30,
    46,  78,  142, 54,  86,  150, 102, 166, 198, 58,  90,  154, 106, 170, 202, 114, 178, 210,
    226, 60,  92,  156, 108, 172, 204, 116, 180, 212, 228, 120, 184, 216, 232, 240};  // geister.py:189
// geister.py:178-181 start cells of piece idx 0..7
__constant__ uint8_t c_geister_opos[2][8] = {{7, 13, 19, 25, 6, 12, 18, 24}, {28, 22, 16, 10, 29, 23, 17, 11}};

#define G_M36 0xFFFFFFFFFull
#define G_COL_Y0 0x041041041ull  // cells with y == 0 (i % 6 == 0)
#define G_COL_Y5 0x820820820ull  // cells with y == 5

__device__ __forceinline__ uint64_t g_rev36(uint64_t x) { return __brevll(x) >> 28; }

template <bool CI>
struct GeisterGame {
    // bb[c*2+t]: 36-bit occupancy of colour c (0 BLACK, 1 WHITE) type t (0 BLUE, 1 RED)
    // pos: 16 bytes, piece id c*8+idx -> cell, 0xFF = absent
    // tc: turn_count + 2 (0,1 = setup plies); win: 0 none, 1 BLACK, 2 WHITE, 3 draw
    // aux: bits 0-1 captured_type+1 of the last ply, bits 8-15 / 16-23 layout+1 of BLACK / WHITE
    struct State {
        uint64_t bb[4];
        uint32_t pos[4];
        uint32_t tc, win, aux, spare;
    };
    static constexpr int A = 214, LW = 7, OBITS = 252, OW = 8, SBITS = 18, LMAX = 70, STATE_U32 = 16, PLAYERS = 2;
    static constexpr bool HAS_REWARD = true;
    static constexpr bool SIMULTANEOUS = false;
    static constexpr int MAX_STEPS = 202;

    __device__ __forceinline__ static State load(const uint32_t* __restrict__ st, int64_t N, int64_t i) {
        const uint4* p = reinterpret_cast<const uint4*>(st);
        uint4 a = p[i], b = p[N + i], c = p[2 * N + i], d = p[3 * N + i];
        State s;
        s.bb[0] = a.x | ((uint64_t)a.y << 32);
        s.bb[1] = a.z | ((uint64_t)a.w << 32);
        s.bb[2] = b.x | ((uint64_t)b.y << 32);
        s.bb[3] = b.z | ((uint64_t)b.w << 32);
        s.pos[0] = c.x; s.pos[1] = c.y; s.pos[2] = c.z; s.pos[3] = c.w;
        s.tc = d.x; s.win = d.y; s.aux = d.z; s.spare = d.w;
        return s;
    }
    __device__ __forceinline__ static State load_cg(const uint32_t* st, int64_t N, int64_t i) {
        const uint4* p = reinterpret_cast<const uint4*>(st);
        uint4 a = __ldcg(p + i), b = __ldcg(p + N + i), c = __ldcg(p + 2 * N + i), d = __ldcg(p + 3 * N + i);
        State s;
        s.bb[0] = a.x | ((uint64_t)a.y << 32);
        s.bb[1] = a.z | ((uint64_t)a.w << 32);
        s.bb[2] = b.x | ((uint64_t)b.y << 32);
        s.bb[3] = b.z | ((uint64_t)b.w << 32);
        s.pos[0] = c.x; s.pos[1] = c.y; s.pos[2] = c.z; s.pos[3] = c.w;
        s.tc = d.x; s.win = d.y; s.aux = d.z; s.spare = d.w;
        return s;
    }
    __device__ __forceinline__ static void store(uint32_t* __restrict__ st, int64_t N, int64_t i, const State& s) {
        uint4* p = reinterpret_cast<uint4*>(st);
        p[i] = make_uint4((uint32_t)s.bb[0], (uint32_t)(s.bb[0] >> 32), (uint32_t)s.bb[1], (uint32_t)(s.bb[1] >> 32));
        p[N + i] = make_uint4((uint32_t)s.bb[2], (uint32_t)(s.bb[2] >> 32), (uint32_t)s.bb[3], (uint32_t)(s.bb[3] >> 32));
        p[2 * N + i] = make_uint4(s.pos[0], s.pos[1], s.pos[2], s.pos[3]);
        p[3 * N + i] = make_uint4(s.tc, s.win, s.aux, s.spare);
    }
    __device__ __forceinline__ static void reset(State& s) {  // geister.py:195-206
        s.bb[0] = s.bb[1] = s.bb[2] = s.bb[3] = 0;
        s.pos[0] = s.pos[1] = s.pos[2] = s.pos[3] = 0xFFFFFFFFu;
        s.tc = 0; s.win = 0; s.aux = 0; s.spare = 0;
    }
    __device__ __forceinline__ static bool terminal(const State& s) { return s.win != 0; }  // geister.py:430-432
    __device__ __forceinline__ static int turn(const State& s) { return s.tc & 1; }          // geister.py:427-428
    __device__ __forceinline__ static int steps(const State& s) { return s.tc; }
    __device__ __forceinline__ static float reward(const State&) { return -0.01f; }          // geister.py:434-436
    __device__ __forceinline__ static void outcome(const State& s, float& o0, float& o1) {   // geister.py:438-445
        o0 = s.win == 1 ? 1.f : (s.win == 2 ? -1.f : 0.f);
        o1 = s.win == 1 ? -1.f : (s.win == 2 ? 1.f : 0.f);
    }

    // Plane / position-word selection without dynamic indexing (keeps State in registers).
    __device__ __forceinline__ static uint64_t bbsel(const State& s, int k) {
        const uint64_t lo = (k & 1) ? s.bb[1] : s.bb[0], hi = (k & 1) ? s.bb[3] : s.bb[2];
        return (k & 2) ? hi : lo;
    }
    __device__ __forceinline__ static uint32_t possel(const State& s, int k) {
        const uint32_t lo = (k & 1) ? s.pos[1] : s.pos[0], hi = (k & 1) ? s.pos[3] : s.pos[2];
        return (k & 2) ? hi : lo;
    }

    // piece_cnt[k] of the reference.  `spare` holds four signed byte corrections that are non-zero only
    // in network-match mirror environments, where update() rewrites a captured piece's type on the
    // board without touching piece_cnt (geister.py:416-424).
    __device__ __forceinline__ static int cnt(const State& s, int k) {
        return __popcll(bbsel(s, k)) + (int)(int8_t)((s.spare >> (8 * k)) & 0xFF);
    }

    __device__ __forceinline__ static void pos_replace(State& s, uint32_t cell, uint32_t newv) {
        const uint32_t rep = cell * 0x01010101u, nv = newv * 0x01010101u;
#pragma unroll
        for (int w = 0; w < 4; ++w) {
            uint32_t m = __vcmpeq4(s.pos[w], rep);
            s.pos[w] = (s.pos[w] & ~m) | (nv & m);
        }
    }

    // geister.py:359-396 (play) and 351-357 / 229-236 (_set / set_pieces)
    __device__ __forceinline__ static void step(State& s, int action, int /*sel*/) {
        if (action < 0 || action >= A) return;
        const int c = s.tc & 1;
        if (s.tc < 2) {
            int layout = action - 144;
            if (layout < 0 || layout >= 70) return;
            const uint32_t blue = c_geister_oseq[layout];
            uint64_t bbl = 0, bbr = 0;
            uint32_t p0 = 0, p1 = 0;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const uint32_t cell = c_geister_opos[c][k];
                if ((blue >> k) & 1) bbl |= 1ull << cell; else bbr |= 1ull << cell;
                if (k < 4) p0 |= cell << (8 * k); else p1 |= cell << (8 * (k - 4));
            }
            if (c == 0) {
                s.bb[0] |= bbl; s.bb[1] |= bbr; s.pos[0] = p0; s.pos[1] = p1;
            } else {
                s.bb[2] |= bbl; s.bb[3] |= bbr; s.pos[2] = p0; s.pos[3] = p1;
            }
            s.aux = (s.aux & ~(0xFFu << (8 + 8 * c))) | ((uint32_t)(layout + 1) << (8 + 8 * c));
            s.tc += 1;
            return;
        }
        int from = action % 36, d = action / 36;
        if (c == 1) { from = 35 - from; d = 3 - d; }  // geister.py:282-292
        const int x = from / 6, y = from - x * 6;
        const int nx = x + (d == 0 ? -1 : (d == 3 ? 1 : 0));
        const int ny = y + (d == 1 ? -1 : (d == 2 ? 1 : 0));
        const bool onboard = nx >= 0 && nx < 6 && ny >= 0 && ny < 6;
        const uint64_t fbit = 1ull << from;
        int piece = -1;
#pragma unroll
        for (int k = 0; k < 4; ++k)
            if (s.bb[k] & fbit) piece = k;
        if (piece < 0) return;  // no piece on the from-cell: not a legal action, ignored
        uint32_t captured = 0;
        if (!onboard) {  // finish by goal (geister.py:370-373)
#pragma unroll
            for (int k = 0; k < 4; ++k) s.bb[k] &= ~fbit;  // only plane `piece` holds the bit
            pos_replace(s, from, 0xFF);
            s.win = c + 1;
        } else {
            const int to = nx * 6 + ny;
            const uint64_t tbit = 1ull << to;
            int cap = -1;
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (s.bb[k] & tbit) cap = k;
            if (cap >= 0) {  // geister.py:376-385
#pragma unroll
                for (int k = 0; k < 4; ++k) s.bb[k] &= ~tbit;  // only plane `cap` holds the bit
                pos_replace(s, to, 0xFF);
                if (cnt(s, cap) == 0) s.win = ((cap & 1) == 0) ? (c + 1) : (2 - c);
                captured = (cap & 1) + 1;
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) s.bb[k] = (s.bb[k] & ~fbit) | (k == piece ? tbit : 0ull);
            pos_replace(s, from, to);
        }
        s.aux = (s.aux & ~3u) | captured;
        s.tc += 1;
        if ((int)s.tc - 2 >= 200 && s.win == 0) s.win = 3;  // geister.py:395-396
    }

    // From-cell sets per ABSOLUTE direction d (geister.py:447-470).
    __device__ __forceinline__ static void from_sets(const State& s, uint64_t (&F)[4]) {
        const int c = s.tc & 1;
        const uint64_t own = c ? (s.bb[2] | s.bb[3]) : (s.bb[0] | s.bb[1]);
        const uint64_t fre = ~own & G_M36;
        F[0] = own & (fre << 6) & G_M36;
        F[1] = own & (fre << 1) & ~G_COL_Y0 & G_M36;
        F[2] = own & (fre >> 1) & ~G_COL_Y5;
        F[3] = own & (fre >> 6);
        const uint64_t blue = c ? s.bb[2] : s.bb[0];
        // goal exits: BLACK leaves sideways from (0,5)/(5,5), WHITE from (0,0)/(5,0) (geister.py:183-186)
        if (c == 0) {
            F[0] |= blue & (1ull << 5);
            F[3] |= blue & (1ull << 35);
        } else {
            F[0] |= blue & (1ull << 0);
            F[3] |= blue & (1ull << 30);
        }
    }

    // 214-bit legal mask in action-id order (geister.py:275-280, 472-486).
    __device__ __forceinline__ static void legal_bits(const State& s, uint32_t (&m)[LW]) {
        if (s.tc < 2) {
            m[0] = m[1] = m[2] = m[3] = 0;
            m[4] = 0xFFFF0000u; m[5] = 0xFFFFFFFFu; m[6] = 0x003FFFFFu;
            return;
        }
        uint64_t F[4], Aq[4];
        from_sets(s, F);
        if ((s.tc & 1) == 0) {
            Aq[0] = F[0]; Aq[1] = F[1]; Aq[2] = F[2]; Aq[3] = F[3];
        } else {  // WHITE ids are 180-degree rotated: cell 35-i, direction 3-d
            Aq[0] = g_rev36(F[3]); Aq[1] = g_rev36(F[2]); Aq[2] = g_rev36(F[1]); Aq[3] = g_rev36(F[0]);
        }
        m[0] = (uint32_t)Aq[0];
        m[1] = (uint32_t)(Aq[0] >> 32) | (uint32_t)(Aq[1] << 4);
        m[2] = (uint32_t)(Aq[1] >> 28) | (uint32_t)(Aq[2] << 8);
        m[3] = (uint32_t)(Aq[2] >> 24) | (uint32_t)(Aq[3] << 12);
        m[4] = (uint32_t)(Aq[3] >> 20);
        m[5] = 0; m[6] = 0;
    }

    // Legal list in the reference's order: piece-id major, absolute direction minor.
    __device__ __forceinline__ static int legal_list(const State& s, int* out) {
        int n = 0;
        if (s.tc < 2) {
            for (int i = 0; i < 70; ++i) out[n++] = 144 + i;
            return n;
        }
        uint64_t F[4];
        from_sets(s, F);
        const int c = s.tc & 1;
        for (int k = 0; k < 8; ++k) {
            const uint32_t p = (possel(s, c * 2 + (k >> 2)) >> (8 * (k & 3))) & 0xFF;
            if (p == 0xFF) continue;
            for (int d = 0; d < 4; ++d)
                if ((F[d] >> p) & 1) out[n++] = c ? ((3 - d) * 36 + 35 - (int)p) : (d * 36 + (int)p);
        }
        return n;
    }

    __device__ __forceinline__ static void pack_record(const State& s, uint32_t (&r)[8]) {
        r[0] = (uint32_t)s.bb[0]; r[1] = (uint32_t)s.bb[1]; r[2] = (uint32_t)s.bb[2]; r[3] = (uint32_t)s.bb[3];
        r[4] = (uint32_t)(s.bb[0] >> 32) | ((uint32_t)(s.bb[1] >> 32) << 4) | ((uint32_t)(s.bb[2] >> 32) << 8) |
               ((uint32_t)(s.bb[3] >> 32) << 12) | (s.tc << 16);
    }
    __device__ __forceinline__ static State unpack_record(const uint32_t (&r)[8]) {
        State s;
        reset(s);
        s.bb[0] = r[0] | ((uint64_t)(r[4] & 0xF) << 32);
        s.bb[1] = r[1] | ((uint64_t)((r[4] >> 4) & 0xF) << 32);
        s.bb[2] = r[2] | ((uint64_t)((r[4] >> 8) & 0xF) << 32);
        s.bb[3] = r[3] | ((uint64_t)((r[4] >> 12) & 0xF) << 32);
        s.tc = (r[4] >> 16) & 0xFF;
        return s;
    }

    __device__ __forceinline__ static uint32_t onehot4(int n) { return (n >= 1 && n <= 4) ? (1u << (n - 1)) : 0u; }

    // geister.py:495-538 / ci_geister.py:518-573: 7x36 board bits (plane-major) + 18 scalar bits.
    __device__ __forceinline__ static void obs_bits(const State& s, int view, uint32_t (&ob)[OW], uint32_t& scalar) {
        const int tp = s.tc & 1;
        const bool turn_view = (view == B200RL_VIEW_NONE) || (view == B200RL_VIEW_TURN) || (view == tp);
        const int vc = turn_view ? tp : 1 - tp;
        const int oc = 1 - vc;
        const uint64_t cb = vc ? s.bb[2] : s.bb[0], cr = vc ? s.bb[3] : s.bb[1];
        const uint64_t obl = vc ? s.bb[0] : s.bb[2], ore = vc ? s.bb[1] : s.bb[3];
        scalar = (vc == 0 ? 1u : 0u) | (turn_view ? 2u : 0u) | (onehot4(cnt(s, vc * 2)) << 2) |
                 (onehot4(cnt(s, vc * 2 + 1)) << 6) | (onehot4(cnt(s, oc * 2)) << 10) |
                 (onehot4(cnt(s, oc * 2 + 1)) << 14);
        const bool show = CI || (view == B200RL_VIEW_NONE);
        uint64_t pl[7] = {G_M36, cb | cr, obl | ore, cb, cr, show ? obl : 0ull, show ? ore : 0ull};
        if (vc == 1) {
#pragma unroll
            for (int k = 1; k < 7; ++k) pl[k] = g_rev36(pl[k]);
        }
        // concatenate 7 x 36 bits: plane k occupies bits [36k, 36k+36)
        ob[0] = (uint32_t)pl[0];
        ob[1] = (uint32_t)(pl[0] >> 32) | (uint32_t)(pl[1] << 4);
        ob[2] = (uint32_t)(pl[1] >> 28) | (uint32_t)(pl[2] << 8);
        ob[3] = (uint32_t)(pl[2] >> 24) | (uint32_t)(pl[3] << 12);
        ob[4] = (uint32_t)(pl[3] >> 20) | (uint32_t)(pl[4] << 16);
        ob[5] = (uint32_t)(pl[4] >> 16) | (uint32_t)(pl[5] << 20);
        ob[6] = (uint32_t)(pl[5] >> 12) | (uint32_t)(pl[6] << 24);
        ob[7] = (uint32_t)(pl[6] >> 8);
    }
};
