/*
 * CPU oracle for the game rules — TEST INFRASTRUCTURE (see oracle/__init__.py).
 *
 * Plain-C restatement of the reference environments, keeping the reference's
 * own data structures (cell arrays, piece_position, piece_cnt ...) rather than
 * the bitboards of the CUDA engine, so the two implementations are independent:
 *   TicTacToe          handyrl/envs/tictactoe.py:72-172
 *   ParallelTicTacToe  handyrl/envs/parallel_tictactoe.py:13-58
 *   Geister            handyrl/envs/geister.py:170-538
 *   CIGeister          handyrl/envs/ci_geister.py:518-573 (observation only differs)
 * Pinned against tapes recorded from the imported reference (tests/golden/tape_*.npz,
 * tools/make_golden.py) by tests/test_games_oracle.py.
 *
 * Also holds the CPU baseline driver (orc_selfplay): the body of
 * handyrl/generation.py:31-68 with the epoch-0 uniform policy, one game at a time.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

enum { G_TTT = 0, G_PTTT = 1, G_GEISTER = 2, G_CIGEISTER = 3 };

typedef struct {
    int game;
    /* tictactoe */
    int tboard[3][3]; /* 0, +1 BLACK, -1 WHITE */
    int tcolor, twin, trecord;
    /* geister */
    int board[6][6];       /* -1 empty or colour*2+type */
    int board_index[6][6]; /* piece id or -1 */
    int piece_position[16][2];
    int piece_cnt[4];
    int color, turn_count, win_color; /* win_color: -1 None, 0, 1, 2 draw */
    int captured_type;                /* -1 None */
    int layouts[2];
} orc_env;

static const int GD[4][2] = {{-1, 0}, {0, -1}, {0, 1}, {1, 0}};            /* geister.py:188 */
static const int GGOAL[2][2][2] = {{{-1, 5}, {6, 5}}, {{-1, 0}, {6, 0}}};  /* geister.py:183-186 */
/* geister.py:178-181, as (x, y) = (letter, digit) */
static const int GOPOS[2][8][2] = {{{1, 1}, {2, 1}, {3, 1}, {4, 1}, {1, 0}, {2, 0}, {3, 0}, {4, 0}},
                                   {{4, 4}, {3, 4}, {2, 4}, {1, 4}, {4, 5}, {3, 5}, {2, 5}, {1, 5}}};

int orc_sizeof(void) { return (int)sizeof(orc_env); }

void orc_reset(orc_env* e, int game) {
    memset(e, 0, sizeof(*e));
    e->game = game;
    e->tcolor = 1; /* BLACK (tictactoe.py:83) */
    for (int x = 0; x < 6; ++x)
        for (int y = 0; y < 6; ++y) e->board[x][y] = e->board_index[x][y] = -1;
    e->color = 0;
    e->turn_count = -2; /* geister.py:199 */
    e->win_color = -1;
    e->captured_type = -1;
    e->layouts[0] = e->layouts[1] = -1;
}

static int is_geister(const orc_env* e) { return e->game >= G_GEISTER; }

/* ---------------- tictactoe ---------------- */
static void ttt_place(orc_env* e, int action, int color) { /* tictactoe.py:103-119 */
    int x = action / 3, y = action % 3;
    e->tboard[x][y] = color;
    int row = e->tboard[x][0] + e->tboard[x][1] + e->tboard[x][2];
    int col = e->tboard[0][y] + e->tboard[1][y] + e->tboard[2][y];
    int d0 = e->tboard[0][0] + e->tboard[1][1] + e->tboard[2][2];
    int d1 = e->tboard[2][0] + e->tboard[1][1] + e->tboard[0][2];
    int win = row == 3 * color || col == 3 * color || (x == y && d0 == 3 * color) || (x == 2 - y && d1 == 3 * color);
    if (win) e->twin = color;
}

/* ---------------- geister ---------------- */
static int onboard(int x, int y) { return 0 <= x && x < 6 && 0 <= y && y < 6; }

static void put_piece(orc_env* e, int piece, int x, int y, int idx) { /* geister.py:208-212 */
    e->board[x][y] = piece;
    e->piece_position[idx][0] = x;
    e->piece_position[idx][1] = y;
    e->board_index[x][y] = idx;
    e->piece_cnt[piece] += 1;
}
static void remove_piece(orc_env* e, int piece, int x, int y) { /* geister.py:214-219 */
    e->board[x][y] = -1;
    int idx = e->board_index[x][y];
    e->board_index[x][y] = -1;
    e->piece_position[idx][0] = e->piece_position[idx][1] = -1;
    e->piece_cnt[piece] -= 1;
}
static void move_piece(orc_env* e, int piece, int ox, int oy, int nx, int ny) { /* geister.py:221-227 */
    e->board[ox][oy] = -1;
    e->board[nx][ny] = piece;
    int idx = e->board_index[ox][oy];
    e->board_index[ox][oy] = -1;
    e->board_index[nx][ny] = idx;
    e->piece_position[idx][0] = nx;
    e->piece_position[idx][1] = ny;
}

/* s-th lexicographic 4-subset of {0..7} (geister.py:189) -> membership flags */
static void oseq(int s, int chosen[8]) {
    int cnt = 0;
    memset(chosen, 0, 8 * sizeof(int));
    for (int a = 0; a < 8; ++a)
        for (int b = a + 1; b < 8; ++b)
            for (int c = b + 1; c < 8; ++c)
                for (int d = c + 1; d < 8; ++d) {
                    if (cnt == s) { chosen[a] = chosen[b] = chosen[c] = chosen[d] = 1; return; }
                    ++cnt;
                }
}

static void geister_set(orc_env* e, int layout) { /* geister.py:351-357, 229-236 */
    e->layouts[e->color] = layout;
    int chosen[8];
    oseq(layout, chosen);
    for (int idx = 0; idx < 8; ++idx) {
        int t = chosen[idx] ? 0 : 1;
        put_piece(e, e->color * 2 + t, GOPOS[e->color][idx][0], GOPOS[e->color][idx][1], e->color * 8 + idx);
    }
    e->color = 1 - e->color;
    e->turn_count += 1;
}

static void action2from(int a, int c, int* x, int* y) { /* geister.py:281-286 */
    int pos1d = a % 36;
    *x = pos1d / 6;
    *y = pos1d % 6;
    if (c == 1) { *x = 5 - *x; *y = 5 - *y; }
}
static int action2direction(int a, int c) { /* geister.py:288-292 */
    int d = a / 36;
    return c == 1 ? 3 - d : d;
}
static int fromdirection2action(int x, int y, int d, int c) { /* geister.py:275-279 */
    if (c == 1) { x = 5 - x; y = 5 - y; d = 3 - d; }
    return d * 36 + x * 6 + y;
}

static void geister_play(orc_env* e, int action) { /* geister.py:359-396 */
    if (e->turn_count < 0) { geister_set(e, action - 144); return; }
    int ox, oy;
    action2from(action, e->color, &ox, &oy);
    int d = action2direction(action, e->color);
    int nx = ox + GD[d][0], ny = oy + GD[d][1];
    int piece = e->board[ox][oy];
    e->captured_type = -1;
    if (!onboard(nx, ny)) {
        remove_piece(e, piece, ox, oy);
        e->win_color = e->color;
    } else {
        int cap = e->board[nx][ny];
        if (cap != -1) {
            remove_piece(e, cap, nx, ny);
            if (e->piece_cnt[cap] == 0) e->win_color = (cap % 2 == 0) ? e->color : 1 - e->color;
            e->captured_type = cap % 2;
        }
        move_piece(e, piece, ox, oy, nx, ny);
    }
    e->color = 1 - e->color;
    e->turn_count += 1;
    if (e->turn_count >= 200 && e->win_color == -1) e->win_color = 2;
}

static int geister_legal(const orc_env* e, int c, int t, int nx, int ny) { /* geister.py:463-470 */
    if (onboard(nx, ny)) {
        int cap = e->board[nx][ny];
        int capc = cap == -1 ? -1 : cap / 2;
        return capc != c;
    }
    if (t != 0) return 0;
    for (int g = 0; g < 2; ++g)
        if (GGOAL[c][g][0] == nx && GGOAL[c][g][1] == ny) return 1;
    return 0;
}

/* ---------------- common interface ---------------- */
int orc_turn(const orc_env* e) {
    if (is_geister(e)) return ((e->turn_count % 2) + 2) % 2; /* Python modulo, geister.py:427-428 */
    if (e->game == G_PTTT) return 0;                          /* not a player (parallel_tictactoe.py:54-55) */
    return e->trecord % 2;
}

int orc_terminal(const orc_env* e) {
    if (is_geister(e)) return e->win_color != -1;
    return e->twin != 0 || e->trecord == 9;
}

int orc_steps(const orc_env* e) { return is_geister(e) ? e->turn_count + 2 : e->trecord; }

/* step({p: a}); sel = selected player for ParallelTicTacToe (parallel_tictactoe.py:20-24) */
void orc_step(orc_env* e, int action, int sel) {
    if (is_geister(e)) {
        geister_play(e, action);
    } else if (e->game == G_PTTT) {
        ttt_place(e, action, sel == 0 ? 1 : -1);
        e->trecord += 1;
    } else {
        ttt_place(e, action, e->tcolor);
        e->tcolor = -e->tcolor;
        e->trecord += 1;
    }
}

float orc_reward(const orc_env* e) { return is_geister(e) ? -0.01f : 0.0f; } /* geister.py:434-436 */

void orc_outcome(const orc_env* e, float* out) {
    out[0] = out[1] = 0.f;
    if (is_geister(e)) { /* geister.py:438-445 */
        if (e->win_color == 0) { out[0] = 1; out[1] = -1; }
        else if (e->win_color == 1) { out[0] = -1; out[1] = 1; }
    } else { /* tictactoe.py:140-147 */
        if (e->twin > 0) { out[0] = 1; out[1] = -1; }
        if (e->twin < 0) { out[0] = -1; out[1] = 1; }
    }
}

int orc_action_length(const orc_env* e) { return is_geister(e) ? 214 : 9; }

/* legal_actions() in the reference's list order; returns the count */
int orc_legal_actions(const orc_env* e, int* out) {
    int n = 0;
    if (!is_geister(e)) { /* tictactoe.py:149-151 */
        for (int a = 0; a < 9; ++a)
            if (e->tboard[a / 3][a % 3] == 0) out[n++] = a;
        return n;
    }
    if (e->turn_count < 0) { /* geister.py:474-475 */
        for (int i = 0; i < 70; ++i) out[n++] = 144 + i;
        return n;
    }
    for (int k = 0; k < 8; ++k) { /* geister.py:477-484 */
        const int* pos = e->piece_position[e->color * 8 + k];
        if (pos[0] == -1) continue;
        int t = e->board[pos[0]][pos[1]] % 2;
        for (int d = 0; d < 4; ++d)
            if (geister_legal(e, e->color, t, pos[0] + GD[d][0], pos[1] + GD[d][1]))
                out[n++] = fromdirection2action(pos[0], pos[1], d, e->color);
    }
    return n;
}

/* observation(player); player = -1 for None.  main: 27 (ttt) / 252 (geister) floats, scalar: 18 floats */
void orc_observation(const orc_env* e, int player, float* main, float* scalar) {
    if (!is_geister(e)) { /* tictactoe.py:163-172 */
        int turn_view = player < 0 || (e->game != G_PTTT && player == orc_turn(e));
        int color = turn_view ? e->tcolor : -e->tcolor;
        for (int x = 0; x < 3; ++x)
            for (int y = 0; y < 3; ++y) {
                main[0 * 9 + x * 3 + y] = turn_view ? 1.f : 0.f;
                main[1 * 9 + x * 3 + y] = e->tboard[x][y] == color ? 1.f : 0.f;
                main[2 * 9 + x * 3 + y] = e->tboard[x][y] == -color ? 1.f : 0.f;
            }
        return;
    }
    /* geister.py:495-538 */
    int turn_view = player < 0 || player == orc_turn(e);
    int color = turn_view ? e->color : 1 - e->color;
    int opp = 1 - color;
    int nb = e->piece_cnt[color * 2], nr = e->piece_cnt[color * 2 + 1];
    int ob = e->piece_cnt[opp * 2], orr = e->piece_cnt[opp * 2 + 1];
    scalar[0] = color == 0 ? 1.f : 0.f;
    scalar[1] = turn_view ? 1.f : 0.f;
    for (int i = 1; i <= 4; ++i) {
        scalar[1 + i] = nb == i ? 1.f : 0.f;
        scalar[5 + i] = nr == i ? 1.f : 0.f;
        scalar[9 + i] = ob == i ? 1.f : 0.f;
        scalar[13 + i] = orr == i ? 1.f : 0.f;
    }
    int show = (e->game == G_CIGEISTER) || player < 0; /* ci_geister.py:562-566 vs geister.py:531-532 */
    float b[7][6][6];
    for (int x = 0; x < 6; ++x)
        for (int y = 0; y < 6; ++y) {
            int p = e->board[x][y];
            int bc = p == color * 2, rc = p == color * 2 + 1, bo = p == opp * 2, ro = p == opp * 2 + 1;
            b[0][x][y] = 1.f;
            b[1][x][y] = (bc || rc) ? 1.f : 0.f;
            b[2][x][y] = (bo || ro) ? 1.f : 0.f;
            b[3][x][y] = bc ? 1.f : 0.f;
            b[4][x][y] = rc ? 1.f : 0.f;
            b[5][x][y] = (show && bo) ? 1.f : 0.f;
            b[6][x][y] = (show && ro) ? 1.f : 0.f;
        }
    for (int k = 0; k < 7; ++k)
        for (int x = 0; x < 6; ++x)
            for (int y = 0; y < 6; ++y)
                main[k * 36 + x * 6 + y] = color == 1 ? b[k][5 - x][5 - y] : b[k][x][y]; /* rot90 k=2 */
}

/* ---------------- batch helpers (tests and the timed CPU baseline) ---------------- */
static uint64_t rng_next(uint64_t* s) { /* xorshift64* */
    uint64_t x = *s;
    x ^= x >> 12; x ^= x << 25; x ^= x >> 27;
    *s = x;
    return x * 0x2545F4914F6CDD1Dull;
}

static void pack_bits(const float* f, int n, uint8_t* out, int nbytes) {
    memset(out, 0, nbytes);
    for (int i = 0; i < n; ++i)
        if (f[i] != 0.f) out[i >> 3] |= (uint8_t)(1u << (i & 7));
}

/*
 * Play n_games random games and record every state (same columns as
 * tests/golden/tape_*.npz).  Returns the number of states written (<= cap).
 */
int64_t orc_random_tapes(int game, int n_games, uint64_t seed, int64_t cap, int32_t* game_id, int32_t* action,
                         int32_t* sel, int32_t* turn, int32_t* terminal, float* reward, int32_t* legal_count,
                         int32_t* legal_ids, int lmax, uint8_t* obs0, uint8_t* obs1, uint8_t* obsN, uint8_t* sc0,
                         uint8_t* sc1, uint8_t* scN, float* outcome, int32_t* steps) {
    const int geister = game >= G_GEISTER;
    const int nmain = geister ? 252 : 27, mb = geister ? 32 : 4, sb = 3;
    int64_t n = 0;
    float main[252], scalar[18];
    int legal[80];
    for (int g = 0; g < n_games; ++g) {
        uint64_t rs = seed * 0x9E3779B97F4A7C15ull + (uint64_t)g * 0xD1B54A32D192ED03ull + 1;
        orc_env e;
        orc_reset(&e, game);
        int act = -1, sl = 0, st = 0;
        float rw = 0.f;
        for (;;) {
            if (n >= cap) return n;
            game_id[n] = g; action[n] = act; sel[n] = sl; turn[n] = orc_turn(&e);
            terminal[n] = orc_terminal(&e); reward[n] = rw;
            int cnt = orc_legal_actions(&e, legal);
            legal_count[n] = cnt;
            for (int k = 0; k < lmax; ++k) legal_ids[n * lmax + k] = k < cnt ? legal[k] : -1;
            memset(scalar, 0, sizeof(scalar));
            orc_observation(&e, 0, main, scalar);
            pack_bits(main, nmain, obs0 + n * mb, mb); pack_bits(scalar, 18, sc0 + n * sb, sb);
            orc_observation(&e, 1, main, scalar);
            pack_bits(main, nmain, obs1 + n * mb, mb); pack_bits(scalar, 18, sc1 + n * sb, sb);
            orc_observation(&e, -1, main, scalar);
            pack_bits(main, nmain, obsN + n * mb, mb); pack_bits(scalar, 18, scN + n * sb, sb);
            ++n;
            if (orc_terminal(&e)) break;
            act = legal[rng_next(&rs) % (uint64_t)cnt];
            sl = (int)(rng_next(&rs) & 1);
            if (game == G_PTTT && sl == 1) act = legal[rng_next(&rs) % (uint64_t)cnt]; /* the other player's pick */
            orc_step(&e, act, sl);
            rw = orc_reward(&e);
            ++st;
        }
        orc_outcome(&e, outcome + 2 * g);
        steps[g] = st;
    }
    return n;
}

/*
 * CPU baseline: the self-play loop of handyrl/generation.py:31-68 with the epoch-0
 * uniform policy, for `n_steps` env-steps on each of `n_threads` pthreads.
 * Per ply it produces what the reference stores in a moment: observation (turn
 * player), action_mask (1e32 / 0), policy (zero logits - mask) and the action.
 * Returns the total number of env-steps executed; `checksum` defeats dead-code removal.
 */
#include <pthread.h>

typedef struct {
    int game, tid;
    int64_t n_steps, done;
    uint64_t seed;
    double csum;
} sp_job;

static void* sp_worker(void* arg) {
    sp_job* j = (sp_job*)arg;
    const int game = j->game;
    uint64_t rs = j->seed * 0x9E3779B97F4A7C15ull + (uint64_t)j->tid * 0xD1B54A32D192ED03ull + 1;
    const int A = game >= G_GEISTER ? 214 : 9;
    enum { RING = 256, MSZ = 270 + 214 + 214 + 2 };
    float* moments = (float*)malloc(sizeof(float) * RING * MSZ);
    orc_env e;
    orc_reset(&e, game);
    int legal[80];
    double csum = 0.0;
    for (int64_t k = 0; k < j->n_steps; ++k) {
        if (orc_terminal(&e)) orc_reset(&e, game);
        float* m = moments + (k % RING) * MSZ;
        float *obs = m, *sc = m + 252, *amask = m + 270, *pol = m + 270 + 214;
        orc_observation(&e, orc_turn(&e), obs, sc);
        int cnt = orc_legal_actions(&e, legal);
        for (int a = 0; a < A; ++a) amask[a] = 1e32f;
        for (int q = 0; q < cnt; ++q) amask[legal[q]] = 0.f;
        for (int a = 0; a < A; ++a) pol[a] = 0.f - amask[a];
        int act = legal[rng_next(&rs) % (uint64_t)cnt];
        int sl = (int)(rng_next(&rs) & 1);
        orc_step(&e, act, sl);
        m[MSZ - 2] = orc_reward(&e);
        m[MSZ - 1] = (float)act;
        csum += obs[(k * 7) % 27] + amask[k % A];
        j->done += 1;
    }
    free(moments);
    j->csum = csum;
    return NULL;
}

int64_t orc_selfplay(int game, int64_t n_steps, uint64_t seed, int n_threads, double* checksum) {
    if (n_threads < 1) n_threads = 1;
    pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * n_threads);
    sp_job* jobs = (sp_job*)calloc(n_threads, sizeof(sp_job));
    for (int t = 0; t < n_threads; ++t) {
        jobs[t].game = game; jobs[t].tid = t; jobs[t].n_steps = n_steps; jobs[t].seed = seed;
        pthread_create(&th[t], NULL, sp_worker, &jobs[t]);
    }
    int64_t total = 0;
    double csum = 0.0;
    for (int t = 0; t < n_threads; ++t) {
        pthread_join(th[t], NULL);
        total += jobs[t].done;
        csum += jobs[t].csum;
    }
    free(th);
    free(jobs);
    if (checksum) *checksum = csum;
    return total;
}
