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

/*
 * Tuning knobs for A/B measurements and tests; not part of the reference's interface.  Every knob defaults to its
 * environment variable (B200RL_<NAME IN CAPITALS>), read once per process; `set` = 0 returns the knob to its built-in
 * default.  Knobs: rollout_chunk (plies per rollout ticket, 0 = static schedule), target_ws_stages (2..8),
 * target_ws_row (256 | 512 bytes per row and fetch), target_ws_wide_rows (batch rows from which 512 is used),
 * target_ws_rowlane (0 = one (row, player) sequence per lane, 1 = one row per lane in the long-row kernel),
 * target_generic, target_no_dual, target_persistent (flags), emit_tma (0 = STG.128 emit path, 1 = staged bulk
 * stores; default: bulk stores for the Geister-sized tiles).  Results never depend on a knob.
 */
int b200rl_set_tuning(const char* name, int64_t value, int set);

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
 * The heads of one loss in one call.  handyrl/train.py:248-253 runs the target recursion once for the value
 * head (bootstrap = outcome, rewards None, gamma 1) and once for the return head (rewards, gamma) with the same
 * modes, lambda, shape and rhos/cs; for rows of <= 64 steps both heads run in ONE kernel launch (a head is a
 * 13-16 us kernel at B = 65,536, T = 32: two launches pay ramp-up and drain twice).  Per-head fields have the
 * meaning of the b200rl_target_scan arguments of the same name; rewards == NULL is `rewards is None`.
 * n_heads is 1 or 2.  Results are bit-identical to n_heads calls of b200rl_target_scan_fused.
 */
typedef struct b200rl_target_head {
    const float* values; int64_t v_sb, v_st, v_sp;
    const float* boot; int64_t boot_sb, boot_sp;
    const float* rewards; int64_t r_sb, r_st, r_sp;
    const float* rhos; int64_t rho_sb, rho_st, rho_sp;
    const float* cs; int64_t c_sb, c_st, c_sp;
    float* out_targets;
    float* out_adv;
    double gamma;
} b200rl_target_head;

int b200rl_target_scan_heads(int mode_targets, int mode_adv, const b200rl_target_head* heads, int n_heads,
                             int64_t B, int64_t T, int64_t P, double lmb, void* stream);

/*
 * monte_carlo (handyrl/losses.py:16-17): out_adv = returns - values with
 * broadcasting; the target is `returns` itself (aliased by the Python shim).
 */
int b200rl_target_mc(const float* values, int64_t v_sb, int64_t v_st, int64_t v_sp,
                     const float* returns, int64_t ret_sb, int64_t ret_st, int64_t ret_sp,
                     float* out_adv, int64_t B, int64_t T, int64_t P, void* stream);

/* ------------------------------------------------------------------ */
/* Part 1: batched environments — handyrl/environment.py:43-146        */
/* ------------------------------------------------------------------ */

enum {
    B200RL_GAME_TICTACTOE = 0,          /* handyrl/envs/tictactoe.py:72-172 */
    B200RL_GAME_PARALLEL_TICTACTOE = 1, /* handyrl/envs/parallel_tictactoe.py:13-58 */
    B200RL_GAME_GEISTER = 2,            /* handyrl/envs/geister.py:170-538 */
    B200RL_GAME_CIGEISTER = 3           /* handyrl/envs/ci_geister.py (observation always shows opponent types) */
};

/* Observation view selector: observation(player) of the reference. */
enum { B200RL_VIEW_NONE = -1, B200RL_VIEW_PLAYER0 = 0, B200RL_VIEW_PLAYER1 = 1, B200RL_VIEW_TURN = 2 };

/*
 * Game state lives in a caller-owned, 16-byte-aligned device buffer of
 * N * b200rl_env_state_bytes(game) bytes (struct-of-arrays planes over the N
 * games; opaque except through these calls and the layout documented in
 * DESIGN.md).  Returns -1 for an unknown game.
 */
int64_t b200rl_env_state_bytes(int game);

/* out[8] = action_length, observation floats (main tensor), scalar-observation floats,
 * longest legal list, legal-mask words, players, max plies per episode, has_reward. */
int b200rl_env_dims(int game, int64_t* out);

/* Environment.reset() for all games, or for those with slot_mask[i] != 0. */
int b200rl_env_reset(int game, void* state, int64_t N, const uint8_t* slot_mask, void* stream);

/*
 * Environment.step({player: action}) / play(action, player) for N games.
 * actions[i] < 0 means "no action" (None).  `selected` (ParallelTicTacToe only,
 * else NULL) is the player whose move lands (parallel_tictactoe.py:22).
 * out_reward (N,2) = reward() after the ply, out_terminal (N) = terminal(); either may be NULL.
 */
int b200rl_env_step(int game, void* state, int64_t N, const int32_t* actions, const int32_t* selected,
                    float* out_reward, uint8_t* out_terminal, void* stream);

/* legal_actions() as an ordered list (reference order), ids (N, LMAX) padded with -1, count (N). */
int b200rl_env_legal_list(int game, const void* state, int64_t N, int32_t* out_ids, int32_t* out_count,
                          void* stream);

/*
 * observation(player) and the legal mask of the player to move.
 *   view / views : one selector for all games, or per-game selectors (N) when views != NULL
 *   out_main     : (N, 7,6,6) Geister 'board' / (N, 3,3,3) TicTacToe, fp32, or NULL
 *   out_scalar   : (N, 18) Geister 'scalar', or NULL
 *   out_amask    : (N, A) fp32, 0 on legal and 1e32 on illegal actions (generation.py:50-51), or NULL
 *   out_bits     : (N, LW) uint32 bit mask of legal action ids, or NULL
 */
int b200rl_env_observe(int game, const void* state, int64_t N, int view, const int32_t* views, float* out_main,
                       float* out_scalar, float* out_amask, uint32_t* out_bits, void* stream);

/* outcome() -> (N, 2) fp32. */
int b200rl_env_outcome(int game, const void* state, int64_t N, float* out, void* stream);

/* (N, 4) int32: turn(), terminal(), plies played (incl. setup), winner code (0 none, 1 player0, 2 player1, 3 draw). */
int b200rl_env_info(int game, const void* state, int64_t N, int32_t* out, void* stream);

/*
 * One fused ply of the self-play loop (generation.py:31-68) for N games: apply
 * `actions` (NULL = observe only), write reward / terminal / outcome of the
 * resulting state, restart finished games when auto_reset != 0, then write the
 * turn player's observation and action mask of the state the next ply starts from.
 */
int b200rl_env_step_observe(int game, void* state, int64_t N, const int32_t* actions, const int32_t* selected,
                            int auto_reset, float* out_main, float* out_scalar, float* out_amask,
                            float* out_reward, uint8_t* out_terminal, float* out_outcome, int32_t* out_turn,
                            void* stream);

/*
 * Persistent self-play with the epoch-0 policy (RandomModel: uniform over legal
 * actions, handyrl/model.py:58-64; handyrl/train.py:438-439) for n_plies plies per
 * game with auto-reset; every game performs exactly one env-step per ply.
 *   seed, game_id_base, it0 : counter-RNG key (seed, game_id_base + slot, it0 + ply)
 *   out_main/out_scalar/out_amask/out_action : moment data of ply `it`, written to wave
 *       (it % waves) of [waves][N][...] buffers; any may be NULL.  With waves > 1 every wave base must be 16-byte
 *       aligned: N * floats-per-game a multiple of 4 (Geister: N even; TicTacToe: N a multiple of 4), else ERR_ARG
 *   traj : [N][D] slot-major ring of 32-byte compact records (8 x uint32, 32-byte aligned): record of
 *          (iteration it, slot i) at (i * D + it % D) * 32 bytes, or NULL
 *   ep_table : (E, 4) int32 ring of finished episodes {slot, steps | win << 16, end_it lo, hi}
 *   counters : uint64[8] {episodes, env steps, player0 wins, player1 wins, draws, ep_table cursor, -, -}
 *   scratch : caller-owned device buffer of >= b200rl_env_rollout_scratch_bytes(N) bytes, 16-byte aligned, for
 *          the (32-game group, 4-ply chunk) ticket counters; cleared by the launcher on `stream`, so one buffer
 *          serves one stream at a time.  The library allocates nothing: the call is graph-capturable.
 * Results depend on (game, iteration) only.  Tuning knob rollout_chunk = 0 disables chunking (A/B measurements).
 */
int b200rl_env_rollout(int game, void* state, int64_t N, int64_t n_plies, uint64_t seed, uint64_t game_id_base,
                       uint64_t it0, float* out_main, float* out_scalar, float* out_amask, int32_t* out_action,
                       int64_t waves, void* traj, int64_t D, int32_t* ep_table, int64_t E, uint64_t* counters,
                       void* scratch, int64_t scratch_bytes, void* stream);

/* Bytes of ticket scratch b200rl_env_rollout needs for N games (-1 for N < 0). */
int64_t b200rl_env_rollout_scratch_bytes(int64_t N);

/*
 * b200rl_env_rollout without any allocation on either side of the ABI.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same arguments and results as b200rl_env_rollout minus the scratch: the
 * ticket counters live in a file-scope `static __device__` array of the library, sized for at most
 * B200RL_ROLLOUT_STATIC_MAX_GAMES games (32,896 bytes, held by every GPU that loads the library until the process
 * exits).  N beyond that maximum is rejected with B200RL_ERR_ARG and a message; the call performs no cudaMalloc,
 * cudaFree or other allocating routine (it resolves the array's device address once per device with
 * cudaGetSymbolAddress and clears the words it uses with cudaMemsetAsync on `stream`).
 * All calls on a device SHARE the array: calls issued on one stream are safe back to back, calls that overlap on the
 * GPU (different streams, concurrent graph launches) race on the ticket counters and must not be made.
 */
#define B200RL_ROLLOUT_STATIC_MAX_GAMES 262144
int b200rl_env_rollout_static(int game, void* state, int64_t N, int64_t n_plies, uint64_t seed, uint64_t game_id_base,
                              uint64_t it0, float* out_main, float* out_scalar, float* out_amask, int32_t* out_action,
                              int64_t waves, void* traj, int64_t D, int32_t* ep_table, int64_t E, uint64_t* counters,
                              void* stream);

/* ------------------------------------------------------------------ */
/* Part 2: episode -> batch assembly — handyrl/train.py:33-133          */
/* ------------------------------------------------------------------ */

/*
 * One ply driven by caller-chosen actions (net-in-the-loop self-play, tape replay):
 * Environment.step() as in b200rl_env_step, plus the 32-byte trajectory record of the
 * ply (pre-move state, action, episode flags) written to ring slot it % D, and the
 * episode-table / counter bookkeeping of b200rl_env_rollout.  A slot with actions[i] < 0
 * does not move.  `both` (N,2) carries both players' chosen actions for
 * ParallelTicTacToe (make_batch needs them), else NULL.  With auto_reset a finished
 * game is restarted before the ply (generation.py:27).
 *
 * Net outputs of the ply (generation.py:40-57: moment['policy'], moment['value']) are recorded beside the
 * record when both the input and its ring are given (all four may be NULL):
 *   logits     (N, PL, A) fp32, PL = 2 for ParallelTicTacToe else 1: the acting player's raw logits
 *   values     (N, 2) fp32: moment['value'][player], 0 where a player did not observe
 *   logit_ring [N][D][PL][LS] fp32, slot-major like the record ring; LS = b200rl_env_logit_slots(game).  Only the
 *              logits of the LEGAL actions are kept, in ascending action-id order: every other entry of
 *              moment['policy'] = logits - action_mask is -1e32 in fp32 whatever the logit
 *   value_ring [N][D][2] fp32
 */
int b200rl_env_step_record(int game, void* state, int64_t N, const int32_t* actions, const int32_t* selected,
                           const int32_t* both, int auto_reset, void* traj, int64_t D, uint64_t it,
                           int32_t* ep_table, int64_t E, uint64_t* counters, uint8_t* out_terminal,
                           const float* logits, const float* values, float* logit_ring, float* value_ring,
                           void* stream);

/* Floats per (ply, acting player) of the logit ring: the longest legal list (tictactoe.py:149-151: 9;
 * geister.py:472-486: the 70 setup actions) rounded up to a multiple of 4.  -1 for an unknown game. */
int64_t b200rl_env_logit_slots(int game);

/*
 * The acting half of one self-play ply with a real policy (generation.py:48-61) for N games:
 * restart finished games, take legal_actions in the reference's list order, weights =
 * softmax(logits[legal]) (util.py:61-63, numpy's pairwise fp32 sum), draw like random.choices (cumulative
 * weights as an fp32 running sum - they are numpy float32 scalars -, first index whose cumulative weight
 * exceeds u * total), apply the action, write the trajectory record / episode table / counters and the
 * logit / value rings like b200rl_env_step_record (values, logit_ring, value_ring may be NULL).
 *   logits (N, PL, A) fp32, PL = 2 for ParallelTicTacToe (both players act, parallel_tictactoe.py:20-29) else 1
 *   u      (N, PL) doubles in [0,1) or NULL (counter RNG keyed (seed, game_id_base + slot, it))
 *   out_action (N, PL) int32, out_terminal (N) uint8; either may be NULL; traj may be NULL
 */
int b200rl_env_sample_record(int game, void* state, int64_t N, const float* logits, const double* u, uint64_t seed,
                             uint64_t game_id_base, void* traj, int64_t D, uint64_t it, int32_t* ep_table, int64_t E,
                             uint64_t* counters, int32_t* out_action, uint8_t* out_terminal, const float* values,
                             float* logit_ring, float* value_ring, void* stream);

enum { B200RL_BATCH_TURN = 0, B200RL_BATCH_ALL = 1, B200RL_BATCH_SOLO = 2 };

/*
 * make_batch(episodes, args) (handyrl/train.py:33-133) over windows of the device
 * trajectory ring.
 *   traj     : [N][D] x 32 B slot-major ring written by b200rl_env_rollout / b200rl_env_step_record
 *   values   : [N][D][2] fp32 value ring of b200rl_env_step_record, or NULL (RandomModel: 0)
 *   logits   : [N][D][PL][LS] fp32 logit ring of b200rl_env_step_record, or NULL (RandomModel: zero logits)
 *   windows  : (B,8) int32 {slot, first_it lo, first_it hi, len = end-start, start, total steps,
 *              winner code, solo player}; first_it is the ring iteration of ply `start`
 *   T        : args['forward_steps']; rows t >= len are padded exactly like train.py:92-104
 *   mode     : TURN = turn_based_training and not observation (train.py:64-68);
 *              ALL  = turn_based_training and observation (both players, train.py:69-73);
 *              SOLO = not turn_based_training (one player per episode, train.py:57-58)
 *   obs_all  : args['observation'] (every player observes every ply)
 *   ret_table: ret_table[m] = discounted return of a ply with m plies left (generation.py:73-77,
 *              accumulated in double by the caller, rounded to fp32)
 * Outputs are the tensors of the returned dict, contiguous, shapes as in train.py:109-133
 * with Po = 2 for ALL else 1 and Pv = 1 for SOLO else 2:
 *   obs_main (B,T,Po,...) obs_scalar (B,T,Po,18)|NULL policy/amask (B,T,Po,A) action (B,T,Po,1) int64
 *   value/reward/ret/tmask/omask (B,T,Pv,1) outcome (B,1,Pv,1) emask (B,T,1,1) progress (B,T,1)
 */
int b200rl_make_batch(int game, const void* traj, int64_t D, int64_t N, const float* values, const float* logits,
                      const int32_t* windows, int64_t B, int64_t T, int mode, int obs_all, const float* ret_table,
                      int64_t ret_len, float* obs_main, float* obs_scalar, float* policy, float* amask,
                      int64_t* action, float* value, float* reward, float* ret, float* outcome, float* emask,
                      float* tmask, float* omask, float* progress, void* stream);

/*
 * Batcher.select_episode (handyrl/train.py:284-302) for B windows: recency-biased rejection sampling
 * over the last min(finished, E, maximum_episodes) rows of the finished-episode table (index 0 =
 * oldest), then a uniform window start.  Rows whose first ply left the trajectory ring (older than D
 * iterations before it_now) are rejected.  Draws: counter RNG keyed (seed, b, draw0 + 2*try + j), two 32-bit draws per 64-bit hash.
 *   out_windows (B,8) int32 in the layout b200rl_make_batch consumes (solo player = 1 random bit).
 * Reads counters[5] on the device: no host synchronisation between rollout and window selection.
 */
int b200rl_select_windows(const int32_t* ep_table, int64_t E, const uint64_t* counters, int64_t maximum_episodes,
                          int64_t forward_steps, uint64_t seed, uint64_t draw0, uint64_t it_now, int64_t D,
                          int32_t* out_windows, int64_t B, void* stream);

/*
 * The same selection on the HOST (replay index kept in host memory, like the reference's Learner): identical
 * rule, RNG and output as b200rl_select_windows, so the rows are bit-identical for equal arguments.
 * ep_table_host / out_windows_host are host pointers (pin them for the upload); `cursor` = episodes finished so
 * far (counters[5]); n_threads host threads share the B windows.
 */
int b200rl_select_windows_host(const int32_t* ep_table_host, int64_t E, uint64_t cursor, int64_t maximum_episodes,
                               int64_t forward_steps, uint64_t seed, uint64_t draw0, uint64_t it_now, int64_t D,
                               int32_t* out_windows_host, int64_t B, int n_threads);

/*
 * Padding + stacking step of make_batch (train.py:92-122) for episodes that were decoded on the
 * host (reference wire format): `src` holds the windows of all B episodes back to back, rows of
 * `row_words` 32-bit words, episode b owning rows offsets[b]..offsets[b+1]; out is (B, T, row)
 * with rows t >= length filled from fill_rows[b] (B,row) when given, else with fill_bits.
 */
int b200rl_pad_stack(const void* src, const int64_t* offsets, const void* fill_rows, uint32_t fill_bits, void* out,
                     int64_t B, int64_t T, int64_t row_words, void* stream);

/* ------------------------------------------------------------------ */
/* Loss pre-processing around the recursion — handyrl/train.py:224-239  */
/* ------------------------------------------------------------------ */

/*
 * Per row r of A logits (rows = B*T*Po, contiguous (rows, A) fp32):
 *   out_logp_target[r] = log_softmax(policy_target[r])[action[r]] * emask[r]          (train.py:226)
 *   ratio = exp(out_logp_target[r] - log_softmax(policy_behaviour[r])[action[r]] * emask[r])  (train.py:225-229)
 *   out_rho[r] = clamp(ratio, 0, clip_rho), out_c[r] = clamp(ratio, 0, clip_c)         (train.py:230-231)
 * action is int64 (rows), emask fp32 (rows) (episode_mask broadcast over players by the caller).
 */
int b200rl_policy_terms(const float* policy_behaviour, const float* policy_target, const int64_t* action,
                        const float* emask, int64_t rows, int64_t A, double clip_rho, double clip_c,
                        float* out_logp_target, float* out_rho, float* out_c, void* stream);

/* Backward of out_logp_target w.r.t. policy_target: grad[r][j] = grad_logp[r] * emask[r] * ([j == a] - softmax_j). */
int b200rl_policy_terms_backward(const float* policy_target, const int64_t* action, const float* emask,
                                 const float* grad_logp, int64_t rows, int64_t A, float* out_grad_policy, void* stream);

/*
 * Value-head input of compute_target (train.py:236-239) on contiguous (B,T,P) fp32:
 *   symmetric && P == 2: v <- (v + (-swap_players(v))) / (omask.sum(players) + 1e-8)
 *   out = v * emask + outcome * (1 - emask)        emask (B,T), outcome (B,P)
 * Same op order as the torch expression (bit-identical).
 */
int b200rl_value_blend(const float* values, const float* omask, const float* emask, const float* outcome, int64_t B,
                       int64_t T, int64_t P, int symmetric, float* out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200RL_H */
