/*
 * CPU port of the WHOLE hot path for Geister — TEST INFRASTRUCTURE / timed CPU baseline only
 * (see oracle/__init__.py; nothing under handyrl_b200/ links or calls this).
 *
 * One "step" per thread, mirroring what bench.py times on the GPU:
 *   1. generation  handyrl/generation.py:31-88 with RandomModel: G games advance `plies` plies each; every ply
 *                  stores the moment (observation, policy, action_mask, action, value, reward) in the game's
 *                  episode buffer; a finished game computes its discounted returns (generation.py:73-77) and
 *                  joins the thread's replay list (no bz2/pickle: the compression of the reference is left
 *                  out, which favours this baseline).
 *   2. batching    Batcher.select_episode (train.py:284-302) + make_batch (train.py:33-133), turn-based mode:
 *                  B = G*plies/T windows gathered into (B,T,...) arrays with the reference's padding rules.
 *   3. targets     compute_target x2 per head as compute_loss calls it (train.py:245-253) with
 *                  value_target UPGO / policy_target VTRACE on the value head (bootstrap = outcome) and the
 *                  return head (returns/rewards), losses.py:31-58, over (B,T,2).
 * Game rules come from games_oracle.c (same translation unit via #include).
 */
#include "games_oracle.c"

#include <math.h>
#include <pthread.h>
#include <time.h>

static double hp_now(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

#define HP_A 214
#define HP_OBS 270 /* 252 board + 18 scalar */
#define HP_MAXLEN 204

typedef struct {
    float obs[HP_OBS];
    float policy[HP_A];
    float amask[HP_A];
    float value[2], reward[2], ret[2];
    int action, turn;
} hp_moment;

typedef struct {
    int steps;
    float outcome[2];
    hp_moment* m; /* steps moments */
} hp_episode;

typedef struct {
    int game, tid, G, plies, T, replay_cap;
    uint64_t seed;
    int64_t env_steps, batch_cells;
    double csum, sec[3]; /* seconds spent in generation / batching / targets */
    double gamma, lmb;   /* args['gamma'], args['lambda'] of the training configuration */
    /* persistent state */
    orc_env* envs;
    hp_moment** cur;   /* per game: moments of the running episode */
    int* cur_len;
    hp_episode* replay; /* ring of finished episodes */
    int64_t n_finished;
    uint64_t rs;
    /* batch + target buffers */
    float *b_obs, *b_policy, *b_amask, *b_value, *b_reward, *b_ret, *b_outcome, *b_emask, *b_tmask, *b_omask, *b_prog;
    int64_t* b_action;
    float *t_v, *t_rho, *t_tv, *t_adv;
    int B;
} hp_thread;

static double hp_rand01(uint64_t* s) { return (double)(rng_next(s) >> 11) * (1.0 / 9007199254740992.0); }

static void hp_finish_episode(hp_thread* t, int g) {
    orc_env* e = &t->envs[g];
    int n = t->cur_len[g];
    hp_moment* ms = t->cur[g];
    for (int p = 0; p < 2; ++p) { /* generation.py:73-77, accumulated in double like the Python float */
        double ret = 0.0;
        for (int i = n - 1; i >= 0; --i) {
            ret = (double)ms[i].reward[p] + t->gamma * ret;
            ms[i].ret[p] = (float)ret;
        }
    }
    hp_episode* ep = &t->replay[t->n_finished % t->replay_cap];
    /* hand the buffer over instead of copying it; the evicted episode's buffer becomes the new scratch */
    hp_moment* spare = ep->m ? ep->m : (hp_moment*)malloc(sizeof(hp_moment) * HP_MAXLEN);
    ep->m = ms;
    t->cur[g] = spare;
    ep->steps = n;
    orc_outcome(e, ep->outcome);
    t->n_finished += 1;
    t->cur_len[g] = 0;
    orc_reset(e, t->game);
}

static void hp_generate(hp_thread* t) {
    int legal[80];
    for (int g = 0; g < t->G; ++g) {
        orc_env* e = &t->envs[g];
        for (int k = 0; k < t->plies; ++k) {
            if (orc_terminal(e)) hp_finish_episode(t, g);
            hp_moment* m = &t->cur[g][t->cur_len[g]++];
            int tp = orc_turn(e);
            m->turn = tp;
            orc_observation(e, tp, m->obs, m->obs + 252);
            int cnt = orc_legal_actions(e, legal);
            for (int a = 0; a < HP_A; ++a) m->amask[a] = 1e32f;
            for (int q = 0; q < cnt; ++q) m->amask[legal[q]] = 0.f;
            for (int a = 0; a < HP_A; ++a) m->policy[a] = 0.f - m->amask[a]; /* RandomModel: zero logits */
            m->action = legal[rng_next(&t->rs) % (uint64_t)cnt];
            m->value[0] = m->value[1] = 0.f;
            orc_step(e, m->action, 0);
            m->reward[0] = m->reward[1] = orc_reward(e);
            t->env_steps += 1;
        }
    }
}

/* Batcher.select_episode (train.py:284-302) + make_batch in turn-based mode (train.py:33-133) */
static void hp_make_batch(hp_thread* t) {
    const int T = t->T, B = t->B;
    int64_t L = t->n_finished < t->replay_cap ? t->n_finished : t->replay_cap;
    for (int b = 0; b < B; ++b) {
        float* obs = t->b_obs + (size_t)b * T * HP_OBS;
        float* pol = t->b_policy + (size_t)b * T * HP_A;
        float* am = t->b_amask + (size_t)b * T * HP_A;
        if (L == 0) continue;
        int64_t idx;
        for (;;) { /* train.py:285-289 */
            idx = (int64_t)(hp_rand01(&t->rs) * (double)L);
            double accept = 1.0 - (double)(L - 1 - idx) / (double)t->replay_cap;
            if (hp_rand01(&t->rs) < accept) break;
        }
        /* index 0 = oldest of the last L finished episodes */
        const hp_episode* ep = &t->replay[(t->n_finished - L + idx) % t->replay_cap];
        int cand = 1 + (ep->steps - T > 0 ? ep->steps - T : 0);
        int st = (int)(hp_rand01(&t->rs) * cand);
        int ed = st + T < ep->steps ? st + T : ep->steps;
        t->b_outcome[b * 2] = ep->outcome[0];
        t->b_outcome[b * 2 + 1] = ep->outcome[1];
        for (int k = 0; k < T; ++k) {
            size_t c = (size_t)b * T + k;
            if (st + k < ed) {
                const hp_moment* m = &ep->m[st + k];
                memcpy(obs + (size_t)k * HP_OBS, m->obs, sizeof(m->obs));
                memcpy(pol + (size_t)k * HP_A, m->policy, sizeof(m->policy));
                memcpy(am + (size_t)k * HP_A, m->amask, sizeof(m->amask));
                t->b_action[c] = m->action;
                for (int p = 0; p < 2; ++p) {
                    t->b_value[c * 2 + p] = p == m->turn ? m->value[p] : 0.f;
                    t->b_reward[c * 2 + p] = m->reward[p];
                    t->b_ret[c * 2 + p] = m->ret[p];
                    t->b_tmask[c * 2 + p] = p == m->turn ? 1.f : 0.f;
                    t->b_omask[c * 2 + p] = p == m->turn ? 1.f : 0.f;
                }
                t->b_emask[c] = 1.f;
                t->b_prog[c] = (float)(st + k) / (float)ep->steps;
            } else { /* train.py:92-104 */
                memset(obs + (size_t)k * HP_OBS, 0, sizeof(float) * HP_OBS);
                memset(pol + (size_t)k * HP_A, 0, sizeof(float) * HP_A);
                for (int a = 0; a < HP_A; ++a) am[(size_t)k * HP_A + a] = 1e32f;
                t->b_action[c] = 0;
                for (int p = 0; p < 2; ++p) {
                    t->b_value[c * 2 + p] = ep->outcome[p];
                    t->b_reward[c * 2 + p] = 0.f;
                    t->b_ret[c * 2 + p] = 0.f;
                    t->b_tmask[c * 2 + p] = 0.f;
                    t->b_omask[c * 2 + p] = 0.f;
                }
                t->b_emask[c] = 0.f;
                t->b_prog[c] = 1.f;
            }
            t->batch_cells += 1;
        }
    }
}

/* one head: UPGO targets (losses.py:31-40) + V-Trace advantages (losses.py:43-58) over (B,T,2) */
static void hp_targets_head(hp_thread* t, const float* v, const float* boot, int boot_bt, const float* rew,
                            float lmb, float gamma, float* out_tv, float* out_adv) {
    const int T = t->T, B = t->B;
    const float oml = (float)(1.0 - (double)lmb), gl = (float)((double)gamma * (double)lmb);
    for (int b = 0; b < B; ++b)
        for (int p = 0; p < 2; ++p) {
            const float R = boot_bt ? boot[((size_t)b * T + (T - 1)) * 2 + p] : boot[b * 2 + p];
            float G = R, acc = 0.f, v1 = 0.f, vs1 = 0.f;
            for (int k = T - 1; k >= 0; --k) {
                size_t i = ((size_t)b * T + k) * 2 + p;
                float vv = v[i], r = rew ? rew[i] : 0.f, rho = t->t_rho[(size_t)b * T + k];
                int last = k == T - 1;
                if (!last) {
                    float mix = oml * v1 + lmb * G;
                    G = r + gamma * (v1 > mix ? v1 : mix);
                }
                float vt1 = last ? R : v1, vst1 = last ? R : vs1;
                float delta = rho * ((r + gamma * vt1) - vv);
                acc = last ? delta : delta + (gl * rho) * acc;
                float vs = acc + vv;
                out_tv[i] = G;
                out_adv[i] = (r + gamma * vst1) - vv;
                v1 = vv;
                vs1 = vs;
            }
        }
}

static void* hp_worker(void* arg) {
    hp_thread* t = (hp_thread*)arg;
    double t0 = hp_now();
    hp_generate(t);
    double t1 = hp_now();
    hp_make_batch(t);
    double t2 = hp_now();
    const size_t n2 = (size_t)t->B * t->T * 2;
    hp_targets_head(t, t->t_v, t->b_outcome, 0, NULL, (float)t->lmb, 1.0f, t->t_tv, t->t_adv); /* train.py:245 */
    hp_targets_head(t, t->t_v + n2, t->b_ret, 1, t->b_reward, (float)t->lmb, (float)t->gamma, t->t_tv + n2, t->t_adv + n2);
    double t3 = hp_now();
    t->sec[0] += t1 - t0; t->sec[1] += t2 - t1; t->sec[2] += t3 - t2;
    t->csum += t->t_tv[0] + t->t_adv[(size_t)t->B * t->T] + t->b_obs[7] + t->b_amask[3];
    return NULL;
}

/* ---- handle-based driver: state persists across steps ---- */
typedef struct {
    int n_threads;
    hp_thread* th;
} hp_ctx;

void* orc_hotpath_create(int G, int plies, int T, int replay_cap, uint64_t seed, int n_threads, double gamma,
                         double lmb) {
    hp_ctx* c = (hp_ctx*)calloc(1, sizeof(hp_ctx));
    c->n_threads = n_threads;
    c->th = (hp_thread*)calloc((size_t)n_threads, sizeof(hp_thread));
    for (int i = 0; i < n_threads; ++i) {
        hp_thread* t = &c->th[i];
        t->gamma = gamma; t->lmb = lmb;
        t->game = G_GEISTER; t->tid = i; t->G = G; t->plies = plies; t->T = T; t->replay_cap = replay_cap;
        t->rs = (seed + 1) * 0x9E3779B97F4A7C15ull + (uint64_t)i * 0xD1B54A32D192ED03ull + 1;
        t->envs = (orc_env*)calloc((size_t)G, sizeof(orc_env));
        t->cur = (hp_moment**)calloc((size_t)G, sizeof(hp_moment*));
        t->cur_len = (int*)calloc((size_t)G, sizeof(int));
        for (int g = 0; g < G; ++g) {
            orc_reset(&t->envs[g], G_GEISTER);
            t->cur[g] = (hp_moment*)malloc(sizeof(hp_moment) * HP_MAXLEN);
        }
        t->replay = (hp_episode*)calloc((size_t)replay_cap, sizeof(hp_episode));
        int B = t->B = G * plies / T;
        size_t n = (size_t)B * T;
        t->b_obs = (float*)malloc(sizeof(float) * n * HP_OBS);
        t->b_policy = (float*)malloc(sizeof(float) * n * HP_A);
        t->b_amask = (float*)malloc(sizeof(float) * n * HP_A);
        t->b_action = (int64_t*)malloc(sizeof(int64_t) * n);
        t->b_value = (float*)malloc(sizeof(float) * n * 2);
        t->b_reward = (float*)malloc(sizeof(float) * n * 2);
        t->b_ret = (float*)malloc(sizeof(float) * n * 2);
        t->b_tmask = (float*)malloc(sizeof(float) * n * 2);
        t->b_omask = (float*)malloc(sizeof(float) * n * 2);
        t->b_emask = (float*)malloc(sizeof(float) * n);
        t->b_prog = (float*)malloc(sizeof(float) * n);
        t->b_outcome = (float*)calloc((size_t)B * 2, sizeof(float));
        t->t_v = (float*)malloc(sizeof(float) * n * 4); /* value-head and return-head net outputs (synthetic) */
        t->t_rho = (float*)malloc(sizeof(float) * n);
        t->t_tv = (float*)malloc(sizeof(float) * n * 4); /* [head][B][T][2] */
        t->t_adv = (float*)malloc(sizeof(float) * n * 4);
        memset(t->b_obs, 0, sizeof(float) * n * HP_OBS);
        for (size_t j = 0; j < n * 4; ++j) t->t_v[j] = (float)tanh(2.0 * hp_rand01(&t->rs) - 1.0);
        for (size_t j = 0; j < n; ++j) {
            double x = exp(hp_rand01(&t->rs) - 0.5);
            t->t_rho[j] = (float)(x > 1.0 ? 1.0 : x);
        }
    }
    return c;
}

/* cumulative seconds per phase, summed over threads */
void orc_hotpath_phase_seconds(void* ctx, double* out) {
    hp_ctx* c = (hp_ctx*)ctx;
    out[0] = out[1] = out[2] = 0.0;
    for (int i = 0; i < c->n_threads; ++i)
        for (int k = 0; k < 3; ++k) out[k] += c->th[i].sec[k];
}

/* one step on every thread; returns env-steps done in this step; out[0] = batch cells, out[1] = finished episodes */
int64_t orc_hotpath_step(void* ctx, int64_t* out) {
    hp_ctx* c = (hp_ctx*)ctx;
    pthread_t* th = (pthread_t*)malloc(sizeof(pthread_t) * (size_t)c->n_threads);
    int64_t before = 0, after = 0, cells = 0, fin = 0;
    for (int i = 0; i < c->n_threads; ++i) {
        before += c->th[i].env_steps;
        c->th[i].batch_cells = 0;
        pthread_create(&th[i], NULL, hp_worker, &c->th[i]);
    }
    for (int i = 0; i < c->n_threads; ++i) {
        pthread_join(th[i], NULL);
        after += c->th[i].env_steps;
        cells += c->th[i].batch_cells;
        fin += c->th[i].n_finished;
    }
    free(th);
    if (out) { out[0] = cells; out[1] = fin; }
    return after - before;
}

/* test access: thread 0's batch / target arrays */
void orc_hotpath_peek(void* ctx, float** tv, float** adv, float** v, float** rho, float** outcome, float** ret,
                      float** reward, int* B) {
    hp_thread* t = &((hp_ctx*)ctx)->th[0];
    *tv = t->t_tv; *adv = t->t_adv; *v = t->t_v; *rho = t->t_rho; *outcome = t->b_outcome; *ret = t->b_ret;
    *reward = t->b_reward; *B = t->B;
}
