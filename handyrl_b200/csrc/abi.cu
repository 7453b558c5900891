// Error reporting and version for the C-ABI (include/b200rl.h).
#include "common.cuh"
#include "../../include/b200rl.h"

static thread_local char g_err[512] = "";

void b200rl_set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

extern "C" const char* b200rl_last_error(void) { return g_err; }
extern "C" int b200rl_version(void) { return 1; }

// Tuning knobs for A/B measurements and tests.  Each knob's default comes from its environment variable, read ONCE
// per process (first use); b200rl_set_tuning overrides it at run time, so no launcher calls getenv on the hot path.
#include <atomic>
#include <cstdlib>
#include <cstring>

namespace {
struct Knob {
    const char* name;
    const char* env;
    std::atomic<long long> value;
    std::atomic<int> state;   // 0 = unread, 1 = unset (use the built-in default), 2 = set
};
Knob g_knobs[B200RL_TUNE_COUNT] = {
    {"rollout_chunk", "B200RL_ROLLOUT_CHUNK", {0}, {0}},
    {"target_ws_stages", "B200RL_TARGET_WS_STAGES", {0}, {0}},
    {"target_ws_row", "B200RL_TARGET_WS_ROW", {0}, {0}},
    {"target_ws_wide_rows", "B200RL_TARGET_WS_WIDE_ROWS", {0}, {0}},
    {"target_generic", "B200RL_TARGET_GENERIC", {0}, {0}},
    {"target_no_dual", "B200RL_TARGET_NO_DUAL", {0}, {0}},
    {"target_persistent", "B200RL_TARGET_PERSISTENT", {0}, {0}},
    {"emit_tma", "B200RL_EMIT_TMA", {0}, {0}},
    {"target_ws_rowlane", "B200RL_TARGET_WS_ROWLANE", {0}, {0}},
};
}  // namespace

bool b200rl_tuning(int knob, long long* out) {
    Knob& k = g_knobs[knob];
    int st = k.state.load(std::memory_order_acquire);
    if (st == 0) {
        const char* e = getenv(k.env);
        if (e) k.value.store(e[0] ? atoll(e) : 1, std::memory_order_relaxed);
        st = e ? 2 : 1;
        k.state.store(st, std::memory_order_release);
    }
    if (st == 2 && out) *out = k.value.load(std::memory_order_relaxed);
    return st == 2;
}

extern "C" int b200rl_set_tuning(const char* name, int64_t value, int set) {
    B200RL_REQUIRE(name != nullptr, "set_tuning: null name");
    for (int i = 0; i < B200RL_TUNE_COUNT; ++i) {
        if (strcmp(name, g_knobs[i].name) == 0) {
            g_knobs[i].value.store((long long)value, std::memory_order_relaxed);
            g_knobs[i].state.store(set ? 2 : 1, std::memory_order_release);
            return B200RL_OK;
        }
    }
    b200rl_set_error("set_tuning: unknown knob '%s'", name);
    return B200RL_ERR_ARG;
}
