"""Python restatement of `Batcher.select_episode` (handyrl/train.py:284-302) driven by the
device's counter RNG — TEST INFRASTRUCTURE (see oracle/__init__.py).

The reference draws from Python's global Mersenne Twister, so parity with it is defined on the
RULE (recency-biased rejection sampling + uniform window start), not on the stream.  This
restatement feeds the same rule with the counter RNG of handyrl_b200/csrc/common.cuh
(b200rl_rand32) so that the CUDA kernel can be checked bit-exactly.
"""
import numpy as np

M64 = (1 << 64) - 1


def mix64(z):
    z = (z + 0x9E3779B97F4A7C15) & M64
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M64
    return z ^ (z >> 31)


def rand64(seed, stream, counter):
    return mix64(seed ^ mix64((stream * 0xD6E8FEB86659FD93 + counter) & M64))


def rand32(seed, stream, counter):
    return rand64(seed, stream, counter) >> 32


def select_windows(ep_table, cursor, maximum_episodes, forward_steps, seed, draw0, it_now, D, B):
    """ep_table (E,4) int32 ring, cursor = episodes finished so far.  Returns (B,8) int32."""
    E = ep_table.shape[0]
    L = min(cursor, E, maximum_episodes)    # self.episodes is trimmed to maximum_episodes (train.py:502-503)
    out = np.zeros((B, 8), dtype=np.int64)
    for b in range(B):
        if L <= 0:
            out[b] = (0, 0, 0, 0, 0, 1, 3, 0)
            continue
        n_draw = min(L, maximum_episodes)
        for tr in range(64):
            c = draw0 + 2 * tr                       # one 64-bit hash per try: index draw | acceptance draw
            x, y = rand64(seed, b, c), rand64(seed, b, c + 1)
            r0, r1, r2, r3 = x >> 32, x & 0xFFFFFFFF, y >> 32, y & 0xFFFFFFFF
            ep_idx = (r0 * n_draw) >> 32                                   # train.py:286
            accept = 1.0 - (L - 1 - ep_idx) / maximum_episodes              # train.py:287
            last_try = tr == 63
            if last_try:
                ep_idx = L - 1
            e = ep_table[(cursor - L + ep_idx) % E]
            slot, steps, win = int(e[0]), int(e[1]) & 0xFFFF, (int(e[1]) >> 16) & 3
            end_it = int(np.uint32(e[2])) | (int(np.uint32(e[3])) << 32)
            alive = it_now - (end_it + 1 - steps) <= D
            if last_try or (alive and r1 * (1.0 / 4294967296.0) < accept):  # train.py:288
                break
        cand = 1 + max(0, steps - forward_steps)                           # train.py:291
        st = (r2 * cand) >> 32                                             # train.py:292
        ed = min(st + forward_steps, steps)                                # train.py:293
        first = end_it + 1 - steps + st
        out[b] = (slot, first & 0xFFFFFFFF, first >> 32, ed - st, st, steps, win, r3 >> 31)
    return (out & 0xFFFFFFFF).astype(np.uint32).view(np.int32)
