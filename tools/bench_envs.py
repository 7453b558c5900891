"""Self-play rollout rate of every built-in environment (BASELINE.json configs 1-3: TicTacToe, ParallelTicTacToe,
Geister / CIGeister), 65,536 games per GPU, uniform policy, full reference-format moment output per ply, plus the
pure-write ceiling of this GPU (torch fill_) for context.  One JSON line on stdout."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from handyrl_b200.generation import BatchGenerator  # noqa: E402

N = int(os.environ.get('BENCH_ENVS_N', 65536))
BYTES = {"TicTacToe": 108 + 36 + 4 + 32, "ParallelTicTacToe": 108 + 36 + 4 + 32, "Geister": 1940 + 32,
         "CIGeister": 1940 + 32}


def timed(fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    peak = float(json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]) \
        if os.path.exists(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")) else 6548.8
    out = {"games": N, "peak_gbs": peak, "envs": {}}
    big = torch.empty(1 << 30, dtype=torch.float32, device="cuda")       # 4 GiB
    ms = timed(lambda: big.fill_(1.0), 10)
    out["write_ceiling_gbs"] = big.numel() * 4 / (ms * 1e-3) / 1e9
    del big
    for game, plies in (("TicTacToe", 256), ("ParallelTicTacToe", 256), ("Geister", 32), ("CIGeister", 32)):
        gen = BatchGenerator(game, N, seed=1, depth=512, max_episodes=1 << 20)
        env = gen.env
        W = 2 if "Geister" in game else 32
        main_w, scalar_w = env.alloc_obs(waves=W)
        waves = dict(main=main_w, scalar=scalar_w, action_mask=torch.empty(W, N, env.A, device="cuda"))
        act = torch.empty(W, N, dtype=torch.int32, device="cuda")
        ms = timed(lambda: gen.rollout(plies, waves=waves, out_action=act), 10)
        rate = N * plies / (ms * 1e-3)
        gbs = rate * BYTES[game] / 1e9
        st = gen.stats()
        out["envs"][game] = {"plies_per_launch": plies, "ms": ms, "env_steps_per_s": rate, "bytes_per_step": BYTES[game],
                             "GBps": gbs, "frac_of_peak": gbs / peak,
                             "mean_episode_len": st["env_steps"] / max(st["episodes"], 1),
                             "wins0_wins1_draws": [st["wins0"], st["wins1"], st["draws"]]}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
