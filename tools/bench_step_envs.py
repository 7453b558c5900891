"""The whole hot-path step of bench.py (rollout -> device select_episode -> make_batch -> targets) for EVERY built-in
environment, phase by phase: bench.py times it for Geister (BASELINE configs[1]); this gives configs[0] / [2] (TicTacToe,
ParallelTicTacToe) and CIGeister the same breakdown.  Inputs larger than L2 (the batch alone is >= 0.5 GB), CUDA events on
the launching stream, fraction of the measured copy peak from MEASURED_PEAKS.json.  One JSON line.  The two short phases
(select: ~15 us, targets: ~20 us) are replayed from a captured CUDA graph, because Python launches them slower than they
run; replaying them back to back on the same buffers leaves part of their inputs in L2, so their fractions here are an
upper bound - bench.py's targets figure, taken with the rest of the step in between, is the one to quote.

    python tools/bench_step_envs.py            # 65,536 games x 32 plies per step; T = 32 (Geister) / 8 (TicTacToe:
                                               # episodes last <= 9 plies), B = games x 32 / T
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from handyrl_b200.generation import BatchGenerator  # noqa: E402
from handyrl_b200.losses import compute_target_heads  # noqa: E402
from handyrl_b200.train import make_batch_windows  # noqa: E402

N = int(os.environ.get('BENCH_ENVS_N', 65536))
ITERS = int(os.environ.get('BENCH_ITERS', 20))


def peak():
    try:
        return float(json.load(open(os.path.join(os.path.dirname(__file__), '..', 'MEASURED_PEAKS.json')))['hbm_gbs'])
    except Exception:
        return 6552.0


def timed(fn, iters, graph=False):
    """ms per call; graph=True replays the `iters` calls from one captured CUDA graph (kernels of a few tens of
    microseconds are otherwise timed at the rate Python can launch them)"""
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    if graph:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(iters):
                fn()
        g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / iters
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def one(game, T, plies, B):
    geister = 'Geister' in game
    args = {'turn_based_training': True, 'observation': False, 'gamma': 0.8, 'forward_steps': T, 'compress_steps': 4,
            'maximum_episodes': 1 << 20, 'lambda': 0.7}
    gen = BatchGenerator(game, N, seed=1, depth=512, max_episodes=1 << 21)
    env = gen.env
    W = 2 if geister else 32
    main_w, scalar_w = env.alloc_obs(waves=W)
    waves = dict(main=main_w, scalar=scalar_w, action_mask=torch.empty(W, N, env.A, device='cuda'))
    act = torch.empty(W, N, dtype=torch.int32, device='cuda')
    for _ in range(max(1, 512 // plies)):           # fill the ring: windows need finished episodes
        gen.rollout(plies, waves=waves, out_action=act)
    t_roll = timed(lambda: gen.rollout(plies, waves=waves, out_action=act), ITERS)
    win = gen.select_windows(B, args)
    wout = torch.empty(B, 8, dtype=torch.int32, device='cuda')
    t_sel = timed(lambda: gen.select_windows(B, args, out=wout), ITERS, graph=True)
    out = make_batch_windows(gen.store, win, args)
    t_mb = timed(lambda: make_batch_windows(gen.store, win, args, out=out), ITERS)
    # bytes a cell writes: every output tensor of the batch + the 32-byte record it reads
    cell_bytes = 32 + sum(t.numel() * t.element_size() for k, t in _flat(out).items() if k != 'outcome') / (B * T)
    # targets as compute_loss asks for them (train.py:245-253): UPGO value / VTRACE policy on the value head, and on the
    # return head where the game has rewards
    v = torch.tanh(torch.randn(B, T, 2, 1, device='cuda'))
    rho = torch.rand(B, T, 1, 1, device='cuda')
    heads = [(v, out['outcome'], None, 1.0)]
    tb = 36 / 2
    if env.has_reward:
        heads.append((v.clone(), out['return'], out['reward'], args['gamma']))
        tb = 36
    t_tg = timed(lambda: compute_target_heads('UPGO', 'VTRACE', heads, 0.7, rho, rho), ITERS, graph=True)
    pk = peak()
    steps = N * plies
    roll_b = {'Geister': 1972, 'CIGeister': 1972}.get(game, 180)
    live = float(out['episode_mask'].mean())
    r = {
        'games': N, 'plies': plies, 'B': B, 'T': T, 'live_cell_fraction': round(live, 4),
        'rollout_ms': t_roll, 'select_ms': t_sel, 'make_batch_ms': t_mb, 'targets_ms': t_tg,
        'rollout_frac': steps * roll_b / (t_roll * 1e-3) / 1e9 / pk,
        'make_batch_bytes_per_cell': cell_bytes,
        'make_batch_frac': B * T * cell_bytes / ((t_mb + t_sel) * 1e-3) / 1e9 / pk,
        'targets_bytes_per_element': tb,
        'targets_frac': B * T * 2 * tb / (t_tg * 1e-3) / 1e9 / pk,
        'step_ms': t_roll + t_sel + t_mb + t_tg,
        'env_steps_per_s': steps / ((t_roll + t_sel + t_mb + t_tg) * 1e-3),
    }
    return r


def _flat(out):
    d = {}
    for k, t in out.items():
        if isinstance(t, dict):
            for kk, tt in t.items():
                d[k + '.' + kk] = tt
        elif t is not None:
            d[k] = t
    return d


if __name__ == '__main__':
    res = {'peak_gbs': peak(), 'envs': {}}
    for game in ('TicTacToe', 'ParallelTicTacToe', 'Geister', 'CIGeister'):
        geister = 'Geister' in game
        T = 32 if geister else 8
        plies = 32
        B = N * plies // T                                      # every generated ply is batched once: 2 M cells
        res['envs'][game] = one(game, T, plies, B)
        torch.cuda.empty_cache()
    print(json.dumps(res))
