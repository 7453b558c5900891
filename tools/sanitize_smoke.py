"""Every kernel of libb200rl.so at small, ragged sizes in one process: a fault smoke (compute-sanitizer is closed on the
GPU pool this was developed on - "find a bad access with bounds checks and asserts of your own, small cases, and a
comparison with the CPU reference" - so this runs plain, with CUDA_LAUNCH_BLOCKING=1 to locate a fault; where the tool is
available: compute-sanitizer --tool memcheck | racecheck | synccheck python tools/sanitize_smoke.py).

Covers: reset / step / observe / legal_list / outcome / info / step_observe, the persistent rollout (ticketed and static
schedule, both emit paths, caller scratch and the library's static scratch), step_record / sample_record with the value
and logit rings, select_windows (device and host), make_batch in its three modes with both emit paths and with recorded
logits, pad_stack through reference-wire episodes, the target kernels (short rows, two heads in one launch, the
warp-specialised long-row kernel in both lane mappings, the generic strided kernel, Monte Carlo), and the loss terms.
Sizes are ragged on purpose (games not a multiple of 32, rows not a multiple of the tile)."""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from handyrl_b200 import _lib  # noqa: E402
from handyrl_b200.engine import BatchedEnv  # noqa: E402
from handyrl_b200.generation import BatchGenerator, export_episode  # noqa: E402
from handyrl_b200.losses import compute_target, compute_target_heads, compute_target_pair  # noqa: E402
from handyrl_b200.train import make_batch, make_batch_windows  # noqa: E402

dev = torch.device("cuda")
ARGS = {'turn_based_training': True, 'observation': False, 'gamma': 0.8, 'forward_steps': 8, 'compress_steps': 4,
        'maximum_episodes': 1000}


def engine(game, n):
    env = BatchedEnv(game, n)
    for _ in range(3):
        ids, cnt = env.legal_list()
        act = ids[:, 0].contiguous()
        env.step(act, torch.zeros(n, dtype=torch.int32, device=dev), want_reward=True, want_terminal=True)
        env.observe(view=2, amask=True, bits=True)
        env.observe(view=-1)
        env.outcome()
        env.info()
    out = env.step_observe(None)
    env.step_observe(out['action_mask'].argmin(1).int(), auto_reset=True, out=out)


def rollouts(game, n, plies):
    for knob in ((), (("emit_tma", 1),), (("rollout_chunk", 0),)):
        for k, v in knob:
            _lib.set_tuning(k, v)
        gen = BatchGenerator(game, n, seed=3, depth=64, max_episodes=4096)
        main, scalar = gen.env.alloc_obs(waves=2)
        waves = dict(main=main, scalar=scalar, action_mask=torch.empty(2, n, gen.env.A, device=dev))
        act = torch.empty(2, n, dtype=torch.int32, device=dev)
        gen.rollout(plies, waves=waves, out_action=act)
        gen.rollout(plies)
        for k, _ in knob:
            _lib.set_tuning(k, None)
    # the library's static scratch
    env = BatchedEnv(game, n)
    ctr = torch.zeros(8, dtype=torch.int64, device=dev)
    p = lambda t: ctypes.c_void_p(0 if t is None else t.data_ptr())
    for it0 in (0, plies):
        _lib.call("b200rl_env_rollout_static", env.game, p(env.state), n, plies, 5, 0, it0, None, None, None, None, 0, None, 0,
                  None, 0, p(ctr), _lib.stream_ptr())
    return gen


def batches(game, n, plies):
    gen = BatchGenerator(game, n, seed=4, depth=256, max_episodes=4096, record_values=True, record_logits=True)
    PL = gen.store.PL
    g = torch.Generator(device=dev).manual_seed(1)
    for _ in range(plies):
        gen.observe_turn()
        logits = torch.randn(n, PL, gen.env.A, device=dev, generator=g)
        gen.act(logits if PL == 2 else logits[:, 0].contiguous(), values=torch.rand(n, 2, device=dev, generator=g))
    for mode in (dict(), dict(observation=True), dict(turn_based_training=False)):
        a = dict(ARGS, **mode)
        w = gen.select_windows(37, a)
        for knob in (0, 1):
            _lib.set_tuning("emit_tma", knob)
            make_batch_windows(gen.store, w, a)
            _lib.set_tuning("emit_tma", None)
    eps = gen.episodes()[:6]
    if eps:
        ref = [export_episode(e, ARGS) for e in eps]            # device -> reference wire format
        wins = [{'args': {}, 'outcome': e['outcome'], 'moment': e['moment'], 'base': 0, 'start': 0,
                 'end': min(8, e['steps']), 'total': e['steps']} for e in ref]
        make_batch(wins, ARGS)                                   # host decode + b200rl_pad_stack


def targets():
    g = torch.Generator(device=dev).manual_seed(2)
    for B, T, P in ((77, 8, 2), (300, 32, 2), (45, 200, 2), (33, 132, 1), (19, 12, 4)):
        v = torch.tanh(torch.randn(B, T, P, 1, device=dev, generator=g))
        ret = torch.randn(B, T, P, 1, device=dev, generator=g)
        oc = torch.randint(-1, 2, (B, 1, P, 1), device=dev, generator=g).float()
        rew = torch.full((B, T, P, 1), -0.01, device=dev)
        rho = torch.rand(B, T, 1, 1, device=dev, generator=g)
        for rowlane in (0, 1):
            _lib.set_tuning("target_ws_rowlane", rowlane)
            for mode in ("MC", "TD", "UPGO", "VTRACE"):
                compute_target(mode, v, ret, rew, 0.7, 0.8, rho, rho)
                compute_target(mode, v, oc, None, 0.7, 1.0, rho, rho.clone())
            compute_target_pair("VTRACE", "UPGO", v, ret, rew, 0.7, 0.8, rho, rho)
            compute_target_heads("UPGO", "VTRACE", [(v, oc, None, 1.0), (v.clone(), ret, rew, 0.8)], 0.7, rho, rho)
            _lib.set_tuning("target_ws_rowlane", None)
        # non-contiguous / broadcast views -> the generic kernel
        compute_target("TD", v.transpose(0, 1).contiguous().transpose(0, 1), oc, None, 0.7, 1.0, rho, rho)


def loss_terms():
    from handyrl_b200 import nets
    from handyrl_b200.learner import compute_loss
    gen = BatchGenerator("TicTacToe", 200, seed=6, depth=64)
    gen.rollout(40)
    a = dict(ARGS, **{'lambda': 0.7, 'value_target': 'VTRACE', 'policy_target': 'UPGO', 'entropy_regularization': 0.1,
                      'entropy_regularization_decay': 0.1})
    b = make_batch_windows(gen.store, gen.select_windows(24, a), a)
    net = nets.SimpleConv2dModel().to(dev)
    losses, _ = compute_loss(b, net, None, a)
    losses['total'].backward()


if __name__ == "__main__":
    for game, n in (("TicTacToe", 70), ("ParallelTicTacToe", 33), ("Geister", 45), ("CIGeister", 97)):
        engine(game, n)
    rollouts("Geister", 9632, 9)      # 301 groups of 32 games: above the ticket threshold (2 x 148 SMs), ragged last group
    rollouts("TicTacToe", 132, 9)     # several waves need N * 27 floats to be a multiple of 4
    for game, n, plies in (("Geister", 45, 14), ("TicTacToe", 130, 12), ("ParallelTicTacToe", 66, 12)):
        batches(game, n, plies)
    targets()
    loss_terms()
    torch.cuda.synchronize()
    print("sanitize_smoke: all entry points ran")
