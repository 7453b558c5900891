"""GPU: the persistent self-play kernel (uniform epoch-0 policy).  Its action tape is
replayed through the C oracle: observations, action masks, terminal flags, winners,
episode table and counters must match bit-exactly."""
import numpy as np
import pytest
import torch

from oracle import games_c

pytestmark = pytest.mark.gpu
ENVS = {'tictactoe': 'TicTacToe', 'parallel_tictactoe': 'ParallelTicTacToe', 'geister': 'Geister',
        'ci_geister': 'CIGeister'}


def _rollout(env_key, N, K, chunks=1, seed=7):
    from handyrl_b200.engine import BatchedEnv
    env = BatchedEnv(ENVS[env_key], N)
    main, scalar = env.alloc_obs(waves=K)
    waves = dict(main=main, scalar=scalar, action_mask=torch.empty(K, N, env.A, device='cuda'))
    act = torch.full((K, N), -9, dtype=torch.int32, device='cuda')
    traj = torch.zeros(N, K, 8, dtype=torch.int32, device='cuda')     # slot-major ring [N][D]
    ept = torch.zeros(4 * N * K // 2 + 16, 4, dtype=torch.int32, device='cuda')
    ctr = torch.zeros(8, dtype=torch.int64, device='cuda')
    per = K // chunks
    for c in range(chunks):
        env.rollout(per, seed=seed, waves=waves, traj=traj, ep_table=ept, counters=ctr, out_action=act)
    torch.cuda.synchronize()
    # tests index the records as [ply][game]
    return env, waves, act.cpu().numpy(), traj.cpu().numpy().transpose(1, 0, 2).copy(), ept.cpu().numpy(), ctr.cpu().numpy()


@pytest.mark.parametrize("env_key,N,K", [('tictactoe', 96, 64), ('parallel_tictactoe', 96, 64), ('geister', 70, 420),
                                         ('ci_geister', 34, 250)])
def test_rollout_replays_through_oracle(env_key, N, K):
    env, waves, act, traj, ept, ctr = _rollout(env_key, N, K)
    geister = 'geister' in env_key
    main = waves['main'].cpu().numpy().reshape(K, N, -1)
    scalar = waves['scalar'].cpu().numpy() if geister else None
    amask = waves['action_mask'].cpu().numpy()
    flags = traj[:, :, 5].view(np.uint32)
    n_ep = 0
    wins = [0, 0, 0, 0]
    episodes = []
    for g in range(N):
        e = games_c.Env(env_key)
        steps = 0
        for k in range(K):
            f = int(flags[k, g])
            a, sel, first, term, win = f & 0xFFFF, (f >> 16) & 1, (f >> 24) & 1, (f >> 25) & 1, (f >> 26) & 3
            assert a == act[k, g]
            if e.terminal():
                e.reset()
                steps = 0
            assert first == (steps == 0)
            m, sc = e.observation(e.turn() if env_key != 'parallel_tictactoe' else 0)
            np.testing.assert_array_equal(main[k, g], m)
            if geister:
                np.testing.assert_array_equal(scalar[k, g], sc)
            legal = e.legal_actions()
            ref_am = np.full(env.A, 1e32, dtype=np.float32)
            ref_am[legal] = 0
            np.testing.assert_array_equal(amask[k, g], ref_am)
            assert a in legal
            e.step(a, sel)
            steps += 1
            assert term == int(e.terminal())
            assert int(traj[k, g, 6]) == steps
            if term:
                oc = e.outcome()
                code = 1 if oc[0] > 0 else (2 if oc[1] > 0 else 3)
                assert win == code
                wins[code] += 1
                n_ep += 1
                episodes.append((g, steps | (code << 16), k))
    assert ctr[0] == n_ep and ctr[1] == N * K and ctr[5] == n_ep
    assert list(ctr[2:5]) == wins[1:]
    got = sorted((int(r[0]), int(r[1]), int(r[2])) for r in ept[:n_ep])
    assert got == sorted(episodes)
    # final device state == oracle state (via observation of every view)
    ob = env.observe(view=-1)
    assert ob['main'].shape[0] == N


def test_rollout_is_chunk_invariant_and_seeded():
    a = _rollout('geister', 64, 128, chunks=1, seed=3)
    b = _rollout('geister', 64, 128, chunks=4, seed=3)
    c = _rollout('geister', 64, 128, chunks=1, seed=4)
    assert np.array_equal(a[2], b[2]) and np.array_equal(a[3], b[3])
    assert not np.array_equal(a[2], c[2])
    assert torch.equal(a[0].state, b[0].state)
    # launches that start and end inside a 4-ply record group (the rollout combines 4 records per 128-byte line)
    d = _rollout('geister', 64, 126, chunks=1, seed=3)
    e = _rollout('geister', 64, 126, chunks=6, seed=3)      # 21 plies per launch
    f = _rollout('geister', 64, 126, chunks=126, seed=3)    # one ply per launch
    assert np.array_equal(d[3], e[3]) and np.array_equal(d[3], f[3]) and np.array_equal(d[2], f[2])
    assert np.array_equal(d[3], a[3][:126])


def test_uniform_policy_statistics():
    """First TicTacToe move is uniform over 9 cells; Geister setup over 70 layouts."""
    from handyrl_b200.engine import BatchedEnv
    N = 90000
    for game, A, lo in (('TicTacToe', 9, 0), ('Geister', 70, 144)):
        env = BatchedEnv(game, N)
        act = torch.zeros(1, N, dtype=torch.int32, device='cuda')
        env.rollout(1, seed=5, out_action=act)
        h = torch.bincount(act[0].long() - lo, minlength=A).cpu().numpy()
        assert h.sum() == N and len(h) == A
        exp = N / A
        chi2 = ((h - exp) ** 2 / exp).sum()
        assert chi2 < 3 * A, (chi2, h)


def test_full_size_rollout_invariants():
    """BASELINE size: 65,536 concurrent Geister games; invariants that need no oracle."""
    from handyrl_b200.engine import BatchedEnv
    N, K = 65536, 256
    env = BatchedEnv('Geister', N)
    ctr = torch.zeros(8, dtype=torch.int64, device='cuda')
    main, scalar = env.alloc_obs(waves=1)
    waves = dict(main=main, scalar=scalar, action_mask=torch.empty(1, N, env.A, device='cuda'))
    env.rollout(K, seed=1, waves=waves, counters=ctr)
    c = ctr.cpu().numpy()
    assert c[1] == N * K
    assert c[0] == c[2] + c[3] + c[4] and c[0] > N // 2   # every finished game has exactly one result
    info = env.info().cpu().numpy()
    assert info[:, 2].max() <= 202 and info[:, 2].min() >= 1
    m = waves['main'][0]
    assert bool((m[:, 0] == 1).all())                        # plane 0 is all ones
    assert bool(((m == 0) | (m == 1)).all())
    assert bool((m[:, 1] + m[:, 2] <= 1).all())              # a cell is not both mine and theirs
    assert bool((m[:, 3] + m[:, 4] == m[:, 1]).all())        # my blue + my red == my pieces
    assert bool((m[:, 5:7] == 0).all())                      # hidden opponent types (Geister, player view)
    am = waves['action_mask'][0]
    assert bool(((am == 0) | (am == 1e32)).all()) and bool(((am == 0).sum(1) >= 1).all())


_DIGEST_SCRIPT = r"""
import hashlib, sys, torch
sys.path.insert(0, %r)
from handyrl_b200.engine import BatchedEnv
N, K = 16384, 42
out = []
for game in ('Geister', 'TicTacToe'):
    env = BatchedEnv(game, N)
    main, scalar = env.alloc_obs(waves=2)
    waves = dict(main=main, scalar=scalar, action_mask=torch.empty(2, N, env.A, device='cuda'))
    act = torch.full((K, N), -9, dtype=torch.int32, device='cuda')
    traj = torch.zeros(N, 64, 8, dtype=torch.int32, device='cuda')
    ept = torch.zeros(1 << 18, 4, dtype=torch.int32, device='cuda')   # > finished episodes: no ring wrap
    ctr = torch.zeros(8, dtype=torch.int64, device='cuda')
    for n in (21, 21):                                  # launches that end inside a 4-ply record group
        env.rollout(n, seed=11, waves=waves, traj=traj, ep_table=ept, counters=ctr, out_action=act)
    torch.cuda.synchronize()
    n_ep = int(ctr[5])
    assert n_ep <= ept.shape[0]
    rows = ept[:n_ep].cpu().numpy()
    rows = rows[rows[:, 0].argsort(kind='stable')]      # append order is not deterministic; content is
    rows = rows[(rows[:, 2].astype('int64') * N + rows[:, 0]).argsort(kind='stable')]
    h = hashlib.sha256()
    for t in (act, traj, env.state, main, waves['action_mask'], ctr[:5]):
        h.update(t.cpu().numpy().tobytes())
    if scalar is not None:
        h.update(scalar.cpu().numpy().tobytes())
    h.update(rows.tobytes())
    out.append(h.hexdigest())
print(' '.join(out))
"""


def test_ticketed_rollout_equals_static_rollout():
    """N = 16,384 games is above the ticket threshold (2 x SM count groups): the dynamic (group, chunk) schedule
    must produce exactly what the static one-warp-per-group schedule (B200RL_ROLLOUT_CHUNK=0) produces — moments,
    records, states, counters and the set of finished episodes."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    res = {}
    for name, chunk in (('tickets', None), ('static', '0'), ('tickets8', '8')):
        env = dict(os.environ)
        env.pop('B200RL_ROLLOUT_CHUNK', None)
        if chunk is not None:
            env['B200RL_ROLLOUT_CHUNK'] = chunk
        r = subprocess.run([sys.executable, '-c', _DIGEST_SCRIPT % root], env=env, capture_output=True, text=True,
                           timeout=600)
        assert r.returncode == 0, r.stderr[-2000:]
        res[name] = r.stdout.strip().splitlines()[-1]
    assert res['tickets'] == res['static'] == res['tickets8'], res


@pytest.mark.parametrize("game,N", [('Geister', 16384), ('Geister', 1000), ('CIGeister', 77), ('TicTacToe', 4096)])
def test_bulk_store_rollout_equals_stg_rollout(game, N):
    """k_rollout<G, TMA>: the staged bulk-store emit path (cp.async.bulk.global.shared::cta) writes the same
    moments, records, states and counters as the STG.128 path - full tickets, ragged last groups, odd row counts."""
    import torch
    from handyrl_b200 import _lib
    from handyrl_b200.engine import BatchedEnv
    res = []
    for knob in (0, 1):
        _lib.set_tuning('emit_tma', knob)
        try:
            env = BatchedEnv(game, N)
            W = 2 if N % 2 == 0 else 1               # several waves need an even N (16-byte aligned wave bases)
            main, scalar = env.alloc_obs(waves=W)
            main.fill_(-7)
            waves = dict(main=main, scalar=scalar, action_mask=torch.full((W, N, env.A), -7.0, device='cuda'))
            if scalar is not None:
                scalar.fill_(-7)
            act = torch.full((W, N), -9, dtype=torch.int32, device='cuda')
            traj = torch.zeros(N, 64, 8, dtype=torch.int32, device='cuda')
            ctr = torch.zeros(8, dtype=torch.int64, device='cuda')
            for n in (9, 12):
                env.rollout(n, seed=5, waves=waves, traj=traj, counters=ctr, out_action=act)
            torch.cuda.synchronize()
            res.append([t.clone() for t in (main, waves['action_mask'], act, traj, env.state, ctr[:5])] +
                       ([scalar.clone()] if scalar is not None else []))
        finally:
            _lib.set_tuning('emit_tma', None)
    for a, b in zip(*res):
        assert torch.equal(a, b)
    assert float(res[0][0].min()) >= 0            # every byte of the wave buffers was written
