"""CPU: the C oracle of the game rules against tapes recorded from the reference itself."""
import os

import numpy as np
import pytest

from oracle import games_c

ENVS = ['tictactoe', 'parallel_tictactoe', 'geister', 'ci_geister']


def _bits(a, n):
    return np.unpackbits(a, bitorder='little')[:n].astype(np.float32)


@pytest.mark.parametrize("env", ENVS)
def test_c_oracle_replays_reference_tapes(golden_dir, env):
    t = dict(np.load(os.path.join(golden_dir, "tape_%s.npz" % env)))   # NpzFile re-inflates on every access
    geister = 'geister' in env
    nmain = 252 if geister else 27
    e = games_c.Env(env)
    game = -1
    n_states = len(t['game'])
    for i in range(n_states):
        if t['game'][i] != game:
            game = t['game'][i]
            e.reset()
            assert t['action'][i] == -1
        else:
            e.step(t['action'][i], t['sel'][i])
            assert abs(e.reward() - t['reward'][i]) < 1e-9 or not geister
        assert e.terminal() == bool(t['terminal'][i])
        if env != 'parallel_tictactoe':
            assert e.turn() == t['turn'][i]
        legal = list(t['legal_flat'][t['legal_off'][i]:t['legal_off'][i + 1]])
        assert e.legal_actions() == legal, (i, e.legal_actions(), legal)
        for key, pl in (('0', 0), ('1', 1), ('N', None)):
            main, sc = e.observation(pl)
            np.testing.assert_array_equal(main, _bits(t['obs' + key][i], nmain))
            if geister:
                np.testing.assert_array_equal(sc, _bits(t['sc' + key][i], 18))
        last = i + 1 == n_states or t['game'][i + 1] != game
        if last:
            assert e.terminal()
            assert e.outcome() == list(t['outcome'][game])


@pytest.mark.parametrize("env", ENVS)
def test_c_random_tapes_are_self_consistent(env):
    d = games_c.random_tapes(env, 20, seed=3)
    assert d['terminal'].sum() == 20
    e = games_c.Env(env)
    # replaying the recorded actions reproduces the recorded legal lists
    game = -1
    for i in range(len(d['game'])):
        if d['game'][i] != game:
            game = d['game'][i]
            e.reset()
        else:
            e.step(d['action'][i], d['sel'][i])
        n = d['legal_count'][i]
        assert e.legal_actions() == list(d['legal_ids'][i][:n])


def test_selfplay_baseline_driver_runs():
    assert games_c.selfplay('geister', 2000, seed=1, n_threads=2) == 4000


def test_hotpath_port_targets_match_numpy_oracle():
    """oracle/hotpath_oracle.c (the timed CPU baseline of bench.py) against oracle/targets_np.py, bit-exact"""
    import numpy as np
    from oracle import games_c, targets_np
    hp = games_c.HotPath(32, 32, 16, 2, replay_cap=128, seed=4)
    steps = cells = 0
    for _ in range(10):
        n, c, fin = hp.step()
        steps += n
        cells += c
    # no windows are cut before the first episode finishes; afterwards every generated ply is batched once
    assert steps == 10 * 2 * 32 * 32 and c == 2 * 32 * 32 and cells <= steps and fin > 0
    d = hp.peek()
    tv, _ = targets_np.upgo(d['v'][0], d['outcome'], None, 0.7, 1.0)
    _, adv = targets_np.vtrace(d['v'][0], d['outcome'], None, 0.7, 1.0, d['rho'], d['rho'])
    assert np.array_equal(tv, d['tv'][0]) and np.array_equal(adv, d['adv'][0])
    tv, _ = targets_np.upgo(d['v'][1], d['ret'], d['reward'], 0.7, 0.8)
    _, adv = targets_np.vtrace(d['v'][1], d['ret'], d['reward'], 0.7, 0.8, d['rho'], d['rho'])
    assert np.array_equal(tv, d['tv'][1]) and np.array_equal(adv, d['adv'][1])
    # batch columns: rewards are -0.01 on live rows and 0 on padding; returns follow generation.py:73-77
    assert set(np.unique(d['reward']).tolist()) <= {0.0, float(np.float32(-0.01))}
    assert float(d['ret'].min()) >= -0.05 - 1e-6 and float(d['ret'].max()) <= 0.0
    assert set(np.unique(d['outcome']).tolist()) <= {-1.0, 0.0, 1.0}
    assert sum(hp.phase_seconds()) > 0
